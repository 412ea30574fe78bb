// join_noalloc.cu — the owner join for callers that must not allocate.
//
// nbg_owner_join_device asks its caller for device scratch memory (the join's
// hash table), and nbg_owner_join allocates that table itself.  Neither suits a
// caller that is not allowed to touch the CUDA allocator while it runs: a
// reconcile tick that has promised a latency bound (cudaMalloc and cudaFree
// synchronise the whole device and can block behind unrelated work), or a
// controller that shares its GPU with the notebooks it manages and comes up
// after they have claimed the memory, so an allocation made "when needed" may
// simply fail.  For those callers the table lives in this file, in a fixed-size
// array that is part of the library image, and a call costs no allocator call.
//
// When the memory is claimed: CUDA loads a module lazily, and nothing in this
// file runs on the device, so the first cudaGetSymbolAddress on a device — the
// set-up inside the first nbg_owner_join_static call there — is what loads the
// module and reserves the array.  That one call can block or run out of memory
// (NBG_E_NOMEM); callers make it once at start-up, as a warm-up, and every later
// call finds the address cached.  The table then stays until the process exits.
//
// The price is a hard limit (NBG_JOIN_STATIC_MAX_STS StatefulSets per call) and
// 3 MiB of device memory on every GPU the entry point has been used on.
#include <cuda_runtime.h>

#include <mutex>

#include "../../../include/nbgraft.h"

extern "C" {
unsigned long long nbg_join_table_slots(unsigned long long m);
cudaError_t nbg_launch_owner_join(const unsigned long long*, unsigned long long, const unsigned long long*,
                                  unsigned long long, uint32_t*, void*, cudaStream_t);
}

// 2 x NBG_JOIN_STATIC_MAX_STS slots, 8-byte key + 4-byte value each (join.cu's layout).
constexpr unsigned long long kStaticSlots = 2ull * NBG_JOIN_STATIC_MAX_STS;
constexpr size_t kStaticBytes = (size_t)kStaticSlots * 12u;

// The table.  In host code this NAME is only a handle for the runtime, not a device
// address: never pass `g_join_table` itself to a kernel — resolve it per device first.
static __device__ __align__(256) unsigned char g_join_table[kStaticBytes];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_mu;
void* g_table_addr[kMaxDevices] = {};   // device address of g_join_table, per CUDA device ordinal

// Once per device, with that device current: ask the runtime where this device's
// copy of the array lives.
int table_for_current_device(void** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return NBG_E_CUDA;
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_table_addr[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_join_table);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      // this lookup is where the module is loaded and the array gets its memory: a full device fails here
      return e == cudaErrorMemoryAllocation ? NBG_E_NOMEM : NBG_E_CUDA;
    }
    g_table_addr[dev] = p;
  }
  *out = g_table_addr[dev];
  return NBG_OK;
}

}  // namespace

extern "C" int nbg_owner_join_static(const uint64_t* nb_key, uint64_t n, const uint64_t* sts_key, uint64_t m,
                                     uint32_t* match, void* stream) {
  if (n == 0) return NBG_OK;
  if (!nb_key || !match || (m && !sts_key)) return NBG_E_INVAL;
  if (m > NBG_JOIN_STATIC_MAX_STS) return NBG_E_TOOBIG;   // the built-in table holds no more than this
  if (nbg_join_table_slots(m) > kStaticSlots) return NBG_E_TOOBIG;
  void* table = nullptr;
  if (int rc = table_for_current_device(&table)) return rc;
  // The launcher empties the table before building it, so whatever the previous call left
  // in the shared array is gone before anyone reads it.
  if (nbg_launch_owner_join((const unsigned long long*)nb_key, n, (const unsigned long long*)sts_key, m, match, table,
                            (cudaStream_t)stream) != cudaSuccess)
    return NBG_E_CUDA;
  return NBG_OK;
}
