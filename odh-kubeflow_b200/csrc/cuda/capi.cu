// capi.cu — the C ABI of libnbgraft.so (include/nbgraft.h).
//
// Host-buffer entry points run a two-slot pipeline per device: chunk c+1 is
// copied host->device on the copy stream while chunk c is compared on the run
// stream, and the (tiny) results return on the run stream.  Device staging is
// a grow-only arena per slot, so steady-state calls allocate nothing.
// There is no CPU fallback anywhere in this file: without a device nbg_init
// fails and every other entry point rejects the null context.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <queue>
#include <vector>

#include "../../../include/nbgraft.h"
#include "../hash.h"

extern "C" {
cudaError_t nbg_launch_cull(const nbg_cull_batch*, unsigned long long, const nbg_cull_params*,
                            unsigned, unsigned, uint8_t*, uint32_t*, uint64_t*, uint64_t*,
                            cudaStream_t);
cudaError_t nbg_launch_diff_sts(const nbg_sts_batch*, unsigned long long, const unsigned*, int, uint8_t*,
                                uint64_t*, cudaStream_t);
cudaError_t nbg_launch_diff_sts_sparse(const nbg_sts_sparse_batch*, unsigned long long, const unsigned*, uint8_t*,
                                       uint64_t*, cudaStream_t);
cudaError_t nbg_launch_diff_sts_gen(const nbg_sts_batch*, const uint64_t*, const uint8_t*, const nbg_sts_gen_consts*,
                                    unsigned long long, const unsigned*, uint8_t*, uint64_t*, cudaStream_t);
cudaError_t nbg_launch_diff_svc(const nbg_svc_batch*, unsigned long long, const unsigned*, uint8_t*,
                                uint64_t*, cudaStream_t);
cudaError_t nbg_launch_diff_vs(const nbg_vs_batch*, unsigned long long, const unsigned*, uint8_t*,
                               uint64_t*, cudaStream_t);
cudaError_t nbg_launch_diff_generic(const nbg_generic_batch*, unsigned long long, int, const unsigned*, uint8_t*,
                                    uint64_t*, cudaStream_t);
cudaError_t nbg_launch_diff_svc_gen(const nbg_svc_gen_batch*, unsigned long long, const unsigned*, uint8_t*, uint64_t*, cudaStream_t);
cudaError_t nbg_launch_diff_vs_gen(const nbg_vs_gen_batch*, const nbg_vs_gen_consts*, unsigned long long, const unsigned*, uint8_t*,
                                   uint64_t*, cudaStream_t);
cudaError_t nbg_launch_cull_text_fixup(const uint8_t*, unsigned long long, uint32_t*, uint32_t*, uint8_t*, cudaStream_t);
cudaError_t nbg_launch_cull_text_zone(const uint8_t*, const uint32_t*, const uint32_t*, const int32_t*, unsigned,
                                      const uint32_t*, const int32_t*, unsigned, unsigned long long, int32_t*, cudaStream_t);
cudaError_t nbg_launch_count_running(const uint32_t*, const uint64_t*, const uint64_t*, const uint8_t*, unsigned long long,
                                     unsigned, uint32_t*, cudaStream_t);
cudaError_t nbg_launch_restart_block(const nbg_restart_batch*, unsigned long long, const unsigned*, uint8_t*, uint64_t*,
                                     cudaStream_t);
cudaError_t nbg_launch_parse_records(const uint32_t*, const unsigned char*, const uint32_t*, const unsigned char*,
                                     unsigned long long, unsigned, unsigned, int64_t*, uint32_t*, int32_t*, cudaStream_t);
cudaError_t nbg_launch_hash_strings(const uint32_t*, const unsigned char*, unsigned long long, unsigned, uint64_t*,
                                    cudaStream_t);
unsigned long long nbg_join_table_slots(unsigned long long m);
cudaError_t nbg_launch_owner_join(const unsigned long long*, unsigned long long, const unsigned long long*,
                                  unsigned long long, uint32_t*, void*, cudaStream_t);
}

namespace {

thread_local char g_cuda_err[256] = "";

bool cuda_ok(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return true;
  snprintf(g_cuda_err, sizeof g_cuda_err, "%.120s: %.120s", what, cudaGetErrorString(e));   // `what` may be a whole call
  return false;
}
#define CU(call) do { if (!cuda_ok((call), #call)) return NBG_E_CUDA; } while (0)

struct Slot {
  char* dev = nullptr;
  size_t cap = 0;
  cudaEvent_t in_done = nullptr, free_ev = nullptr;
  bool used = false;
};

struct Dev {
  int id = 0;
  cudaStream_t s_in = nullptr, s_run = nullptr;
  cudaStream_t s_side[3] = {nullptr, nullptr, nullptr};   // nbg_step_device: the three small predicates
  cudaEvent_t ev_fork = nullptr, ev_join[3] = {nullptr, nullptr, nullptr};
  Slot slot[2];
  char* scratch = nullptr;   // grow-only kernel scratch (the owner join's hash table)
  size_t scratch_cap = 0;
  std::mutex mu;
};

}  // namespace

struct nbg_ctx {
  std::vector<Dev*> devs;
  std::atomic<uint64_t> launches{0};
};

namespace {

constexpr size_t kAlign = 256;
size_t align_up(size_t x) { return (x + kAlign - 1) / kAlign * kAlign; }

// Lay out one chunk in a slot's arena.
struct Layout {
  struct In { const void* host; size_t bytes; size_t off; };
  struct Out { void* host; size_t bytes; size_t off; };
  std::vector<In> ins;
  std::vector<Out> outs;
  size_t total = 0;
  size_t in(const void* host, size_t bytes) {
    const size_t off = total;
    ins.push_back({host, bytes, off});
    total += align_up(bytes ? bytes : 1);
    return off;
  }
  size_t scratch(size_t bytes) {            // device-only region of the arena (neither uploaded nor downloaded)
    const size_t off = total;
    total += align_up(bytes ? bytes : 1);
    return off;
  }
  size_t out(void* host, size_t bytes) {
    const size_t off = total;
    outs.push_back({host, bytes, off});
    total += align_up(bytes ? bytes : 1);
    return off;
  }
};

int slot_reserve(Slot& sl, size_t bytes) {
  if (sl.cap >= bytes) return NBG_OK;
  if (sl.used) CU(cudaEventSynchronize(sl.free_ev));
  if (sl.dev) { CU(cudaFree(sl.dev)); sl.dev = nullptr; sl.cap = 0; }
  const size_t want = bytes + bytes / 4;
  if (cudaMalloc(&sl.dev, want) != cudaSuccess) {
    cudaGetLastError();
    if (!cuda_ok(cudaMalloc(&sl.dev, bytes), "cudaMalloc(staging)")) { sl.dev = nullptr; return NBG_E_NOMEM; }
    sl.cap = bytes;
  } else {
    sl.cap = want;
  }
  return NBG_OK;
}

int slot_upload(Dev& d, Slot& sl, const Layout& lay) {
  if (int rc = slot_reserve(sl, lay.total)) return rc;
  if (sl.used) CU(cudaStreamWaitEvent(d.s_in, sl.free_ev, 0));
  for (const auto& in : lay.ins)
    if (in.bytes) CU(cudaMemcpyAsync(sl.dev + in.off, in.host, in.bytes, cudaMemcpyHostToDevice, d.s_in));
  CU(cudaEventRecord(sl.in_done, d.s_in));
  CU(cudaStreamWaitEvent(d.s_run, sl.in_done, 0));
  return NBG_OK;
}

int slot_download(Dev& d, Slot& sl, const Layout& lay) {
  for (const auto& o : lay.outs)
    if (o.bytes) CU(cudaMemcpyAsync(o.host, sl.dev + o.off, o.bytes, cudaMemcpyDeviceToHost, d.s_run));
  CU(cudaEventRecord(sl.free_ev, d.s_run));
  sl.used = true;
  return NBG_OK;
}

// Offsets of rows [i0,i1] must be monotone.  Branch-free so it vectorises: the check runs
// per pipeline chunk, right before that chunk is queued, and must not hold the copies up.
bool off_ok(const uint32_t* off, uint64_t i0, uint64_t i1) {
  unsigned bad = 0;
  for (uint64_t i = i0; i < i1; ++i) bad |= (unsigned)(off[i] > off[i + 1]);
  return bad == 0;
}
bool csr_ok(const nbg_kvmap& m, uint64_t i0, uint64_t i1) { return off_ok(m.off, i0, i1); }
// Sparse tiles of rows [i0,i1): offsets even, every run as long as its mask says (popcount, rounded up to even).
// The kernels size their reads of BOTH sides by one side's run when the masks agree, so this is what keeps a
// malformed batch from reading outside the arrays.  Branch-free, runs per chunk like off_ok.
bool sparse_ok(const nbg_podsparse& p, uint64_t i0, uint64_t i1) {
  unsigned bad = p.off[i0] & 1u;
  for (uint64_t i = i0; i < i1; ++i) {
    const unsigned bits = (unsigned)(__builtin_popcountll(p.mask[3 * i]) + __builtin_popcountll(p.mask[3 * i + 1]) +
                                     __builtin_popcountll(p.mask[3 * i + 2]));
    bad |= (unsigned)(p.off[i + 1] - p.off[i] != ((bits + 1u) & ~1u));
  }
  return bad == 0;
}
// host-buffer entry points take CSR maps only (`end` must be NULL: chunking relies on contiguous runs)
bool csr_ptrs_ok(const nbg_kvmap& m, uint64_t n) { return m.off && !m.end && (m.ent || m.off[n] == m.off[0]); }

// Upload rows [i0,i1) of a CSR map; returns device pointers and the base offset.
struct CsrDev { size_t off_at, ent_at; unsigned base; };
CsrDev csr_in(Layout& lay, const nbg_kvmap& m, uint64_t i0, uint64_t i1) {
  CsrDev c;
  c.base = m.off[i0];
  c.off_at = lay.in(m.off + i0, (i1 - i0 + 1) * sizeof(uint32_t));
  const size_t cnt = m.off[i1] - m.off[i0];
  c.ent_at = lay.in(cnt ? m.ent + m.off[i0] : nullptr, cnt * sizeof(nbg_kv));
  return c;
}
nbg_kvmap csr_ptr(const Slot& sl, const CsrDev& c) {
  nbg_kvmap m{};
  m.off = reinterpret_cast<const uint32_t*>(sl.dev + c.off_at);
  m.ent = reinterpret_cast<const nbg_kv*>(sl.dev + c.ent_at);
  return m;
}

// Contiguous, 64-aligned split of [0,n) over the context's devices.
void device_range(uint64_t n, int ndev, int d, uint64_t* lo, uint64_t* hi) {
  const uint64_t words = (n + 63) / 64;
  const uint64_t w_lo = words * d / ndev, w_hi = words * (d + 1) / ndev;
  *lo = w_lo * 64 < n ? w_lo * 64 : n;
  *hi = w_hi * 64 < n ? w_hi * 64 : n;
}

// Set while nbg_step drives several entry points as ONE pipeline: they then only enqueue
// (the caller holds the device locks and waits once at the end).
thread_local bool tl_batched = false;
// Set while nbg_diff_deploy[_device] drives the StatefulSet entry points with CopyDeploymentSetFields' replicas rule.
thread_local bool tl_deployment = false;

template <class F>
int for_each_device(nbg_ctx* ctx, uint64_t n, F&& run_range) {
  // Enqueue every device's chunks first, then wait: devices overlap.
  const int ndev = (int)ctx->devs.size();
  int rc = NBG_OK;
  if (tl_batched) {
    for (int d = 0; d < ndev && rc == NBG_OK; ++d) {
      uint64_t lo, hi;
      device_range(n, ndev, d, &lo, &hi);
      if (hi <= lo) continue;
      if (!cuda_ok(cudaSetDevice(ctx->devs[d]->id), "cudaSetDevice")) return NBG_E_CUDA;
      rc = run_range(*ctx->devs[d], lo, hi);
    }
    return rc;
  }
  std::vector<std::unique_lock<std::mutex>> locks;
  for (int d = 0; d < ndev; ++d) locks.emplace_back(ctx->devs[d]->mu);
  for (int d = 0; d < ndev && rc == NBG_OK; ++d) {
    uint64_t lo, hi;
    device_range(n, ndev, d, &lo, &hi);
    if (hi <= lo) continue;
    if (!cuda_ok(cudaSetDevice(ctx->devs[d]->id), "cudaSetDevice")) { rc = NBG_E_CUDA; break; }
    rc = run_range(*ctx->devs[d], lo, hi);
  }
  for (int d = 0; d < ndev; ++d) {
    cudaSetDevice(ctx->devs[d]->id);
    if (!cuda_ok(cudaStreamSynchronize(ctx->devs[d]->s_run), "cudaStreamSynchronize") && rc == NBG_OK)
      rc = NBG_E_CUDA;
    if (!cuda_ok(cudaStreamSynchronize(ctx->devs[d]->s_in), "cudaStreamSynchronize") && rc == NBG_OK)
      rc = NBG_E_CUDA;
  }
  return rc;
}

// Objects per pipeline chunk (multiples of 64 so bitmap words never straddle chunks).
// NBG_CHUNK_SCALE (environment, power of two, default 4) scales all of them — a tuning knob
// (measured through nbg_step at 1.25 M notebooks: 1 -> 30.5 ms, 2 -> 28.5, 4 -> 27.7, 8 -> 27.6).
uint64_t chunk_scale() {
  static const uint64_t s = [] {
    const char* e = getenv("NBG_CHUNK_SCALE");
    long v = e ? atol(e) : 4;
    return (uint64_t)(v >= 1 && v <= 64 ? v : 1);
  }();
  return s;
}
#define kStsChunk (8192 * chunk_scale())          // ~27 MB   (sizes at scale 1)
#define kStsSparseChunk (32768 * chunk_scale())   // ~25 MB at ~35 non-zero slots per side
#define kSvcChunk (65536 * chunk_scale())         // ~21 MB
#define kVsChunk (131072 * chunk_scale())         // ~20 MB
#define kCullChunk (262144 * chunk_scale())       // notebooks, ~22 MB at 4 records each
#define kTextChunk (1048576 * chunk_scale())      // strings, ~30 MB of timestamps

int check_dev(const nbg_ctx* ctx, int device) {
  if (!ctx) return NBG_E_INVAL;
  if (device < 0 || device >= (int)ctx->devs.size()) return NBG_E_INVAL;
  return NBG_OK;
}

}  // namespace

extern "C" {

int nbg_abi_version(void) { return NBG_ABI_VERSION; }

const char* nbg_last_cuda_error(void) { return g_cuda_err; }

const char* nbg_strerror(int status) {
  switch (status) {
    case NBG_OK: return "ok";
    case NBG_E_NODEVICE: return "no usable CUDA device (this library has no CPU path)";
    case NBG_E_INVAL: return "invalid argument";
    case NBG_E_NOMEM: return "device or pinned-host allocation failed";
    case NBG_E_CUDA: return "CUDA call failed (see nbg_last_cuda_error)";
    case NBG_E_NILREPLICAS: return "an object has a nil Spec.Replicas (the Go reference would panic)";
    case NBG_E_RANGE: return "offsets not monotone, run length disagrees with its mask, or slot out of range";
    case NBG_E_TOOBIG: return "batch exceeds a documented maximum";
    default: return "unknown status";
  }
}

uint64_t nbg_hash64(const void* data, size_t len) { return nbg_hash_bytes(data, len); }
uint32_t nbg_shard_of(uint64_t ns_hash, uint32_t n_shards) { return nbg_shard_of_inline(ns_hash, n_shards); }

int nbg_balance_shards(const uint64_t* ns_hash, const uint64_t* ns_size, uint64_t n_ns, uint32_t n_shards,
                       uint32_t* shard_out, uint64_t* load_out) {
  if (n_shards == 0 || (n_ns && (!ns_hash || !ns_size || !shard_out))) return NBG_E_INVAL;
  std::vector<uint64_t> order(n_ns);
  for (uint64_t k = 0; k < n_ns; ++k) order[k] = k;
  std::sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) {
    if (ns_size[a] != ns_size[b]) return ns_size[a] > ns_size[b];
    if (ns_hash[a] != ns_hash[b]) return ns_hash[a] < ns_hash[b];
    return a < b;
  });
  // lightest shard first; ties go to the lowest shard number
  typedef std::pair<uint64_t, uint32_t> Load;
  std::priority_queue<Load, std::vector<Load>, std::greater<Load>> heap;
  for (uint32_t s = 0; s < n_shards; ++s) heap.push(Load(0, s));
  for (uint64_t k : order) {
    Load l = heap.top();
    heap.pop();
    shard_out[k] = l.second;
    l.first += ns_size[k];
    heap.push(l);
  }
  if (load_out)
    for (; !heap.empty(); heap.pop()) load_out[heap.top().second] = heap.top().first;
  return NBG_OK;
}

int nbg_init(const int* devices, int n_devices, nbg_ctx** out) {
  if (!out) return NBG_E_INVAL;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) {
    cudaGetLastError();
    return NBG_E_NODEVICE;
  }
  const int one = 0;
  if (!devices) { devices = &one; n_devices = 1; }
  if (n_devices <= 0) return NBG_E_INVAL;
  nbg_ctx* ctx = new (std::nothrow) nbg_ctx();
  if (!ctx) return NBG_E_NOMEM;
  for (int k = 0; k < n_devices; ++k) {
    if (devices[k] < 0 || devices[k] >= count) { nbg_destroy(ctx); return NBG_E_NODEVICE; }
    Dev* d = new (std::nothrow) Dev();
    if (!d) { nbg_destroy(ctx); return NBG_E_NOMEM; }
    d->id = devices[k];
    ctx->devs.push_back(d);
    bool ok = cuda_ok(cudaSetDevice(d->id), "cudaSetDevice") &&
              cuda_ok(cudaStreamCreateWithFlags(&d->s_in, cudaStreamNonBlocking), "cudaStreamCreate") &&
              cuda_ok(cudaStreamCreateWithFlags(&d->s_run, cudaStreamNonBlocking), "cudaStreamCreate");
    for (int s = 0; ok && s < 2; ++s)
      ok = cuda_ok(cudaEventCreateWithFlags(&d->slot[s].in_done, cudaEventDisableTiming), "cudaEventCreate") &&
           cuda_ok(cudaEventCreateWithFlags(&d->slot[s].free_ev, cudaEventDisableTiming), "cudaEventCreate");
    ok = ok && cuda_ok(cudaEventCreateWithFlags(&d->ev_fork, cudaEventDisableTiming), "cudaEventCreate");
    for (int k = 0; ok && k < 3; ++k)
      ok = cuda_ok(cudaStreamCreateWithFlags(&d->s_side[k], cudaStreamNonBlocking), "cudaStreamCreate") &&
           cuda_ok(cudaEventCreateWithFlags(&d->ev_join[k], cudaEventDisableTiming), "cudaEventCreate");
    if (!ok) { nbg_destroy(ctx); return NBG_E_CUDA; }
  }
  *out = ctx;
  return NBG_OK;
}

void nbg_destroy(nbg_ctx* ctx) {
  if (!ctx) return;
  for (Dev* d : ctx->devs) {
    cudaSetDevice(d->id);
    if (d->s_run) cudaStreamSynchronize(d->s_run);
    if (d->s_in) cudaStreamSynchronize(d->s_in);
    if (d->scratch) cudaFree(d->scratch);
    for (auto& sl : d->slot) {
      if (sl.dev) cudaFree(sl.dev);
      if (sl.in_done) cudaEventDestroy(sl.in_done);
      if (sl.free_ev) cudaEventDestroy(sl.free_ev);
    }
    for (int k = 0; k < 3; ++k) {
      if (d->s_side[k]) { cudaStreamSynchronize(d->s_side[k]); cudaStreamDestroy(d->s_side[k]); }
      if (d->ev_join[k]) cudaEventDestroy(d->ev_join[k]);
    }
    if (d->ev_fork) cudaEventDestroy(d->ev_fork);
    if (d->s_in) cudaStreamDestroy(d->s_in);
    if (d->s_run) cudaStreamDestroy(d->s_run);
    delete d;
  }
  delete ctx;
}

int nbg_device_count(const nbg_ctx* ctx) { return ctx ? (int)ctx->devs.size() : 0; }
int nbg_device_id(const nbg_ctx* ctx, int device) {
  return (ctx && device >= 0 && device < (int)ctx->devs.size()) ? ctx->devs[device]->id : -1;
}
uint64_t nbg_launch_count(const nbg_ctx* ctx) { return ctx ? ctx->launches.load() : 0; }

void* nbg_alloc_pinned(size_t bytes) {
  void* p = nullptr;
  if (!cuda_ok(cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable), "cudaHostAlloc")) return nullptr;
  return p;
}
void nbg_free_pinned(void* p) { if (p) cudaFreeHost(p); }

void* nbg_ctx_stream(nbg_ctx* ctx, int device) {
  if (check_dev(ctx, device) != NBG_OK) return nullptr;
  return (void*)ctx->devs[device]->s_run;
}

int nbg_sync(nbg_ctx* ctx, int device) {
  if (int rc = check_dev(ctx, device)) return rc;
  CU(cudaSetDevice(ctx->devs[device]->id));
  CU(cudaStreamSynchronize(ctx->devs[device]->s_run));
  return NBG_OK;
}

// ------------------------------------------------------------ device-pointer API
int nbg_cull_device(nbg_ctx* ctx, int device, const nbg_cull_batch* in, uint64_t n,
                    const nbg_cull_params* p, uint8_t* action, uint32_t* argmax,
                    uint64_t* cull_bitmap, uint64_t* period_bitmap, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (!in || !p) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!action || !argmax || !cull_bitmap || !period_bitmap) return NBG_E_INVAL;
  if (p->now_nsec < 0 || p->now_nsec > 999999999 || (p->mode != NBG_CULL_PREDICATES && p->mode != NBG_CULL_RECONCILE))
    return NBG_E_INVAL;
  if (p->cull_idle_min > NBG_CULL_MAX_MINUTES || p->cull_idle_min < -NBG_CULL_MAX_MINUTES ||
      p->check_period_min > NBG_CULL_MAX_MINUTES || p->check_period_min < -NBG_CULL_MAX_MINUTES)
    return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  cudaStream_t s = (cudaStream_t)stream;
  CU(nbg_launch_cull(in, n, p, 0, 0, action, argmax, cull_bitmap, period_bitmap, s));
  ctx->launches++;
  return NBG_OK;
}

int nbg_diff_sts_device(nbg_ctx* ctx, int device, const nbg_sts_batch* in, uint64_t n,
                        uint8_t* reason, uint64_t* bitmap, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (!in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  cudaStream_t s = (cudaStream_t)stream;
  const unsigned bases[4] = {0, 0, 0, 0};
  CU(nbg_launch_diff_sts(in, n, bases, tl_deployment ? 1 : 0, reason, bitmap, s));
  ctx->launches++;
  return NBG_OK;
}

int nbg_diff_sts_gen_device(nbg_ctx* ctx, int device, const nbg_sts_batch* in, const nbg_sts_gen_inputs* gen,
                            const nbg_sts_gen_consts* consts, uint64_t n, uint8_t* reason, uint64_t* bitmap, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (!in || !gen || !consts) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap || !gen->prefix_hash || !gen->flags) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  const unsigned bases[3] = {0, 0, 0};
  CU(nbg_launch_diff_sts_gen(in, gen->prefix_hash, gen->flags, consts, n, bases, reason, bitmap, (cudaStream_t)stream));
  ctx->launches++;
  return NBG_OK;
}

int nbg_diff_deploy_device(nbg_ctx* ctx, int device, const nbg_sts_batch* in, uint64_t n,
                           uint8_t* reason, uint64_t* bitmap, void* stream) {
  tl_deployment = true;
  const int rc = nbg_diff_sts_device(ctx, device, in, n, reason, bitmap, stream);
  tl_deployment = false;
  return rc;
}

int nbg_diff_deploy(nbg_ctx* ctx, const nbg_sts_batch* in, uint64_t n, uint8_t* reason, uint64_t* bitmap) {
  tl_deployment = true;
  const int rc = nbg_diff_sts(ctx, in, n, reason, bitmap);
  tl_deployment = false;
  return rc;
}

int nbg_diff_sts_sparse_device(nbg_ctx* ctx, int device, const nbg_sts_sparse_batch* in, uint64_t n,
                               uint8_t* reason, uint64_t* bitmap, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (!in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  const unsigned bases[6] = {0, 0, 0, 0, 0, 0};
  CU(nbg_launch_diff_sts_sparse(in, n, bases, reason, bitmap, (cudaStream_t)stream));
  ctx->launches++;
  return NBG_OK;
}

int nbg_diff_svc_device(nbg_ctx* ctx, int device, const nbg_svc_batch* in, uint64_t n,
                        uint8_t* reason, uint64_t* bitmap, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (!in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  cudaStream_t s = (cudaStream_t)stream;
  const unsigned bases[6] = {0, 0, 0, 0, 0, 0};
  CU(nbg_launch_diff_svc(in, n, bases, reason, bitmap, s));
  ctx->launches++;
  return NBG_OK;
}

int nbg_diff_vs_device(nbg_ctx* ctx, int device, const nbg_vs_batch* in, uint64_t n,
                       uint8_t* reason, uint64_t* bitmap, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (!in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  cudaStream_t s = (cudaStream_t)stream;
  const unsigned bases[2] = {0, 0};
  CU(nbg_launch_diff_vs(in, n, bases, reason, bitmap, s));
  ctx->launches++;
  return NBG_OK;
}

size_t nbg_owner_join_scratch_bytes(uint64_t m) { return (size_t)nbg_join_table_slots(m) * 12u; }

uint64_t nbg_join_key(uint64_t ns_hash, uint64_t uid_hash) {
  const uint64_t k = nbg_combine(nbg_combine(0x6A6F696E6B657921ull, ns_hash), uid_hash);
  return k ? k : 1;
}

int nbg_owner_join_device(nbg_ctx* ctx, int device, const uint64_t* nb_key, uint64_t n,
                          const uint64_t* sts_key, uint64_t m, uint32_t* match,
                          void* scratch, size_t scratch_bytes, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (n == 0) return NBG_OK;
  if (!nb_key || !match || (m && !sts_key) || !scratch) return NBG_E_INVAL;
  if (m >= (1ull << 31)) return NBG_E_TOOBIG;
  if (scratch_bytes < nbg_owner_join_scratch_bytes(m) || ((uintptr_t)scratch & 7u)) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  CU(nbg_launch_owner_join((const unsigned long long*)nb_key, n, (const unsigned long long*)sts_key, m, match,
                           scratch, (cudaStream_t)stream));
  ctx->launches += m ? 2 : 1;
  return NBG_OK;
}

int nbg_diff_generic_device(nbg_ctx* ctx, int device, const nbg_generic_batch* in, uint64_t n, int mode,
                            uint8_t* reason, uint64_t* bitmap, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (!in || (mode != NBG_GEN_DEEPEQUAL && mode != NBG_GEN_CONFIGMAP)) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  const unsigned bases[4] = {0, 0, 0, 0};
  CU(nbg_launch_diff_generic(in, n, mode, bases, reason, bitmap, (cudaStream_t)stream));
  ctx->launches++;
  return NBG_OK;
}

int nbg_diff_svc_gen_device(nbg_ctx* ctx, int device, const nbg_svc_gen_batch* in, uint64_t n, uint8_t* reason,
                            uint64_t* bitmap, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (!in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  const unsigned bases[3] = {0, 0, 0};
  CU(nbg_launch_diff_svc_gen(in, n, bases, reason, bitmap, (cudaStream_t)stream));
  ctx->launches++;
  return NBG_OK;
}

int nbg_diff_vs_gen_device(nbg_ctx* ctx, int device, const nbg_vs_gen_batch* in, const nbg_vs_gen_consts* consts,
                           uint64_t n, uint8_t* reason, uint64_t* bitmap, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (!in || !consts) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  const unsigned bases[2] = {0, 0};
  CU(nbg_launch_diff_vs_gen(in, consts, n, bases, reason, bitmap, (cudaStream_t)stream));
  ctx->launches++;
  return NBG_OK;
}

int nbg_count_running_device(nbg_ctx* ctx, int device, const uint32_t* ns_id, const uint64_t* name_hash,
                             const uint64_t* label_hash, const uint8_t* flags, uint64_t m, uint32_t n_namespaces,
                             uint32_t* counts, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (n_namespaces == 0) return NBG_OK;
  if (!counts || (m && (!ns_id || !name_hash || !label_hash || !flags))) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  CU(nbg_launch_count_running(ns_id, name_hash, label_hash, flags, m, n_namespaces, counts, (cudaStream_t)stream));
  ctx->launches += m ? 1 : 0;
  return NBG_OK;
}

int nbg_restart_block_device(nbg_ctx* ctx, int device, const nbg_restart_batch* in, uint64_t n, uint8_t* decision,
                             uint64_t* bitmap, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (!in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!decision || !bitmap || !in->flags) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  const unsigned bases[3] = {0, 0, 0};
  CU(nbg_launch_restart_block(in, n, bases, decision, bitmap, (cudaStream_t)stream));
  ctx->launches++;
  return NBG_OK;
}

int nbg_parse_records_device(nbg_ctx* ctx, int device, const nbg_strings* la, const nbg_strings* state, uint64_t R,
                             int64_t* sec, uint32_t* nsf, int32_t* zone, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (R == 0) return NBG_OK;
  if (!la || !la->off || !la->bytes || !sec || !nsf || (state && (!state->off || !state->bytes))) return NBG_E_INVAL;
  if (((uintptr_t)la->bytes & 15u) || (state && ((uintptr_t)state->bytes & 15u))) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  CU(nbg_launch_parse_records(la->off, la->bytes, state ? state->off : nullptr, state ? state->bytes : nullptr, R, 0, 0,
                              sec, nsf, zone, (cudaStream_t)stream));
  ctx->launches++;
  return NBG_OK;
}

int nbg_hash_strings_device(nbg_ctx* ctx, int device, const nbg_strings* strings, uint64_t R, uint64_t* out, void* stream) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (R == 0) return NBG_OK;
  if (!strings || !strings->off || !strings->bytes || !out) return NBG_E_INVAL;
  if ((uintptr_t)strings->bytes & 15u) return NBG_E_INVAL;
  CU(cudaSetDevice(ctx->devs[device]->id));
  CU(nbg_launch_hash_strings(strings->off, strings->bytes, R, 0, out, (cudaStream_t)stream));
  ctx->launches++;
  return NBG_OK;
}

// ------------------------------------------------------------ host-buffer API
int nbg_diff_generic(nbg_ctx* ctx, const nbg_generic_batch* in, uint64_t n, int mode, uint8_t* reason, uint64_t* bitmap) {
  if (!ctx || !in || (mode != NBG_GEN_DEEPEQUAL && mode != NBG_GEN_CONFIGMAP)) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap) return NBG_E_INVAL;
  if (!csr_ptrs_ok(in->labels_a, n) || !csr_ptrs_ok(in->labels_b, n) || !csr_ptrs_ok(in->spec_a, n) ||
      !csr_ptrs_ok(in->spec_b, n))
    return NBG_E_INVAL;
  return for_each_device(ctx, n, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kSvcChunk, ++c) {
      const uint64_t i1 = i0 + kSvcChunk < hi ? i0 + kSvcChunk : hi, m = i1 - i0;
      if (!csr_ok(in->labels_a, i0, i1) || !csr_ok(in->labels_b, i0, i1) || !csr_ok(in->spec_a, i0, i1) ||
          !csr_ok(in->spec_b, i0, i1))
        return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const CsrDev la = csr_in(lay, in->labels_a, i0, i1), lb = csr_in(lay, in->labels_b, i0, i1);
      const CsrDev sa = csr_in(lay, in->spec_a, i0, i1), sb = csr_in(lay, in->spec_b, i0, i1);
      const size_t o_reason = lay.out(reason + i0, m);
      const size_t o_bits = lay.out(bitmap + i0 / 64, (m + 63) / 64 * 8);
      if (int r = slot_upload(d, sl, lay)) return r;
      nbg_generic_batch b{};
      b.labels_a = csr_ptr(sl, la); b.labels_b = csr_ptr(sl, lb);
      b.spec_a = csr_ptr(sl, sa); b.spec_b = csr_ptr(sl, sb);
      const unsigned bases[4] = {la.base, lb.base, sa.base, sb.base};
      CU(nbg_launch_diff_generic(&b, m, mode, bases, (uint8_t*)(sl.dev + o_reason), (uint64_t*)(sl.dev + o_bits), d.s_run));
      ctx->launches++;
      if (int r = slot_download(d, sl, lay)) return r;
    }
    return NBG_OK;
  });
}

int nbg_diff_svc_gen(nbg_ctx* ctx, const nbg_svc_gen_batch* in, uint64_t n, uint8_t* reason, uint64_t* bitmap) {
  if (!ctx || !in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap || !in->name_hash || !in->target_port || !in->flags || !in->ports_to) return NBG_E_INVAL;
  if (!csr_ptrs_ok(in->labels_to, n) || !csr_ptrs_ok(in->ann_to, n) || !csr_ptrs_ok(in->sel_to, n)) return NBG_E_INVAL;
  return for_each_device(ctx, n, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kSvcChunk, ++c) {
      const uint64_t i1 = i0 + kSvcChunk < hi ? i0 + kSvcChunk : hi, m = i1 - i0;
      if (!csr_ok(in->labels_to, i0, i1) || !csr_ok(in->ann_to, i0, i1) || !csr_ok(in->sel_to, i0, i1)) return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const size_t a_nm = lay.in(in->name_hash + i0, m * 8), a_tp = lay.in(in->target_port + i0, m * 4);
      const size_t a_fl = lay.in(in->flags + i0, m), a_pt = lay.in(in->ports_to + i0 * NBG_SVC_SLOTS, m * NBG_SVC_SLOTS * 8);
      const CsrDev lt = csr_in(lay, in->labels_to, i0, i1), at = csr_in(lay, in->ann_to, i0, i1), st = csr_in(lay, in->sel_to, i0, i1);
      const size_t o_reason = lay.out(reason + i0, m), o_bits = lay.out(bitmap + i0 / 64, (m + 63) / 64 * 8);
      if (int r = slot_upload(d, sl, lay)) return r;
      nbg_svc_gen_batch b{};
      b.name_hash = (const uint64_t*)(sl.dev + a_nm); b.target_port = (const int32_t*)(sl.dev + a_tp);
      b.flags = (const uint8_t*)(sl.dev + a_fl); b.ports_to = (const uint64_t*)(sl.dev + a_pt);
      b.labels_to = csr_ptr(sl, lt); b.ann_to = csr_ptr(sl, at); b.sel_to = csr_ptr(sl, st);
      const unsigned bases[3] = {lt.base, at.base, st.base};
      CU(nbg_launch_diff_svc_gen(&b, m, bases, (uint8_t*)(sl.dev + o_reason), (uint64_t*)(sl.dev + o_bits), d.s_run));
      ctx->launches++;
      if (int r = slot_download(d, sl, lay)) return r;
    }
    return NBG_OK;
  });
}

int nbg_diff_vs_gen(nbg_ctx* ctx, const nbg_vs_gen_batch* in, const nbg_vs_gen_consts* consts, uint64_t n, uint8_t* reason,
                    uint64_t* bitmap) {
  if (!ctx || !in || !consts) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap || !in->prefix_hash || !in->rewrite_hash || !in->host_hash || !in->flags || !in->spec_to) return NBG_E_INVAL;
  if (!csr_ptrs_ok(in->hdr_from, n) || !csr_ptrs_ok(in->hdr_to, n)) return NBG_E_INVAL;
  return for_each_device(ctx, n, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kVsChunk, ++c) {
      const uint64_t i1 = i0 + kVsChunk < hi ? i0 + kVsChunk : hi, m = i1 - i0;
      if (!csr_ok(in->hdr_from, i0, i1) || !csr_ok(in->hdr_to, i0, i1)) return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const size_t a_p = lay.in(in->prefix_hash + i0, m * 8), a_r = lay.in(in->rewrite_hash + i0, m * 8), a_h = lay.in(in->host_hash + i0, m * 8);
      const size_t a_fl = lay.in(in->flags + i0, m), a_st = lay.in(in->spec_to + i0 * NBG_VS_SLOTS, m * NBG_VS_SLOTS * 8);
      const CsrDev hf = csr_in(lay, in->hdr_from, i0, i1), ht = csr_in(lay, in->hdr_to, i0, i1);
      const size_t o_reason = lay.out(reason + i0, m), o_bits = lay.out(bitmap + i0 / 64, (m + 63) / 64 * 8);
      if (int r = slot_upload(d, sl, lay)) return r;
      nbg_vs_gen_batch b{};
      b.prefix_hash = (const uint64_t*)(sl.dev + a_p); b.rewrite_hash = (const uint64_t*)(sl.dev + a_r);
      b.host_hash = (const uint64_t*)(sl.dev + a_h); b.flags = (const uint8_t*)(sl.dev + a_fl);
      b.spec_to = (const uint64_t*)(sl.dev + a_st); b.hdr_from = csr_ptr(sl, hf); b.hdr_to = csr_ptr(sl, ht);
      const unsigned bases[2] = {hf.base, ht.base};
      CU(nbg_launch_diff_vs_gen(&b, consts, m, bases, (uint8_t*)(sl.dev + o_reason), (uint64_t*)(sl.dev + o_bits), d.s_run));
      ctx->launches++;
      if (int r = slot_download(d, sl, lay)) return r;
    }
    return NBG_OK;
  });
}

// One pipeline for a whole step: the four batches of the same n notebooks are queued back to
// back, so the copy engine never drains between them, and the call waits once.
int nbg_step(nbg_ctx* ctx, const nbg_sts_sparse_batch* sts, const nbg_svc_gen_batch* svc, const nbg_vs_gen_batch* vs,
             const nbg_vs_gen_consts* vs_consts, const nbg_cull_batch* cull, const nbg_cull_params* params, uint64_t n,
             const nbg_step_outputs* out) {
  if (!ctx || !out) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  const int ndev = (int)ctx->devs.size();
  std::vector<std::unique_lock<std::mutex>> locks;
  for (int d = 0; d < ndev; ++d) locks.emplace_back(ctx->devs[d]->mu);
  int rc = NBG_OK;
  tl_batched = true;
  if (sts && rc == NBG_OK) rc = nbg_diff_sts_sparse(ctx, sts, n, out->sts_reason, out->sts_bitmap);
  if (svc && rc == NBG_OK) rc = nbg_diff_svc_gen(ctx, svc, n, out->svc_reason, out->svc_bitmap);
  if (vs && rc == NBG_OK) rc = nbg_diff_vs_gen(ctx, vs, vs_consts, n, out->vs_reason, out->vs_bitmap);
  if (cull && rc == NBG_OK) rc = nbg_cull(ctx, cull, n, params, out->action, out->argmax, out->cull_bitmap, out->period_bitmap);
  tl_batched = false;
  for (int d = 0; d < ndev; ++d) {
    cudaSetDevice(ctx->devs[d]->id);
    if (!cuda_ok(cudaStreamSynchronize(ctx->devs[d]->s_run), "cudaStreamSynchronize") && rc == NBG_OK) rc = NBG_E_CUDA;
    if (!cuda_ok(cudaStreamSynchronize(ctx->devs[d]->s_in), "cudaStreamSynchronize") && rc == NBG_OK) rc = NBG_E_CUDA;
  }
  if (rc == NBG_OK && sts) {
    unsigned any = 0;
    for (uint64_t i = 0; i < n; ++i) any |= out->sts_reason[i];
    if (any & NBG_STS_NILREPLICAS) rc = NBG_E_NILREPLICAS;
  }
  return rc;
}

// One step over RESIDENT batches.  The StatefulSet kernel streams ~0.8 KB per notebook and owns
// `stream`; the three small predicates are latency-bound on their own, so they go to the context's
// side streams, forked from and joined back into `stream` with events: to the caller the whole step
// is ordered on `stream` like a single launch.
int nbg_step_device(nbg_ctx* ctx, int device, const nbg_sts_sparse_batch* sts, const nbg_svc_gen_batch* svc,
                    const nbg_vs_gen_batch* vs, const nbg_vs_gen_consts* vs_consts, const nbg_cull_batch* cull,
                    const nbg_cull_params* p, uint64_t n, const nbg_step_outputs* out, void* stream,
                    void* ev_sts_start, void* ev_sts_stop) {
  if (int rc = check_dev(ctx, device)) return rc;
  if (!out || (vs && !vs_consts) || (cull && !p)) return NBG_E_INVAL;
  if (n == 0 || !(sts || svc || vs || cull)) return NBG_OK;
  if ((sts && (!out->sts_reason || !out->sts_bitmap)) || (svc && (!out->svc_reason || !out->svc_bitmap)) ||
      (vs && (!out->vs_reason || !out->vs_bitmap)) ||
      (cull && (!out->action || !out->argmax || !out->cull_bitmap || !out->period_bitmap)))
    return NBG_E_INVAL;
  if (cull) {
    if (p->now_nsec < 0 || p->now_nsec > 999999999 || (p->mode != NBG_CULL_PREDICATES && p->mode != NBG_CULL_RECONCILE))
      return NBG_E_INVAL;
    if (p->cull_idle_min > NBG_CULL_MAX_MINUTES || p->cull_idle_min < -NBG_CULL_MAX_MINUTES ||
        p->check_period_min > NBG_CULL_MAX_MINUTES || p->check_period_min < -NBG_CULL_MAX_MINUTES)
      return NBG_E_INVAL;
  }
  Dev& d = *ctx->devs[device];
  std::unique_lock<std::mutex> lock(d.mu);             // the fork/join events are per device: one enqueue at a time
  CU(cudaSetDevice(d.id));
  cudaStream_t s = (cudaStream_t)stream;
  const unsigned zero[6] = {0, 0, 0, 0, 0, 0};
  CU(cudaEventRecord(d.ev_fork, s));
  if (sts) {                                           // first in line for the SMs: it is the long one
    if (ev_sts_start) CU(cudaEventRecord((cudaEvent_t)ev_sts_start, s));
    CU(nbg_launch_diff_sts_sparse(sts, n, zero, out->sts_reason, out->sts_bitmap, s));
    if (ev_sts_stop) CU(cudaEventRecord((cudaEvent_t)ev_sts_stop, s));
    ctx->launches++;
  }
  for (int k = 0; k < 3; ++k) {
    if (!(k == 0 ? (const void*)svc : k == 1 ? (const void*)vs : (const void*)cull)) continue;
    CU(cudaStreamWaitEvent(d.s_side[k], d.ev_fork, 0));
    if (k == 0) CU(nbg_launch_diff_svc_gen(svc, n, zero, out->svc_reason, out->svc_bitmap, d.s_side[k]));
    if (k == 1) CU(nbg_launch_diff_vs_gen(vs, vs_consts, n, zero, out->vs_reason, out->vs_bitmap, d.s_side[k]));
    if (k == 2) CU(nbg_launch_cull(cull, n, p, 0, 0, out->action, out->argmax, out->cull_bitmap, out->period_bitmap, d.s_side[k]));
    CU(cudaEventRecord(d.ev_join[k], d.s_side[k]));
    CU(cudaStreamWaitEvent(s, d.ev_join[k], 0));
    ctx->launches++;
  }
  return NBG_OK;
}

int nbg_count_running(nbg_ctx* ctx, const uint32_t* ns_id, const uint64_t* name_hash, const uint64_t* label_hash,
                      const uint8_t* flags, uint64_t m, uint32_t n_namespaces, uint32_t* counts) {
  if (!ctx) return NBG_E_INVAL;
  if (n_namespaces == 0) return NBG_OK;
  if (!counts || (m && (!ns_id || !name_hash || !label_hash || !flags))) return NBG_E_INVAL;
  Dev& d = *ctx->devs[0];                              // one table of counts: a single device
  std::unique_lock<std::mutex> lock(d.mu);
  CU(cudaSetDevice(d.id));
  Slot& sl = d.slot[0];
  Layout lay;
  const size_t a_ns = lay.in(ns_id, m * 4), a_nm = lay.in(name_hash, m * 8), a_lb = lay.in(label_hash, m * 8);
  const size_t a_fl = lay.in(flags, m);
  const size_t o_c = lay.out(counts, (size_t)n_namespaces * 4);
  if (int rc = slot_upload(d, sl, lay)) return rc;
  CU(nbg_launch_count_running((const uint32_t*)(sl.dev + a_ns), (const uint64_t*)(sl.dev + a_nm), (const uint64_t*)(sl.dev + a_lb),
                              (const uint8_t*)(sl.dev + a_fl), m, n_namespaces, (uint32_t*)(sl.dev + o_c), d.s_run));
  ctx->launches += m ? 1 : 0;
  if (int rc = slot_download(d, sl, lay)) return rc;
  CU(cudaStreamSynchronize(d.s_run));
  return NBG_OK;
}

int nbg_restart_block(nbg_ctx* ctx, const nbg_restart_batch* in, uint64_t n, uint8_t* decision, uint64_t* bitmap) {
  if (!ctx || !in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!decision || !bitmap || !in->flags) return NBG_E_INVAL;
  const nbg_podsparse* sp[3] = {&in->old_spec, &in->updated_spec, &in->mutated_spec};
  for (const nbg_podsparse* p : sp) {
    if (!p->mask || !p->off || p->end || (p->off[n] > p->off[0] && !p->val) || ((uintptr_t)p->val & 15u)) return NBG_E_INVAL;
  }
  return for_each_device(ctx, n, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kStsSparseChunk, ++c) {
      const uint64_t i1 = i0 + kStsSparseChunk < hi ? i0 + kStsSparseChunk : hi, m = i1 - i0;
      for (const nbg_podsparse* p : sp) {
        if (!off_ok(p->off, i0, i1) || !sparse_ok(*p, i0, i1)) return NBG_E_RANGE;
      }
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const size_t a_fl = lay.in(in->flags + i0, m);
      size_t a_m[3], a_o[3], a_v[3];
      unsigned bases[3];
      for (int k = 0; k < 3; ++k) {
        bases[k] = sp[k]->off[i0];
        const size_t nv = sp[k]->off[i1] - bases[k];
        a_m[k] = lay.in(sp[k]->mask + i0 * 3, m * 24);
        a_o[k] = lay.in(sp[k]->off + i0, (m + 1) * 4);
        a_v[k] = lay.in(nv ? sp[k]->val + bases[k] : nullptr, nv * 8);
      }
      const size_t o_dec = lay.out(decision + i0, m);
      const size_t o_bits = lay.out(bitmap + i0 / 64, (m + 63) / 64 * 8);
      if (int r = slot_upload(d, sl, lay)) return r;
      nbg_restart_batch b{};
      b.flags = (const uint8_t*)(sl.dev + a_fl);
      nbg_podsparse* dst[3] = {&b.old_spec, &b.updated_spec, &b.mutated_spec};
      for (int k = 0; k < 3; ++k) {
        dst[k]->mask = (const uint64_t*)(sl.dev + a_m[k]);
        dst[k]->off = (const uint32_t*)(sl.dev + a_o[k]);
        dst[k]->val = (const uint64_t*)(sl.dev + a_v[k]);
      }
      CU(nbg_launch_restart_block(&b, m, bases, (uint8_t*)(sl.dev + o_dec), (uint64_t*)(sl.dev + o_bits), d.s_run));
      ctx->launches++;
      if (int r = slot_download(d, sl, lay)) return r;
    }
    return NBG_OK;
  });
}

int nbg_parse_records(nbg_ctx* ctx, const nbg_strings* la, const nbg_strings* state, uint64_t R, int64_t* sec,
                      uint32_t* nsf, int32_t* zone) {
  if (!ctx || !la) return NBG_E_INVAL;
  if (R == 0) return NBG_OK;
  if (!la->off || !sec || !nsf || (state && !state->off)) return NBG_E_INVAL;
  if ((la->off[R] > la->off[0] && !la->bytes) || (state && state->off[R] > state->off[0] && !state->bytes)) return NBG_E_INVAL;
  return for_each_device(ctx, R, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kTextChunk, ++c) {
      const uint64_t i1 = i0 + kTextChunk < hi ? i0 + kTextChunk : hi, m = i1 - i0;
      if (!off_ok(la->off, i0, i1) || (state && !off_ok(state->off, i0, i1))) return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const unsigned tb = la->off[i0], sb = state ? state->off[i0] : 0;
      const size_t nt = la->off[i1] - tb, ns = state ? state->off[i1] - sb : 0;
      const size_t a_to = lay.in(la->off + i0, (m + 1) * 4), a_tx = lay.in(nt ? la->bytes + tb : nullptr, nt);
      const size_t a_so = state ? lay.in(state->off + i0, (m + 1) * 4) : 0, a_sx = state ? lay.in(ns ? state->bytes + sb : nullptr, ns) : 0;
      const size_t o_sec = lay.out(sec + i0, m * 8), o_nsf = lay.out(nsf + i0, m * 4);
      const size_t o_zone = zone ? lay.out(zone + i0, m * 4) : 0;
      if (int r = slot_upload(d, sl, lay)) return r;
      CU(nbg_launch_parse_records((const uint32_t*)(sl.dev + a_to), (const unsigned char*)(sl.dev + a_tx),
                                  state ? (const uint32_t*)(sl.dev + a_so) : nullptr,
                                  state ? (const unsigned char*)(sl.dev + a_sx) : nullptr, m, tb, sb,
                                  (int64_t*)(sl.dev + o_sec), (uint32_t*)(sl.dev + o_nsf),
                                  zone ? (int32_t*)(sl.dev + o_zone) : nullptr, d.s_run));
      ctx->launches++;
      if (int r = slot_download(d, sl, lay)) return r;
    }
    return NBG_OK;
  });
}

int nbg_cull_text(nbg_ctx* ctx, const nbg_cull_text_batch* in, uint64_t n, const nbg_cull_params* p, uint8_t* action,
                  uint32_t* argmax, int32_t* argmax_zone, uint64_t* cull_bitmap, uint64_t* period_bitmap) {
  if (!ctx || !in || !p) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!action || !argmax || !cull_bitmap || !period_bitmap || !in->k_off || !in->t_off || !in->nb_flags ||
      !in->ann_text.off || !in->chk_text.off)
    return NBG_E_INVAL;
  if (p->now_nsec < 0 || p->now_nsec > 999999999 || (p->mode != NBG_CULL_PREDICATES && p->mode != NBG_CULL_RECONCILE))
    return NBG_E_INVAL;
  if (p->cull_idle_min > NBG_CULL_MAX_MINUTES || p->cull_idle_min < -NBG_CULL_MAX_MINUTES ||
      p->check_period_min > NBG_CULL_MAX_MINUTES || p->check_period_min < -NBG_CULL_MAX_MINUTES)
    return NBG_E_INVAL;
  if ((in->k_off[n] > in->k_off[0] && (!in->k_last_activity.off || !in->k_state.off)) ||
      (in->t_off[n] > in->t_off[0] && !in->t_last_activity.off))
    return NBG_E_INVAL;

  return for_each_device(ctx, n, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kCullChunk, ++c) {
      const uint64_t i1 = i0 + kCullChunk < hi ? i0 + kCullChunk : hi, m = i1 - i0;
      if (!off_ok(in->k_off, i0, i1) || !off_ok(in->t_off, i0, i1) || !off_ok(in->ann_text.off, i0, i1) ||
          !off_ok(in->chk_text.off, i0, i1))
        return NBG_E_RANGE;
      const uint64_t kb = in->k_off[i0], ke = in->k_off[i1], tb = in->t_off[i0], te = in->t_off[i1];
      const uint64_t nk = ke - kb, nt = te - tb;
      if ((nk && (!off_ok(in->k_last_activity.off, kb, ke) || !off_ok(in->k_state.off, kb, ke))) ||
          (nt && !off_ok(in->t_last_activity.off, tb, te)))
        return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      // a string-table slice: offsets [r0..r1] and the bytes they span
      struct Str { size_t off_at = 0, bytes_at = 0; unsigned base = 0; };
      auto str_in = [&](const nbg_strings& t, uint64_t r0, uint64_t r1) {
        Str s;
        if (r1 == r0) { s.off_at = lay.scratch(8); s.bytes_at = lay.scratch(16); return s; }
        s.base = t.off[r0];
        s.off_at = lay.in(t.off + r0, (r1 - r0 + 1) * 4);
        const size_t nb = t.off[r1] - s.base;
        s.bytes_at = lay.in(nb ? t.bytes + s.base : nullptr, nb);
        return s;
      };
      const size_t a_koff = lay.in(in->k_off + i0, (m + 1) * 4), a_toff = lay.in(in->t_off + i0, (m + 1) * 4);
      const Str kla = str_in(in->k_last_activity, kb, ke), kst = str_in(in->k_state, kb, ke);
      const Str tla = str_in(in->t_last_activity, tb, te);
      const Str ann = str_in(in->ann_text, i0, i1), chk = str_in(in->chk_text, i0, i1);
      const size_t a_flags = lay.in(in->nb_flags + i0, m);
      // device-only intermediates: what the host packer would have produced
      const size_t s_ksec = lay.scratch(nk * 8), s_knsf = lay.scratch(nk * 4), s_kzone = lay.scratch(nk * 4);
      const size_t s_tsec = lay.scratch(nt * 8), s_tnsf = lay.scratch(nt * 4), s_tzone = lay.scratch(nt * 4);
      const size_t s_asec = lay.scratch(m * 8), s_ansf = lay.scratch(m * 4), s_csec = lay.scratch(m * 8), s_cnsf = lay.scratch(m * 4);
      const size_t s_nbf = lay.scratch(m);
      const size_t words = (m + 63) / 64;
      const size_t o_act = lay.out(action + i0, m), o_arg = lay.out(argmax + i0, m * 4);
      const size_t o_zone = argmax_zone ? lay.out(argmax_zone + i0, m * 4) : lay.scratch(m * 4);
      const size_t o_cull = lay.out(cull_bitmap + i0 / 64, words * 8), o_per = lay.out(period_bitmap + i0 / 64, words * 8);
      if (int rc = slot_upload(d, sl, lay)) return rc;
      char* D = sl.dev;
      auto u32 = [&](size_t at) { return (uint32_t*)(D + at); };
      if (nk) CU(nbg_launch_parse_records(u32(kla.off_at), (const unsigned char*)(D + kla.bytes_at), u32(kst.off_at),
                                          (const unsigned char*)(D + kst.bytes_at), nk, kla.base, kst.base,
                                          (int64_t*)(D + s_ksec), u32(s_knsf), (int32_t*)(D + s_kzone), d.s_run));
      if (nt) CU(nbg_launch_parse_records(u32(tla.off_at), (const unsigned char*)(D + tla.bytes_at), nullptr, nullptr, nt,
                                          tla.base, 0, (int64_t*)(D + s_tsec), u32(s_tnsf), (int32_t*)(D + s_tzone), d.s_run));
      CU(nbg_launch_parse_records(u32(ann.off_at), (const unsigned char*)(D + ann.bytes_at), nullptr, nullptr, m, ann.base, 0,
                                  (int64_t*)(D + s_asec), u32(s_ansf), nullptr, d.s_run));
      CU(nbg_launch_parse_records(u32(chk.off_at), (const unsigned char*)(D + chk.bytes_at), nullptr, nullptr, m, chk.base, 0,
                                  (int64_t*)(D + s_csec), u32(s_cnsf), nullptr, d.s_run));
      CU(nbg_launch_cull_text_fixup((const uint8_t*)(D + a_flags), m, u32(s_ansf), u32(s_cnsf), (uint8_t*)(D + s_nbf), d.s_run));
      nbg_cull_batch b{};
      b.k_off = u32(a_koff); b.k_sec = (const int64_t*)(D + s_ksec); b.k_nsf = u32(s_knsf);
      b.t_off = u32(a_toff); b.t_sec = (const int64_t*)(D + s_tsec); b.t_nsf = u32(s_tnsf);
      b.ann_sec = (const int64_t*)(D + s_asec); b.ann_nsf = u32(s_ansf);
      b.chk_sec = (const int64_t*)(D + s_csec); b.chk_nsf = u32(s_cnsf);
      b.nb_flags = (const uint8_t*)(D + s_nbf);
      CU(nbg_launch_cull(&b, m, p, (unsigned)kb, (unsigned)tb, (uint8_t*)(D + o_act), u32(o_arg), (uint64_t*)(D + o_cull),
                         (uint64_t*)(D + o_per), d.s_run));
      CU(nbg_launch_cull_text_zone((const uint8_t*)(D + o_act), u32(o_arg), u32(a_koff), (const int32_t*)(D + s_kzone), (unsigned)kb,
                                   u32(a_toff), (const int32_t*)(D + s_tzone), (unsigned)tb, m, (int32_t*)(D + o_zone), d.s_run));
      ctx->launches += 5 + (nk ? 1 : 0) + (nt ? 1 : 0);
      if (int rc = slot_download(d, sl, lay)) return rc;
    }
    return NBG_OK;
  });
}

int nbg_hash_strings(nbg_ctx* ctx, const nbg_strings* strings, uint64_t R, uint64_t* out) {
  if (!ctx || !strings) return NBG_E_INVAL;
  if (R == 0) return NBG_OK;
  if (!strings->off || !out || (strings->off[R] > strings->off[0] && !strings->bytes)) return NBG_E_INVAL;
  return for_each_device(ctx, R, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kTextChunk, ++c) {
      const uint64_t i1 = i0 + kTextChunk < hi ? i0 + kTextChunk : hi, m = i1 - i0;
      if (!off_ok(strings->off, i0, i1)) return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const unsigned b0 = strings->off[i0];
      const size_t nb = strings->off[i1] - b0;
      const size_t a_o = lay.in(strings->off + i0, (m + 1) * 4), a_x = lay.in(nb ? strings->bytes + b0 : nullptr, nb);
      const size_t o_h = lay.out(out + i0, m * 8);
      if (int r = slot_upload(d, sl, lay)) return r;
      CU(nbg_launch_hash_strings((const uint32_t*)(sl.dev + a_o), (const unsigned char*)(sl.dev + a_x), m, b0,
                                 (uint64_t*)(sl.dev + o_h), d.s_run));
      ctx->launches++;
      if (int r = slot_download(d, sl, lay)) return r;
    }
    return NBG_OK;
  });
}

int nbg_owner_join(nbg_ctx* ctx, const uint64_t* nb_key, uint64_t n, const uint64_t* sts_key, uint64_t m,
                   uint32_t* match) {
  if (!ctx) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!nb_key || !match || (m && !sts_key)) return NBG_E_INVAL;
  if (m >= (1ull << 31)) return NBG_E_TOOBIG;
  // One device: the join is a single table; a multi-device context uses its first device.
  Dev& d = *ctx->devs[0];
  std::unique_lock<std::mutex> lock(d.mu);
  CU(cudaSetDevice(d.id));
  const size_t need = nbg_owner_join_scratch_bytes(m);
  if (d.scratch_cap < need) {
    CU(cudaStreamSynchronize(d.s_run));
    if (d.scratch) { CU(cudaFree(d.scratch)); d.scratch = nullptr; d.scratch_cap = 0; }
    if (!cuda_ok(cudaMalloc(&d.scratch, need), "cudaMalloc(join table)")) { d.scratch = nullptr; return NBG_E_NOMEM; }
    d.scratch_cap = need;
  }
  Slot& sl = d.slot[0];
  Layout lay;
  const size_t a_nb = lay.in(nb_key, n * 8), a_sts = lay.in(sts_key, m * 8);
  const size_t o_match = lay.out(match, n * 4);
  if (int rc = slot_upload(d, sl, lay)) return rc;
  CU(nbg_launch_owner_join((const unsigned long long*)(sl.dev + a_nb), n, (const unsigned long long*)(sl.dev + a_sts), m,
                           (uint32_t*)(sl.dev + o_match), d.scratch, d.s_run));
  ctx->launches += m ? 2 : 1;
  if (int rc = slot_download(d, sl, lay)) return rc;
  CU(cudaStreamSynchronize(d.s_run));
  return NBG_OK;
}

int nbg_cull(nbg_ctx* ctx, const nbg_cull_batch* in, uint64_t n, const nbg_cull_params* p,
             uint8_t* action, uint32_t* argmax, uint64_t* cull_bitmap, uint64_t* period_bitmap) {
  if (!ctx || !in || !p) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!action || !argmax || !cull_bitmap || !period_bitmap) return NBG_E_INVAL;
  if (!in->ann_sec || !in->ann_nsf || !in->chk_sec || !in->chk_nsf || !in->nb_flags) return NBG_E_INVAL;
  if (p->now_nsec < 0 || p->now_nsec > 999999999 || (p->mode != NBG_CULL_PREDICATES && p->mode != NBG_CULL_RECONCILE))
    return NBG_E_INVAL;
  if (p->cull_idle_min > NBG_CULL_MAX_MINUTES || p->cull_idle_min < -NBG_CULL_MAX_MINUTES ||
      p->check_period_min > NBG_CULL_MAX_MINUTES || p->check_period_min < -NBG_CULL_MAX_MINUTES)
    return NBG_E_INVAL;
  if (!in->k_off || !in->t_off || in->k_end || in->t_end) return NBG_E_INVAL;
  if ((in->k_off[n] > in->k_off[0] && (!in->k_sec || !in->k_nsf)) ||
      (in->t_off[n] > in->t_off[0] && (!in->t_sec || !in->t_nsf)))
    return NBG_E_INVAL;

  return for_each_device(ctx, n, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kCullChunk, ++c) {
      const uint64_t i1 = i0 + kCullChunk < hi ? i0 + kCullChunk : hi, m = i1 - i0;
      if (!off_ok(in->k_off, i0, i1) || !off_ok(in->t_off, i0, i1)) return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const unsigned kb = in->k_off[i0], tb = in->t_off[i0];
      const size_t nk = in->k_off[i1] - kb, nt = in->t_off[i1] - tb;
      const size_t a_koff = lay.in(in->k_off + i0, (m + 1) * 4);
      const size_t a_ksec = lay.in(nk ? in->k_sec + kb : nullptr, nk * 8);
      const size_t a_knsf = lay.in(nk ? in->k_nsf + kb : nullptr, nk * 4);
      const size_t a_toff = lay.in(in->t_off + i0, (m + 1) * 4);
      const size_t a_tsec = lay.in(nt ? in->t_sec + tb : nullptr, nt * 8);
      const size_t a_tnsf = lay.in(nt ? in->t_nsf + tb : nullptr, nt * 4);
      const size_t a_asec = lay.in(in->ann_sec + i0, m * 8);
      const size_t a_ansf = lay.in(in->ann_nsf + i0, m * 4);
      const size_t a_csec = lay.in(in->chk_sec + i0, m * 8);
      const size_t a_cnsf = lay.in(in->chk_nsf + i0, m * 4);
      const size_t a_flag = lay.in(in->nb_flags + i0, m);
      const size_t words = (m + 63) / 64;
      const size_t o_act = lay.out(action + i0, m);
      const size_t o_arg = lay.out(argmax + i0, m * 4);
      const size_t o_cull = lay.out(cull_bitmap + i0 / 64, words * 8);
      const size_t o_per = lay.out(period_bitmap + i0 / 64, words * 8);
      if (int rc = slot_upload(d, sl, lay)) return rc;
      nbg_cull_batch b{};
      b.k_off = (const uint32_t*)(sl.dev + a_koff); b.k_sec = (const int64_t*)(sl.dev + a_ksec);
      b.k_nsf = (const uint32_t*)(sl.dev + a_knsf); b.t_off = (const uint32_t*)(sl.dev + a_toff);
      b.t_sec = (const int64_t*)(sl.dev + a_tsec); b.t_nsf = (const uint32_t*)(sl.dev + a_tnsf);
      b.ann_sec = (const int64_t*)(sl.dev + a_asec); b.ann_nsf = (const uint32_t*)(sl.dev + a_ansf);
      b.chk_sec = (const int64_t*)(sl.dev + a_csec); b.chk_nsf = (const uint32_t*)(sl.dev + a_cnsf);
      b.nb_flags = (const uint8_t*)(sl.dev + a_flag);
      CU(nbg_launch_cull(&b, m, p, kb, tb, (uint8_t*)(sl.dev + o_act), (uint32_t*)(sl.dev + o_arg),
                         (uint64_t*)(sl.dev + o_cull), (uint64_t*)(sl.dev + o_per), d.s_run));
      ctx->launches++;
      if (int rc = slot_download(d, sl, lay)) return rc;
    }
    return NBG_OK;
  });
}

int nbg_diff_sts(nbg_ctx* ctx, const nbg_sts_batch* in, uint64_t n, uint8_t* reason, uint64_t* bitmap) {
  if (!ctx || !in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap || !in->replicas_from || !in->replicas_to || !in->tmpl_from || !in->tmpl_to ||
      !in->pod_from || !in->pod_to)
    return NBG_E_INVAL;
  if (!csr_ptrs_ok(in->labels_from, n) || !csr_ptrs_ok(in->labels_to, n) || !csr_ptrs_ok(in->ann_from, n) ||
      !csr_ptrs_ok(in->ann_to, n))
    return NBG_E_INVAL;

  int rc = for_each_device(ctx, n, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kStsChunk, ++c) {
      const uint64_t i1 = i0 + kStsChunk < hi ? i0 + kStsChunk : hi, m = i1 - i0;
      if (!csr_ok(in->labels_from, i0, i1) || !csr_ok(in->labels_to, i0, i1) || !csr_ok(in->ann_from, i0, i1) ||
          !csr_ok(in->ann_to, i0, i1))
        return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const size_t a_rf = lay.in(in->replicas_from + i0, m * 4);
      const size_t a_rt = lay.in(in->replicas_to + i0, m * 4);
      const size_t a_tf = lay.in(in->tmpl_from + i0, m * 8);
      const size_t a_tt = lay.in(in->tmpl_to + i0, m * 8);
      const size_t a_pf = lay.in(in->pod_from + i0 * NBG_POD_SLOTS, m * NBG_POD_SLOTS * 8);
      const size_t a_pt = lay.in(in->pod_to + i0 * NBG_POD_SLOTS, m * NBG_POD_SLOTS * 8);
      const CsrDev lf = csr_in(lay, in->labels_from, i0, i1), lt = csr_in(lay, in->labels_to, i0, i1);
      const CsrDev af = csr_in(lay, in->ann_from, i0, i1), at = csr_in(lay, in->ann_to, i0, i1);
      const size_t o_reason = lay.out(reason + i0, m);
      const size_t o_bits = lay.out(bitmap + i0 / 64, (m + 63) / 64 * 8);
      if (int r = slot_upload(d, sl, lay)) return r;
      nbg_sts_batch b{};
      b.replicas_from = (const int32_t*)(sl.dev + a_rf); b.replicas_to = (const int32_t*)(sl.dev + a_rt);
      b.tmpl_from = (const uint64_t*)(sl.dev + a_tf); b.tmpl_to = (const uint64_t*)(sl.dev + a_tt);
      b.pod_from = (const uint64_t*)(sl.dev + a_pf); b.pod_to = (const uint64_t*)(sl.dev + a_pt);
      b.labels_from = csr_ptr(sl, lf); b.labels_to = csr_ptr(sl, lt);
      b.ann_from = csr_ptr(sl, af); b.ann_to = csr_ptr(sl, at);
      const unsigned bases[4] = {lf.base, lt.base, af.base, at.base};
      CU(nbg_launch_diff_sts(&b, m, bases, tl_deployment ? 1 : 0, (uint8_t*)(sl.dev + o_reason), (uint64_t*)(sl.dev + o_bits), d.s_run));
      ctx->launches++;
      if (int r = slot_download(d, sl, lay)) return r;
    }
    return NBG_OK;
  });
  if (rc != NBG_OK) return rc;
  unsigned any = 0;
  for (uint64_t i = 0; i < n; ++i) any |= reason[i];
  return (any & NBG_STS_NILREPLICAS) ? NBG_E_NILREPLICAS : NBG_OK;
}

int nbg_diff_sts_gen(nbg_ctx* ctx, const nbg_sts_batch* in, const nbg_sts_gen_inputs* gen, const nbg_sts_gen_consts* consts,
                     uint64_t n, uint8_t* reason, uint64_t* bitmap) {
  if (!ctx || !in || !gen || !consts) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap || !in->replicas_to || !in->tmpl_from || !in->tmpl_to || !in->pod_from || !in->pod_to ||
      !gen->prefix_hash || !gen->flags)
    return NBG_E_INVAL;
  if (!csr_ptrs_ok(in->labels_from, n) || !csr_ptrs_ok(in->labels_to, n) || !csr_ptrs_ok(in->ann_to, n)) return NBG_E_INVAL;
  int rc = for_each_device(ctx, n, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kStsChunk, ++c) {
      const uint64_t i1 = i0 + kStsChunk < hi ? i0 + kStsChunk : hi, m = i1 - i0;
      if (!csr_ok(in->labels_from, i0, i1) || !csr_ok(in->labels_to, i0, i1) || !csr_ok(in->ann_to, i0, i1)) return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const size_t a_rt = lay.in(in->replicas_to + i0, m * 4);
      const size_t a_tf = lay.in(in->tmpl_from + i0, m * 8), a_tt = lay.in(in->tmpl_to + i0, m * 8);
      const size_t a_pf = lay.in(in->pod_from + i0 * NBG_POD_SLOTS, m * NBG_POD_SLOTS * 8);
      const size_t a_pt = lay.in(in->pod_to + i0 * NBG_POD_SLOTS, m * NBG_POD_SLOTS * 8);
      const size_t a_ph = lay.in(gen->prefix_hash + i0, m * 8), a_gf = lay.in(gen->flags + i0, m);
      const CsrDev lf = csr_in(lay, in->labels_from, i0, i1), lt = csr_in(lay, in->labels_to, i0, i1);
      const CsrDev at = csr_in(lay, in->ann_to, i0, i1);
      const size_t o_reason = lay.out(reason + i0, m);
      const size_t o_bits = lay.out(bitmap + i0 / 64, (m + 63) / 64 * 8);
      if (int r = slot_upload(d, sl, lay)) return r;
      nbg_sts_batch b{};
      b.replicas_to = (const int32_t*)(sl.dev + a_rt);
      b.tmpl_from = (const uint64_t*)(sl.dev + a_tf); b.tmpl_to = (const uint64_t*)(sl.dev + a_tt);
      b.pod_from = (const uint64_t*)(sl.dev + a_pf); b.pod_to = (const uint64_t*)(sl.dev + a_pt);
      b.labels_from = csr_ptr(sl, lf); b.labels_to = csr_ptr(sl, lt); b.ann_to = csr_ptr(sl, at);
      const unsigned bases[3] = {lf.base, lt.base, at.base};
      CU(nbg_launch_diff_sts_gen(&b, (const uint64_t*)(sl.dev + a_ph), (const uint8_t*)(sl.dev + a_gf), consts, m, bases,
                                 (uint8_t*)(sl.dev + o_reason), (uint64_t*)(sl.dev + o_bits), d.s_run));
      ctx->launches++;
      if (int r = slot_download(d, sl, lay)) return r;
    }
    return NBG_OK;
  });
  if (rc != NBG_OK) return rc;
  unsigned any = 0;
  for (uint64_t i = 0; i < n; ++i) any |= reason[i];
  return (any & NBG_STS_NILREPLICAS) ? NBG_E_NILREPLICAS : NBG_OK;
}

int nbg_diff_sts_sparse(nbg_ctx* ctx, const nbg_sts_sparse_batch* in, uint64_t n, uint8_t* reason, uint64_t* bitmap) {
  if (!ctx || !in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap || !in->replicas_from || !in->replicas_to || !in->tmpl_from || !in->tmpl_to ||
      !in->pod_from.mask || !in->pod_to.mask)
    return NBG_E_INVAL;
  if (!csr_ptrs_ok(in->labels_from, n) || !csr_ptrs_ok(in->labels_to, n) || !csr_ptrs_ok(in->ann_from, n) ||
      !csr_ptrs_ok(in->ann_to, n) || !in->pod_from.off || !in->pod_to.off || in->pod_from.end || in->pod_to.end)
    return NBG_E_INVAL;
  if ((in->pod_from.off[n] > in->pod_from.off[0] && !in->pod_from.val) ||
      (in->pod_to.off[n] > in->pod_to.off[0] && !in->pod_to.val))
    return NBG_E_INVAL;
  if (((uintptr_t)in->pod_from.val | (uintptr_t)in->pod_to.val) & 15u) return NBG_E_INVAL;   // 128-bit loads

  int rc = for_each_device(ctx, n, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kStsSparseChunk, ++c) {
      const uint64_t i1 = i0 + kStsSparseChunk < hi ? i0 + kStsSparseChunk : hi, m = i1 - i0;
      if (!csr_ok(in->labels_from, i0, i1) || !csr_ok(in->labels_to, i0, i1) || !csr_ok(in->ann_from, i0, i1) ||
          !csr_ok(in->ann_to, i0, i1) || !off_ok(in->pod_from.off, i0, i1) || !off_ok(in->pod_to.off, i0, i1))
        return NBG_E_RANGE;
      if (!sparse_ok(in->pod_from, i0, i1) || !sparse_ok(in->pod_to, i0, i1)) return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const size_t a_rf = lay.in(in->replicas_from + i0, m * 4), a_rt = lay.in(in->replicas_to + i0, m * 4);
      const size_t a_tf = lay.in(in->tmpl_from + i0, m * 8), a_tt = lay.in(in->tmpl_to + i0, m * 8);
      const size_t a_mf = lay.in(in->pod_from.mask + i0 * 3, m * 24), a_mt = lay.in(in->pod_to.mask + i0 * 3, m * 24);
      const size_t a_of = lay.in(in->pod_from.off + i0, (m + 1) * 4), a_ot = lay.in(in->pod_to.off + i0, (m + 1) * 4);
      const unsigned vfb = in->pod_from.off[i0], vtb = in->pod_to.off[i0];
      const size_t nvf = in->pod_from.off[i1] - vfb, nvt = in->pod_to.off[i1] - vtb;
      const size_t a_vf = lay.in(nvf ? in->pod_from.val + vfb : nullptr, nvf * 8);
      const size_t a_vt = lay.in(nvt ? in->pod_to.val + vtb : nullptr, nvt * 8);
      const CsrDev lf = csr_in(lay, in->labels_from, i0, i1), lt = csr_in(lay, in->labels_to, i0, i1);
      const CsrDev af = csr_in(lay, in->ann_from, i0, i1), at = csr_in(lay, in->ann_to, i0, i1);
      const size_t o_reason = lay.out(reason + i0, m);
      const size_t o_bits = lay.out(bitmap + i0 / 64, (m + 63) / 64 * 8);
      if (int r = slot_upload(d, sl, lay)) return r;
      nbg_sts_sparse_batch b{};
      b.replicas_from = (const int32_t*)(sl.dev + a_rf); b.replicas_to = (const int32_t*)(sl.dev + a_rt);
      b.tmpl_from = (const uint64_t*)(sl.dev + a_tf); b.tmpl_to = (const uint64_t*)(sl.dev + a_tt);
      b.pod_from.mask = (const uint64_t*)(sl.dev + a_mf); b.pod_to.mask = (const uint64_t*)(sl.dev + a_mt);
      b.pod_from.off = (const uint32_t*)(sl.dev + a_of); b.pod_to.off = (const uint32_t*)(sl.dev + a_ot);
      b.pod_from.val = (const uint64_t*)(sl.dev + a_vf); b.pod_to.val = (const uint64_t*)(sl.dev + a_vt);
      b.labels_from = csr_ptr(sl, lf); b.labels_to = csr_ptr(sl, lt);
      b.ann_from = csr_ptr(sl, af); b.ann_to = csr_ptr(sl, at);
      const unsigned bases[6] = {lf.base, lt.base, af.base, at.base, vfb, vtb};
      CU(nbg_launch_diff_sts_sparse(&b, m, bases, (uint8_t*)(sl.dev + o_reason), (uint64_t*)(sl.dev + o_bits), d.s_run));
      ctx->launches++;
      if (int r = slot_download(d, sl, lay)) return r;
    }
    return NBG_OK;
  });
  if (rc != NBG_OK || tl_batched) return rc;
  unsigned any = 0;
  for (uint64_t i = 0; i < n; ++i) any |= reason[i];
  return (any & NBG_STS_NILREPLICAS) ? NBG_E_NILREPLICAS : NBG_OK;
}

int nbg_diff_svc(nbg_ctx* ctx, const nbg_svc_batch* in, uint64_t n, uint8_t* reason, uint64_t* bitmap) {
  if (!ctx || !in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap || !in->flags || !in->ports_from || !in->ports_to) return NBG_E_INVAL;
  if (!csr_ptrs_ok(in->labels_from, n) || !csr_ptrs_ok(in->labels_to, n) || !csr_ptrs_ok(in->ann_from, n) ||
      !csr_ptrs_ok(in->ann_to, n) || !csr_ptrs_ok(in->sel_from, n) || !csr_ptrs_ok(in->sel_to, n))
    return NBG_E_INVAL;

  return for_each_device(ctx, n, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kSvcChunk, ++c) {
      const uint64_t i1 = i0 + kSvcChunk < hi ? i0 + kSvcChunk : hi, m = i1 - i0;
      if (!csr_ok(in->labels_from, i0, i1) || !csr_ok(in->labels_to, i0, i1) || !csr_ok(in->ann_from, i0, i1) ||
          !csr_ok(in->ann_to, i0, i1) || !csr_ok(in->sel_from, i0, i1) || !csr_ok(in->sel_to, i0, i1))
        return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const size_t a_fl = lay.in(in->flags + i0, m);
      const size_t a_pf = lay.in(in->ports_from + i0 * NBG_SVC_SLOTS, m * NBG_SVC_SLOTS * 8);
      const size_t a_pt = lay.in(in->ports_to + i0 * NBG_SVC_SLOTS, m * NBG_SVC_SLOTS * 8);
      const CsrDev lf = csr_in(lay, in->labels_from, i0, i1), lt = csr_in(lay, in->labels_to, i0, i1);
      const CsrDev af = csr_in(lay, in->ann_from, i0, i1), at = csr_in(lay, in->ann_to, i0, i1);
      const CsrDev sf = csr_in(lay, in->sel_from, i0, i1), st = csr_in(lay, in->sel_to, i0, i1);
      const size_t o_reason = lay.out(reason + i0, m);
      const size_t o_bits = lay.out(bitmap + i0 / 64, (m + 63) / 64 * 8);
      if (int r = slot_upload(d, sl, lay)) return r;
      nbg_svc_batch b{};
      b.flags = (const uint8_t*)(sl.dev + a_fl);
      b.ports_from = (const uint64_t*)(sl.dev + a_pf); b.ports_to = (const uint64_t*)(sl.dev + a_pt);
      b.labels_from = csr_ptr(sl, lf); b.labels_to = csr_ptr(sl, lt);
      b.ann_from = csr_ptr(sl, af); b.ann_to = csr_ptr(sl, at);
      b.sel_from = csr_ptr(sl, sf); b.sel_to = csr_ptr(sl, st);
      const unsigned bases[6] = {lf.base, lt.base, af.base, at.base, sf.base, st.base};
      CU(nbg_launch_diff_svc(&b, m, bases, (uint8_t*)(sl.dev + o_reason), (uint64_t*)(sl.dev + o_bits), d.s_run));
      ctx->launches++;
      if (int r = slot_download(d, sl, lay)) return r;
    }
    return NBG_OK;
  });
}

int nbg_diff_vs(nbg_ctx* ctx, const nbg_vs_batch* in, uint64_t n, uint8_t* reason, uint64_t* bitmap) {
  if (!ctx || !in) return NBG_E_INVAL;
  if (n == 0) return NBG_OK;
  if (!reason || !bitmap || !in->flags || !in->spec_from || !in->spec_to) return NBG_E_INVAL;
  if (!csr_ptrs_ok(in->hdr_from, n) || !csr_ptrs_ok(in->hdr_to, n)) return NBG_E_INVAL;

  return for_each_device(ctx, n, [&](Dev& d, uint64_t lo, uint64_t hi) -> int {
    uint64_t c = 0;
    for (uint64_t i0 = lo; i0 < hi; i0 += kVsChunk, ++c) {
      const uint64_t i1 = i0 + kVsChunk < hi ? i0 + kVsChunk : hi, m = i1 - i0;
      if (!csr_ok(in->hdr_from, i0, i1) || !csr_ok(in->hdr_to, i0, i1)) return NBG_E_RANGE;
      Slot& sl = d.slot[c & 1];
      Layout lay;
      const size_t a_fl = lay.in(in->flags + i0, m);
      const size_t a_sf = lay.in(in->spec_from + i0 * NBG_VS_SLOTS, m * NBG_VS_SLOTS * 8);
      const size_t a_st = lay.in(in->spec_to + i0 * NBG_VS_SLOTS, m * NBG_VS_SLOTS * 8);
      const CsrDev hf = csr_in(lay, in->hdr_from, i0, i1), ht = csr_in(lay, in->hdr_to, i0, i1);
      const size_t o_reason = lay.out(reason + i0, m);
      const size_t o_bits = lay.out(bitmap + i0 / 64, (m + 63) / 64 * 8);
      if (int r = slot_upload(d, sl, lay)) return r;
      nbg_vs_batch b{};
      b.flags = (const uint8_t*)(sl.dev + a_fl);
      b.spec_from = (const uint64_t*)(sl.dev + a_sf); b.spec_to = (const uint64_t*)(sl.dev + a_st);
      b.hdr_from = csr_ptr(sl, hf); b.hdr_to = csr_ptr(sl, ht);
      const unsigned bases[2] = {hf.base, ht.base};
      CU(nbg_launch_diff_vs(&b, m, bases, (uint8_t*)(sl.dev + o_reason), (uint64_t*)(sl.dev + o_bits), d.s_run));
      ctx->launches++;
      if (int r = slot_download(d, sl, lay)) return r;
    }
    return NBG_OK;
  });
}

}  // extern "C"
