// store.cu — a resident fleet: the batches of a reconcile step kept in device memory and
// edited in place (include/nbgraft.h, "resident fleet").
//
// Built on the public C ABI only: a step over the store is nbg_step_device on batch structs whose
// variable-length parts are SPANS (off,end) into the store's heaps.
//
//   fixed arrays   one row per slot (replicas, masks, tiles, annotation instants, ...)
//   heaps          one per variable-length array: payload buffer + per-slot start/end, all on the device; the
//                  host keeps three numbers per heap (capacity, used, live).  What a put replaces is added up
//                  by the put's own kernel and read back with it, so no per-slot state lives on the host
//   put            the m rows are gathered into a PINNED staging slot laid out like its
//                  device twin: descriptor table, slot list, every fixed array's rows, every heap's new
//                  spans and payload.  ONE asynchronous H2D copy of the slot and ONE kernel over the
//                  descriptor table then place everything: fixed rows and spans are scattered to their
//                  slots, each heap's payload (contiguous in the caller's CSR batch) lands at the heap's
//                  tail.  Large puts run as a sequence of such chunks over two slots, so the gather of
//                  one chunk overlaps the copy of the one before
//   compaction     when a heap cannot take the next put: the spans are read back, live runs are copied, slot by
//                  slot, into a fresh buffer (twice the live size if more than half of the old one was live)
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "../../../include/nbgraft.h"

namespace {

// ---------------------------------------------------------------- kernels
// One entry of a put's descriptor table (lives at the start of the staging slot).
struct PutDesc {
  unsigned long long dst;   // device address of the array (kinds 0, 1), of the heap's tail (2) or of the heap's start[] (3)
  unsigned long long dst2;  // kind 3: device address of the heap's end[]
  unsigned src;             // byte offset of the source inside the staging slot (kind 3: the new starts)
  unsigned words;           // kind 0: 4-byte words per row; kind 2: 4-byte words in all; kind 3: byte offset of the new ends
  unsigned kind;            // 0 row k -> row idx[k], 1 the same for 1-byte rows, 2 contiguous copy, 3 a heap's spans
  unsigned heap;            // kind 3: which heap (index into the replaced-length counters)
};
static_assert(sizeof(PutDesc) == 32, "descriptor layout");
constexpr unsigned kMaxDesc = 96;

// The whole put in one launch: blockIdx.y picks the descriptor, the blocks of a row stride over its items.
// Kind 3 also adds up the lengths of the runs it replaces (replaced[heap]), which is how the store knows how much
// of a heap is garbage without keeping a host copy of every slot's length.
__global__ void __launch_bounds__(256)
put_kernel(const char* __restrict__ stage, unsigned long long m, unsigned idx_off, unsigned long long* __restrict__ replaced) {
  const PutDesc d = reinterpret_cast<const PutDesc*>(stage)[blockIdx.y];
  const unsigned long long* idx = reinterpret_cast<const unsigned long long*>(stage + idx_off);
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (d.kind == 0u) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(stage + d.src);
    uint32_t* dst = reinterpret_cast<uint32_t*>(d.dst);
    for (const unsigned long long total = m * d.words; t < total; t += stride) {
      const unsigned long long k = t / d.words;
      dst[idx[k] * d.words + (t - k * d.words)] = src[t];
    }
  } else if (d.kind == 1u) {
    const uint8_t* src = reinterpret_cast<const uint8_t*>(stage + d.src);
    uint8_t* dst = reinterpret_cast<uint8_t*>(d.dst);
    for (; t < m; t += stride) dst[idx[t]] = src[t];
  } else if (d.kind == 2u) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(stage + d.src);
    uint32_t* dst = reinterpret_cast<uint32_t*>(d.dst);
    for (; t < d.words; t += stride) dst[t] = src[t];
  } else {
    const uint32_t* ns = reinterpret_cast<const uint32_t*>(stage + d.src);
    const uint32_t* ne = reinterpret_cast<const uint32_t*>(stage + d.words);
    uint32_t* start = reinterpret_cast<uint32_t*>(d.dst);
    uint32_t* end = reinterpret_cast<uint32_t*>(d.dst2);
    unsigned long long gone = 0;
    for (; t < m; t += stride) {                          // a slot is named once per put: no two threads share one
      const unsigned long long j = idx[t];
      gone += end[j] - start[j];
      start[j] = ns[t];
      end[j] = ne[t];
    }
    for (unsigned o = 16; o; o >>= 1) gone += __shfl_xor_sync(0xFFFFFFFFu, gone, o);
    if ((threadIdx.x & 31u) == 0 && gone) atomicAdd(replaced + d.heap, gone);
  }
}
// Compaction: slot j's run moves from old[start[j] .. end[j]) to fresh[new_start[j] ..); a warp per
// slot, `words` 4-byte words per element.  start/end are rewritten in place.
__global__ void __launch_bounds__(256)
compact_runs_kernel(uint32_t* __restrict__ fresh, const uint32_t* __restrict__ old, uint32_t* __restrict__ start,
                    uint32_t* __restrict__ end, const uint32_t* __restrict__ new_start, unsigned long long slots,
                    unsigned words, bool rewrite) {
  const unsigned long long j = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned lane = threadIdx.x & 31u;
  if (j >= slots) return;
  const unsigned s = start[j], e = end[j], ns = new_start[j];
  const unsigned long long n = (unsigned long long)(e - s) * words;
  const uint32_t* from = old + (unsigned long long)s * words;
  uint32_t* to = fresh + (unsigned long long)ns * words;
  for (unsigned long long w = lane; w < n; w += 32u) to[w] = from[w];
  __syncwarp();
  if (rewrite && lane == 0) { start[j] = ns; end[j] = ns + (e - s); }
}

#define CUS(call) do { if ((call) != cudaSuccess) { cudaGetLastError(); return NBG_E_CUDA; } } while (0)

// ---------------------------------------------------------------- data
struct Fixed {                 // one fixed-width per-slot array
  size_t width = 0;            // bytes per slot
  char* data = nullptr;
};

struct Heap {                  // one variable-length array (or two that share an index space)
  int payloads = 1;
  size_t elem[2] = {0, 0};     // bytes per element of each payload
  unsigned align = 1;          // every run starts at a multiple of `align` elements
  char* data[2] = {nullptr, nullptr};
  uint64_t cap = 0, used = 0, live = 0;      // elements
  uint32_t *start = nullptr, *end = nullptr; // device, one per slot
};

enum FixedId {
  F_REP_F, F_REP_T, F_TMPL_F, F_TMPL_T, F_MASK_F, F_MASK_T,          // StatefulSet
  F_SVC_NAME, F_SVC_PORT, F_SVC_FLAGS, F_SVC_PORTS,                  // Service
  F_VS_PREFIX, F_VS_REWRITE, F_VS_HOST, F_VS_FLAGS, F_VS_SPEC,       // VirtualService
  F_ANN_SEC, F_ANN_NSF, F_CHK_SEC, F_CHK_NSF, F_NB_FLAGS,            // culler
  F_COUNT
};
const size_t kFixedWidth[F_COUNT] = {4, 4, 8, 8, 24, 24, 8, 4, 1, NBG_SVC_SLOTS * 8, 8, 8, 8, 1, NBG_VS_SLOTS * 8, 8, 4, 8, 4, 1};

enum HeapId {
  H_POD_F, H_POD_T, H_LAB_F, H_LAB_T, H_ANN_F, H_ANN_T,              // StatefulSet
  H_SVC_LAB, H_SVC_ANN, H_SVC_SEL,                                   // Service
  H_VS_HDR_F, H_VS_HDR_T,                                            // VirtualService
  H_KERNELS, H_TERMINALS,                                            // culler (sec + nsf)
  H_COUNT
};

}  // namespace

struct nbg_store {
  nbg_ctx* ctx = nullptr;
  int device = 0, cuda_id = 0;
  uint64_t capacity = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t step_done = nullptr;   // recorded on the caller's stream after every step: puts are ordered behind it
  bool stepped = false;
  Fixed fixed[F_COUNT];
  Heap heap[H_COUNT];
  struct Slot {                // staging for put: a pinned host buffer and its device twin, same layout
    char* host = nullptr;
    char* dev = nullptr;
    size_t cap = 0;
    cudaEvent_t done = nullptr;   // recorded behind the kernel that read the device twin
    bool used = false;
  } slot[2];
  unsigned next_slot = 0;
  unsigned long long* replaced = nullptr;   // device, H_COUNT counters: lengths of the runs the current put replaces
  std::vector<uint32_t> stamp;    // duplicate check of a put's slot list: stamp[j] == epoch <=> j was named already
  uint32_t epoch = 0;
  char* outs = nullptr;        // grow-only device outputs of nbg_store_step
  size_t outs_cap = 0;
  uint64_t compactions = 0, puts = 0, rows_put = 0;
  std::mutex mu;
  std::mutex step_mu;          // nbg_store_step: one user of `outs` at a time
};

namespace {

constexpr size_t kStageAlign = 256;
size_t stage_round(size_t bytes) { return (bytes + kStageAlign - 1) / kStageAlign * kStageAlign; }

uint64_t round_up(uint64_t x, unsigned a) { return (x + a - 1) / a * a; }

// Make room for `need` more elements at the tail of a heap: compact (and grow) when it is full.
int heap_reserve(nbg_store* st, Heap& h, uint64_t need) {
  if (h.used + need <= h.cap) return NBG_OK;
  // plan the compacted layout on the host from the spans as they are on the device once everything queued so far
  // has run: live runs in slot order, each aligned.  (Compaction is rare and already costs an allocation and a copy
  // of the whole heap; reading the spans back here is what lets a put do without a host copy of every run length.)
  std::vector<uint32_t> new_start(st->capacity), cur_start(st->capacity), cur_end(st->capacity);
  CUS(cudaStreamSynchronize(st->stream));
  CUS(cudaMemcpy(cur_start.data(), h.start, st->capacity * 4, cudaMemcpyDeviceToHost));
  CUS(cudaMemcpy(cur_end.data(), h.end, st->capacity * 4, cudaMemcpyDeviceToHost));
  uint64_t at = 0;
  for (uint64_t j = 0; j < st->capacity; ++j) {
    new_start[j] = (uint32_t)at;
    at = round_up(at + (cur_end[j] - cur_start[j]), h.align);
  }
  uint64_t cap = h.cap;
  if (at + need > cap / 2) cap = 2 * (at + need) + 1024;      // mostly live: grow
  cap = round_up(cap, 4);
  if (cap > 0xFFFFFFF0ull) return NBG_E_TOOBIG;               // spans are 32-bit
  char* fresh[2] = {nullptr, nullptr};
  for (int p = 0; p < h.payloads; ++p)
    if (cudaMalloc(&fresh[p], cap * h.elem[p]) != cudaSuccess) {
      cudaGetLastError();
      if (p) cudaFree(fresh[0]);
      return NBG_E_NOMEM;
    }
  if (at) {
    uint32_t* ns_dev = nullptr;
    if (cudaMalloc(&ns_dev, st->capacity * sizeof(uint32_t)) != cudaSuccess) {
      cudaGetLastError();
      for (int p = 0; p < h.payloads; ++p) cudaFree(fresh[p]);
      return NBG_E_NOMEM;
    }
    cudaError_t e = cudaMemcpyAsync(ns_dev, new_start.data(), st->capacity * sizeof(uint32_t), cudaMemcpyHostToDevice, st->stream);
    const unsigned grid = (unsigned)((st->capacity * 32 + 255) / 256);
    for (int p = 0; p < h.payloads && e == cudaSuccess; ++p) {
      compact_runs_kernel<<<grid, 256, 0, st->stream>>>(
          reinterpret_cast<uint32_t*>(fresh[p]), reinterpret_cast<const uint32_t*>(h.data[p]), h.start, h.end, ns_dev,
          st->capacity, (unsigned)(h.elem[p] / 4), p == h.payloads - 1);     // spans rewritten after the last payload
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(st->stream);            // new_start (host) and ns_dev die here
    cudaFree(ns_dev);
    if (e != cudaSuccess) {
      cudaGetLastError();
      for (int p = 0; p < h.payloads; ++p) cudaFree(fresh[p]);
      return NBG_E_CUDA;
    }
  } else {
    CUS(cudaStreamSynchronize(st->stream));
  }
  if (st->stepped) CUS(cudaEventSynchronize(st->step_done));   // a step still reading the old buffers must finish first
  for (int p = 0; p < h.payloads; ++p) {
    if (h.data[p]) cudaFree(h.data[p]);
    h.data[p] = fresh[p];
  }
  h.cap = cap;
  h.used = at;
  if (at) st->compactions++;
  return NBG_OK;
}

// A chunk of a put while it is being laid out in a staging slot.
struct Gather { char* dst; const char* src; size_t bytes; };
struct Chunk {
  nbg_store* st;
  nbg_store::Slot* slot;
  const uint64_t* idx;          // the chunk's slot list (host)
  uint64_t k0, m;               // rows [k0, k0+m) of the caller's batches
  size_t at = 0;                // bump pointer inside the slot
  std::vector<PutDesc> desc;
  std::vector<Gather> gather;
  size_t take(size_t bytes) { const size_t o = at; at += stage_round(bytes); return o; }
  void add(unsigned kind, void* dst, size_t off, unsigned words) {
    PutDesc d{};
    d.dst = (unsigned long long)dst; d.src = (unsigned)off; d.words = words; d.kind = kind;
    desc.push_back(d);
  }
  // rows [k0, k0+m) of a fixed-width host array -> rows idx[] of a store array
  int fixed(int f, const void* src) {
    if (!src) return NBG_E_INVAL;
    const size_t width = kFixedWidth[f], off = take(m * width);
    gather.push_back({slot->host + off, static_cast<const char*>(src) + k0 * width, m * width});
    add(width == 1 ? 1u : 0u, st->fixed[f].data, off, (unsigned)(width / 4));
    return NBG_OK;
  }
  // the chunk's runs of one heap: payload to the heap's tail, new spans to the slots
  int heap(Heap& h, const uint32_t* off_all, const void* payload0, const void* payload1) {
    const uint32_t* off = off_all + k0;
    const uint64_t total = off[m] - off[0];
    if (total && (!payload0 || (h.payloads > 1 && !payload1))) return NBG_E_INVAL;
    const uint64_t need = round_up(total, h.align);
    if (int rc = heap_reserve(st, h, need)) return rc;
    const void* src[2] = {payload0, payload1};
    for (int p = 0; p < h.payloads && total; ++p) {
      const size_t o = take(total * h.elem[p]);
      gather.push_back({slot->host + o, static_cast<const char*>(src[p]) + (size_t)off[0] * h.elem[p], total * h.elem[p]});
      add(2u, h.data[p] + h.used * h.elem[p], o, (unsigned)(total * h.elem[p] / 4));
    }
    const size_t o_start = take(m * 4), o_end = take(m * 4);
    uint32_t* starts = reinterpret_cast<uint32_t*>(slot->host + o_start);
    uint32_t* ends = reinterpret_cast<uint32_t*>(slot->host + o_end);
    const uint32_t base = (uint32_t)h.used - off[0];
    for (uint64_t k = 0; k < m; ++k) { starts[k] = base + off[k]; ends[k] = base + off[k + 1]; }
    h.used += need;
    h.live += total;                                     // what the put replaces is subtracted when it has run
    PutDesc d{};
    d.dst = (unsigned long long)h.start; d.dst2 = (unsigned long long)h.end;
    d.src = (unsigned)o_start; d.words = (unsigned)o_end; d.kind = 3u; d.heap = (unsigned)(&h - st->heap);
    desc.push_back(d);
    return NBG_OK;
  }
};

// Bytes a chunk of rows [k0, k1) needs in a staging slot (upper bound: every piece rounded up).
size_t chunk_bytes(uint64_t k0, uint64_t k1, const nbg_sts_sparse_batch* sts, const nbg_svc_gen_batch* svc,
                   const nbg_vs_gen_batch* vs, const nbg_cull_batch* cull) {
  const uint64_t m = k1 - k0;
  size_t fixed = 0;
  for (int f = 0; f < F_COUNT; ++f) fixed += m * kFixedWidth[f] + kStageAlign;
  size_t bytes = stage_round(kMaxDesc * sizeof(PutDesc)) + stage_round(m * 8) + fixed + (size_t)H_COUNT * 2 * (m * 4 + kStageAlign);
  auto run = [&](const uint32_t* off, size_t elem) { if (off) bytes += (size_t)(off[k1] - off[k0]) * elem + 2 * kStageAlign; };
  if (sts) {
    run(sts->pod_from.off, 8); run(sts->pod_to.off, 8);
    run(sts->labels_from.off, 16); run(sts->labels_to.off, 16); run(sts->ann_from.off, 16); run(sts->ann_to.off, 16);
  }
  if (svc) { run(svc->labels_to.off, 16); run(svc->ann_to.off, 16); run(svc->sel_to.off, 16); }
  if (vs) { run(vs->hdr_from.off, 16); run(vs->hdr_to.off, 16); }
  if (cull) { run(cull->k_off, 12); run(cull->t_off, 12); }
  return bytes;
}

// Copy the listed pieces into the pinned slot.  Small chunks: one thread — measured on the B200 host, three more
// gather threads cut the gather of a 14 MB put from 1.1 to 0.8 ms but slowed the planning loop and the DMA running
// beside it by more (2.0 ms per put against 1.45).  Chunks of 16 MB and more (a put of tens of thousands of rows)
// are split by bytes over up to four threads started for the chunk: there the copy is the whole cost.
void gather_part(const std::vector<Gather>& g, size_t total, unsigned who, unsigned parts) {
  const size_t lo = total * who / parts, hi = total * (who + 1) / parts;
  size_t at = 0;
  for (const auto& t : g) {
    const size_t b = std::max(lo, at), e = std::min(hi, at + t.bytes);
    if (b < e) memcpy(t.dst + (b - at), t.src + (b - at), e - b);
    at += t.bytes;
    if (at >= hi) break;
  }
}
void run_gather(const std::vector<Gather>& g) {
  size_t total = 0;
  for (const auto& t : g) total += t.bytes;
  const unsigned hc = std::thread::hardware_concurrency();
  const unsigned parts = total >= (16u << 20) ? std::min(4u, std::max(1u, hc / 2)) : 1u;
  if (parts == 1) { gather_part(g, total, 0, 1); return; }
  std::vector<std::thread> pool;
  for (unsigned w = 1; w < parts; ++w) pool.emplace_back(gather_part, std::cref(g), total, w, parts);
  gather_part(g, total, 0, parts);
  for (auto& th : pool) th.join();
}

nbg_kvmap span_map(const Heap& h) {
  nbg_kvmap m{};
  m.off = h.start;
  m.end = h.end;
  m.ent = reinterpret_cast<const nbg_kv*>(h.data[0]);
  return m;
}
nbg_podsparse span_sparse(const Fixed& mask, const Heap& h) {
  nbg_podsparse p{};
  p.mask = reinterpret_cast<const uint64_t*>(mask.data);
  p.off = h.start;
  p.end = h.end;
  p.val = reinterpret_cast<const uint64_t*>(h.data[0]);
  return p;
}

}  // namespace

extern "C" {

void nbg_store_destroy(nbg_store* st) {
  if (!st) return;
  cudaSetDevice(st->cuda_id);
  if (st->stream) cudaStreamSynchronize(st->stream);
  for (auto& f : st->fixed) if (f.data) cudaFree(f.data);
  for (auto& h : st->heap) {
    for (int p = 0; p < h.payloads; ++p) if (h.data[p]) cudaFree(h.data[p]);
    if (h.start) cudaFree(h.start);
    if (h.end) cudaFree(h.end);
  }
  for (auto& sl : st->slot) {
    if (sl.done) { cudaEventSynchronize(sl.done); cudaEventDestroy(sl.done); }
    if (sl.host) cudaFreeHost(sl.host);
    if (sl.dev) cudaFree(sl.dev);
  }
  if (st->outs) cudaFree(st->outs);
  if (st->replaced) cudaFree(st->replaced);
  if (st->step_done) { cudaEventSynchronize(st->step_done); cudaEventDestroy(st->step_done); }
  if (st->stream) cudaStreamDestroy(st->stream);
  delete st;
}

int nbg_store_create(nbg_ctx* ctx, int device, uint64_t capacity, nbg_store** out) {
  if (!ctx || !out || capacity == 0) return NBG_E_INVAL;
  *out = nullptr;
  const int id = nbg_device_id(ctx, device);
  if (id < 0) return NBG_E_INVAL;
  if (capacity > 0x7FFFFFFFull) return NBG_E_TOOBIG;
  nbg_store* st = new (std::nothrow) nbg_store();
  if (!st) return NBG_E_NOMEM;
  st->ctx = ctx; st->device = device; st->cuda_id = id; st->capacity = capacity;
  bool ok = cudaSetDevice(id) == cudaSuccess && cudaStreamCreateWithFlags(&st->stream, cudaStreamNonBlocking) == cudaSuccess &&
            cudaEventCreateWithFlags(&st->step_done, cudaEventDisableTiming) == cudaSuccess;
  for (auto& sl : st->slot) ok = ok && cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming) == cudaSuccess;
  ok = ok && cudaMalloc(&st->replaced, H_COUNT * sizeof(unsigned long long)) == cudaSuccess;
  try { st->stamp.assign(capacity, 0u); } catch (...) { ok = false; }
  for (int f = 0; ok && f < F_COUNT; ++f) {
    st->fixed[f].width = kFixedWidth[f];
    ok = cudaMalloc(&st->fixed[f].data, capacity * kFixedWidth[f]) == cudaSuccess &&
         cudaMemsetAsync(st->fixed[f].data, 0, capacity * kFixedWidth[f], st->stream) == cudaSuccess;
  }
  for (int k = 0; ok && k < H_COUNT; ++k) {
    Heap& h = st->heap[k];
    const bool records = k == H_KERNELS || k == H_TERMINALS, pod = k == H_POD_F || k == H_POD_T;
    h.payloads = records ? 2 : 1;
    h.elem[0] = (records || pod) ? 8 : sizeof(nbg_kv);
    h.elem[1] = records ? 4 : 0;
    h.align = pod ? 2 : 1;                               // sparse runs stay 16-byte aligned
    ok = cudaMalloc(&h.start, capacity * 4) == cudaSuccess && cudaMalloc(&h.end, capacity * 4) == cudaSuccess &&
         cudaMemsetAsync(h.start, 0, capacity * 4, st->stream) == cudaSuccess &&
         cudaMemsetAsync(h.end, 0, capacity * 4, st->stream) == cudaSuccess;
  }
  ok = ok && cudaStreamSynchronize(st->stream) == cudaSuccess;
  if (!ok) { cudaGetLastError(); nbg_store_destroy(st); return NBG_E_NOMEM; }
  *out = st;
  return NBG_OK;
}

int nbg_store_put(nbg_store* st, const uint64_t* idx, uint64_t m, const nbg_sts_sparse_batch* sts,
                  const nbg_svc_gen_batch* svc, const nbg_vs_gen_batch* vs, const nbg_cull_batch* cull) {
  if (!st || (m && !idx)) return NBG_E_INVAL;
  if (m == 0 || !(sts || svc || vs || cull)) return NBG_OK;
  for (uint64_t k = 0; k < m; ++k) if (idx[k] >= st->capacity) return NBG_E_RANGE;
  // spans are not accepted here: the rows arrive as plain CSR batches
  if (sts && (sts->pod_from.end || sts->pod_to.end || sts->labels_from.end || sts->labels_to.end || sts->ann_from.end || sts->ann_to.end))
    return NBG_E_INVAL;
  if (svc && (svc->labels_to.end || svc->ann_to.end || svc->sel_to.end)) return NBG_E_INVAL;
  if (vs && (vs->hdr_from.end || vs->hdr_to.end)) return NBG_E_INVAL;
  if (cull && (cull->k_end || cull->t_end)) return NBG_E_INVAL;
  // every offset array present and monotone, sparse runs aligned and as long as their masks say
  {
    const uint32_t* offs[H_COUNT] = {
        sts ? sts->pod_from.off : nullptr, sts ? sts->pod_to.off : nullptr, sts ? sts->labels_from.off : nullptr,
        sts ? sts->labels_to.off : nullptr, sts ? sts->ann_from.off : nullptr, sts ? sts->ann_to.off : nullptr,
        svc ? svc->labels_to.off : nullptr, svc ? svc->ann_to.off : nullptr, svc ? svc->sel_to.off : nullptr,
        vs ? vs->hdr_from.off : nullptr, vs ? vs->hdr_to.off : nullptr, cull ? cull->k_off : nullptr, cull ? cull->t_off : nullptr};
    const bool wanted[H_COUNT] = {!!sts, !!sts, !!sts, !!sts, !!sts, !!sts, !!svc, !!svc, !!svc, !!vs, !!vs, !!cull, !!cull};
    for (int h = 0; h < H_COUNT; ++h) {
      if (!wanted[h]) continue;
      if (!offs[h]) return NBG_E_INVAL;
      const unsigned align = st->heap[h].align;
      unsigned bad = 0;
      for (uint64_t k = 0; k < m; ++k) bad |= (unsigned)(offs[h][k] > offs[h][k + 1]) | (offs[h][k] % align != 0);
      if (bad) return NBG_E_RANGE;
    }
  }
  if (sts) {
    const nbg_podsparse* sides[2] = {&sts->pod_from, &sts->pod_to};
    for (const nbg_podsparse* p : sides) {
      if (!p->mask) return NBG_E_INVAL;
      unsigned bad = 0;
      for (uint64_t k = 0; k < m; ++k) {
        const unsigned bits = (unsigned)(__builtin_popcountll(p->mask[3 * k]) + __builtin_popcountll(p->mask[3 * k + 1]) +
                                         __builtin_popcountll(p->mask[3 * k + 2]));
        bad |= (unsigned)(p->off[k + 1] - p->off[k] != ((bits + 1u) & ~1u));
      }
      if (bad) return NBG_E_RANGE;
    }
  }
  std::unique_lock<std::mutex> lock(st->mu);
  // a slot named twice would be scattered to by two rows at once (either may win on the device while the host
  // mirror keeps the last), so it is refused before anything is touched
  if (++st->epoch == 0) { std::fill(st->stamp.begin(), st->stamp.end(), 0u); st->epoch = 1; }
  for (uint64_t k = 0; k < m; ++k) {
    if (st->stamp[idx[k]] == st->epoch) return NBG_E_INVAL;
    st->stamp[idx[k]] = st->epoch;
  }
  CUS(cudaSetDevice(st->cuda_id));
  if (st->stepped) CUS(cudaStreamWaitEvent(st->stream, st->step_done, 0));   // rows are rewritten only after the last step has read them
  CUS(cudaMemsetAsync(st->replaced, 0, H_COUNT * sizeof(unsigned long long), st->stream));
  // chunk size: a quarter of the put (so that gathering one chunk overlaps the copy of the one before), within 2..48 MB
  const size_t kChunkBytes = std::min<size_t>(48u << 20, std::max<size_t>(2u << 20, chunk_bytes(0, m, sts, svc, vs, cull) / 4));
  int rc = NBG_OK;
  {
    // room for the WHOLE put in every heap before the first chunk is queued: a large put then costs each heap at
    // most one compaction, however many chunks carry it (each chunk pads its runs' total up to the heap's alignment)
    const uint32_t* offs[H_COUNT] = {
        sts ? sts->pod_from.off : nullptr, sts ? sts->pod_to.off : nullptr, sts ? sts->labels_from.off : nullptr,
        sts ? sts->labels_to.off : nullptr, sts ? sts->ann_from.off : nullptr, sts ? sts->ann_to.off : nullptr,
        svc ? svc->labels_to.off : nullptr, svc ? svc->ann_to.off : nullptr, svc ? svc->sel_to.off : nullptr,
        vs ? vs->hdr_from.off : nullptr, vs ? vs->hdr_to.off : nullptr, cull ? cull->k_off : nullptr, cull ? cull->t_off : nullptr};
    const uint64_t chunks = m / 256 + 2;                // generous bound on the number of chunks (a short reserve only costs a compaction)
    for (int h = 0; h < H_COUNT && rc == NBG_OK; ++h)
      if (offs[h]) rc = heap_reserve(st, st->heap[h], (uint64_t)(offs[h][m] - offs[h][0]) + chunks * st->heap[h].align);
    if (rc != NBG_OK) return rc;
  }
#ifdef NBG_STORE_TIMING
  auto tick = [] { return std::chrono::steady_clock::now(); };
  auto us = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
    return (double)std::chrono::duration_cast<std::chrono::nanoseconds>(b - a).count() * 1e-3; };
  const auto t_begin = tick();
  double t_layout = 0, t_gather = 0, t_queue = 0;
#endif
  for (uint64_t k0 = 0; k0 < m && rc == NBG_OK;) {
    uint64_t rows = std::min<uint64_t>(m - k0, 1u << 16);
    while (rows > 1 && chunk_bytes(k0, k0 + rows, sts, svc, vs, cull) > kChunkBytes) rows = (rows + 1) / 2;
    const size_t need = chunk_bytes(k0, k0 + rows, sts, svc, vs, cull);
    nbg_store::Slot& sl = st->slot[st->next_slot];
    st->next_slot ^= 1u;
    if (sl.used) CUS(cudaEventSynchronize(sl.done));      // the copy and the kernel that read this slot are done
    if (need > sl.cap) {
      if (sl.host) cudaFreeHost(sl.host);
      if (sl.dev) cudaFree(sl.dev);
      sl.host = sl.dev = nullptr; sl.cap = 0;
      const size_t cap = need + need / 4;
      if (cudaHostAlloc(&sl.host, cap, cudaHostAllocDefault) != cudaSuccess || cudaMalloc(&sl.dev, cap) != cudaSuccess) {
        cudaGetLastError();
        if (sl.host) { cudaFreeHost(sl.host); sl.host = nullptr; }
        return NBG_E_NOMEM;
      }
      sl.cap = cap;
    }
#ifdef NBG_STORE_TIMING
    const auto t0 = tick();
#endif
    Chunk c;
    c.st = st; c.slot = &sl; c.idx = idx + k0; c.k0 = k0; c.m = rows;
    c.take(kMaxDesc * sizeof(PutDesc));                   // the descriptor table sits at offset 0
    const size_t idx_off = c.take(rows * 8);
    c.gather.push_back({sl.host + idx_off, reinterpret_cast<const char*>(idx + k0), rows * 8});
    auto fixed = [&](int f, const void* src) { if (rc == NBG_OK) rc = c.fixed(f, src); };
    auto heap = [&](int k, const uint32_t* off, const void* p0, const void* p1 = nullptr) {
      if (rc == NBG_OK) rc = c.heap(st->heap[k], off, p0, p1);
    };
    if (sts) {
      fixed(F_REP_F, sts->replicas_from); fixed(F_REP_T, sts->replicas_to);
      fixed(F_TMPL_F, sts->tmpl_from); fixed(F_TMPL_T, sts->tmpl_to);
      fixed(F_MASK_F, sts->pod_from.mask); fixed(F_MASK_T, sts->pod_to.mask);
      heap(H_POD_F, sts->pod_from.off, sts->pod_from.val); heap(H_POD_T, sts->pod_to.off, sts->pod_to.val);
      heap(H_LAB_F, sts->labels_from.off, sts->labels_from.ent); heap(H_LAB_T, sts->labels_to.off, sts->labels_to.ent);
      heap(H_ANN_F, sts->ann_from.off, sts->ann_from.ent); heap(H_ANN_T, sts->ann_to.off, sts->ann_to.ent);
    }
    if (svc) {
      fixed(F_SVC_NAME, svc->name_hash); fixed(F_SVC_PORT, svc->target_port);
      fixed(F_SVC_FLAGS, svc->flags); fixed(F_SVC_PORTS, svc->ports_to);
      heap(H_SVC_LAB, svc->labels_to.off, svc->labels_to.ent); heap(H_SVC_ANN, svc->ann_to.off, svc->ann_to.ent);
      heap(H_SVC_SEL, svc->sel_to.off, svc->sel_to.ent);
    }
    if (vs) {
      fixed(F_VS_PREFIX, vs->prefix_hash); fixed(F_VS_REWRITE, vs->rewrite_hash); fixed(F_VS_HOST, vs->host_hash);
      fixed(F_VS_FLAGS, vs->flags); fixed(F_VS_SPEC, vs->spec_to);
      heap(H_VS_HDR_F, vs->hdr_from.off, vs->hdr_from.ent); heap(H_VS_HDR_T, vs->hdr_to.off, vs->hdr_to.ent);
    }
    if (cull) {
      fixed(F_ANN_SEC, cull->ann_sec); fixed(F_ANN_NSF, cull->ann_nsf);
      fixed(F_CHK_SEC, cull->chk_sec); fixed(F_CHK_NSF, cull->chk_nsf); fixed(F_NB_FLAGS, cull->nb_flags);
      heap(H_KERNELS, cull->k_off, cull->k_sec, cull->k_nsf); heap(H_TERMINALS, cull->t_off, cull->t_sec, cull->t_nsf);
    }
    if (rc != NBG_OK) break;
    if (c.desc.size() > kMaxDesc || c.at > sl.cap) { rc = NBG_E_TOOBIG; break; }   // cannot happen: sized above
#ifdef NBG_STORE_TIMING
    const auto t1 = tick();
#endif
    run_gather(c.gather);
    memcpy(sl.host, c.desc.data(), c.desc.size() * sizeof(PutDesc));
#ifdef NBG_STORE_TIMING
    const auto t2 = tick();
#endif
    // widest descriptor decides the row of blocks; the others stride or idle
    unsigned long long widest = rows;
    for (const auto& d : c.desc) widest = std::max<unsigned long long>(widest, d.kind == 0u ? rows * d.words : d.kind == 2u ? d.words : rows);
    const unsigned gx = (unsigned)std::min<unsigned long long>((widest + 256ull * 8 - 1) / (256ull * 8), 1024ull);
    if (cudaMemcpyAsync(sl.dev, sl.host, c.at, cudaMemcpyHostToDevice, st->stream) != cudaSuccess) { cudaGetLastError(); rc = NBG_E_CUDA; break; }
    put_kernel<<<dim3(gx ? gx : 1u, (unsigned)c.desc.size()), 256, 0, st->stream>>>(sl.dev, rows, (unsigned)idx_off, st->replaced);
    if (cudaGetLastError() != cudaSuccess || cudaEventRecord(sl.done, st->stream) != cudaSuccess) { cudaGetLastError(); rc = NBG_E_CUDA; break; }
    sl.used = true;
    k0 += rows;
#ifdef NBG_STORE_TIMING
    const auto t3 = tick();
    t_layout += us(t0, t1); t_gather += us(t1, t2); t_queue += us(t2, t3);
#endif
  }
  // the rows are on the device when the call returns (the caller's buffers were only read by the gather)
#ifdef NBG_STORE_TIMING
  const auto t_sync0 = tick();
#endif
  unsigned long long gone[H_COUNT] = {};
  if (cudaMemcpyAsync(gone, st->replaced, sizeof gone, cudaMemcpyDeviceToHost, st->stream) != cudaSuccess ||
      cudaStreamSynchronize(st->stream) != cudaSuccess) { cudaGetLastError(); if (rc == NBG_OK) rc = NBG_E_CUDA; }
  else for (int h = 0; h < H_COUNT; ++h) st->heap[h].live -= gone[h];
#ifdef NBG_STORE_TIMING
  fprintf(stderr, "put m=%llu: layout %.0f us, gather %.0f us, queue %.0f us, final sync %.0f us, total %.0f us\n",
          (unsigned long long)m, t_layout, t_gather, t_queue, us(t_sync0, tick()), us(t_begin, tick()));
#endif
  if (rc == NBG_OK) { st->puts++; st->rows_put += m; }
  return rc;
}

static void store_views(const nbg_store* st, nbg_sts_sparse_batch* sts, nbg_svc_gen_batch* svc, nbg_vs_gen_batch* vs,
                        nbg_cull_batch* cull) {
  const Fixed* f = st->fixed;
  const Heap* h = st->heap;
  *sts = nbg_sts_sparse_batch{};
  sts->replicas_from = reinterpret_cast<const int32_t*>(f[F_REP_F].data);
  sts->replicas_to = reinterpret_cast<const int32_t*>(f[F_REP_T].data);
  sts->tmpl_from = reinterpret_cast<const uint64_t*>(f[F_TMPL_F].data);
  sts->tmpl_to = reinterpret_cast<const uint64_t*>(f[F_TMPL_T].data);
  sts->pod_from = span_sparse(f[F_MASK_F], h[H_POD_F]);
  sts->pod_to = span_sparse(f[F_MASK_T], h[H_POD_T]);
  sts->labels_from = span_map(h[H_LAB_F]); sts->labels_to = span_map(h[H_LAB_T]);
  sts->ann_from = span_map(h[H_ANN_F]); sts->ann_to = span_map(h[H_ANN_T]);
  *svc = nbg_svc_gen_batch{};
  svc->name_hash = reinterpret_cast<const uint64_t*>(f[F_SVC_NAME].data);
  svc->target_port = reinterpret_cast<const int32_t*>(f[F_SVC_PORT].data);
  svc->flags = reinterpret_cast<const uint8_t*>(f[F_SVC_FLAGS].data);
  svc->ports_to = reinterpret_cast<const uint64_t*>(f[F_SVC_PORTS].data);
  svc->labels_to = span_map(h[H_SVC_LAB]); svc->ann_to = span_map(h[H_SVC_ANN]); svc->sel_to = span_map(h[H_SVC_SEL]);
  *vs = nbg_vs_gen_batch{};
  vs->prefix_hash = reinterpret_cast<const uint64_t*>(f[F_VS_PREFIX].data);
  vs->rewrite_hash = reinterpret_cast<const uint64_t*>(f[F_VS_REWRITE].data);
  vs->host_hash = reinterpret_cast<const uint64_t*>(f[F_VS_HOST].data);
  vs->flags = reinterpret_cast<const uint8_t*>(f[F_VS_FLAGS].data);
  vs->spec_to = reinterpret_cast<const uint64_t*>(f[F_VS_SPEC].data);
  vs->hdr_from = span_map(h[H_VS_HDR_F]); vs->hdr_to = span_map(h[H_VS_HDR_T]);
  *cull = nbg_cull_batch{};
  cull->k_off = h[H_KERNELS].start; cull->k_end = h[H_KERNELS].end;
  cull->k_sec = reinterpret_cast<const int64_t*>(h[H_KERNELS].data[0]);
  cull->k_nsf = reinterpret_cast<const uint32_t*>(h[H_KERNELS].data[1]);
  cull->t_off = h[H_TERMINALS].start; cull->t_end = h[H_TERMINALS].end;
  cull->t_sec = reinterpret_cast<const int64_t*>(h[H_TERMINALS].data[0]);
  cull->t_nsf = reinterpret_cast<const uint32_t*>(h[H_TERMINALS].data[1]);
  cull->ann_sec = reinterpret_cast<const int64_t*>(f[F_ANN_SEC].data);
  cull->ann_nsf = reinterpret_cast<const uint32_t*>(f[F_ANN_NSF].data);
  cull->chk_sec = reinterpret_cast<const int64_t*>(f[F_CHK_SEC].data);
  cull->chk_nsf = reinterpret_cast<const uint32_t*>(f[F_CHK_NSF].data);
  cull->nb_flags = reinterpret_cast<const uint8_t*>(f[F_NB_FLAGS].data);
}

int nbg_store_step_device(nbg_store* st, uint64_t n, const nbg_vs_gen_consts* vs_consts, const nbg_cull_params* params,
                          const nbg_step_outputs* out, void* stream) {
  if (!st || !out) return NBG_E_INVAL;
  if (n > st->capacity) return NBG_E_RANGE;
  if (n == 0) return NBG_OK;
  nbg_sts_sparse_batch sts; nbg_svc_gen_batch svc; nbg_vs_gen_batch vs; nbg_cull_batch cull;
  // a predicate takes part when its outputs are given
  const bool do_sts = out->sts_reason && out->sts_bitmap, do_svc = out->svc_reason && out->svc_bitmap,
             do_vs = out->vs_reason && out->vs_bitmap && vs_consts,
             do_cull = out->action && out->argmax && out->cull_bitmap && out->period_bitmap && params;
  std::unique_lock<std::mutex> lock(st->mu);            // held over the enqueue: a put cannot swap a heap under it
  CUS(cudaSetDevice(st->cuda_id));
  store_views(st, &sts, &svc, &vs, &cull);
  const int rc = nbg_step_device(st->ctx, st->device, do_sts ? &sts : nullptr, do_svc ? &svc : nullptr, do_vs ? &vs : nullptr,
                                 vs_consts, do_cull ? &cull : nullptr, params, n, out, stream, nullptr, nullptr);
  if (rc == NBG_OK) {
    CUS(cudaEventRecord(st->step_done, (cudaStream_t)stream));
    st->stepped = true;
  }
  return rc;
}

int nbg_store_step(nbg_store* st, uint64_t n, const nbg_vs_gen_consts* vs_consts, const nbg_cull_params* params,
                   const nbg_step_outputs* out) {
  if (!st || !out) return NBG_E_INVAL;
  if (n > st->capacity) return NBG_E_RANGE;
  if (n == 0) return NBG_OK;
  std::unique_lock<std::mutex> whole(st->step_mu);
  const size_t words = (n + 63) / 64, bm = words * 8, row = (n + 255) / 256 * 256, bmr = (bm + 255) / 256 * 256;
  const size_t need = 3 * (row + bmr) + row + (n * 4 + 255) / 256 * 256 + 2 * bmr;
  char* base;
  {
    std::unique_lock<std::mutex> lock(st->mu);
    CUS(cudaSetDevice(st->cuda_id));
    if (need > st->outs_cap) {
      CUS(cudaStreamSynchronize(st->stream));
      if (st->outs) cudaFree(st->outs);
      st->outs = nullptr; st->outs_cap = 0;
      if (cudaMalloc(&st->outs, need) != cudaSuccess) { cudaGetLastError(); return NBG_E_NOMEM; }
      st->outs_cap = need;
    }
    base = st->outs;
  }
  nbg_step_outputs d{};
  char* at = base;
  auto take = [&](size_t bytes) { char* p = at; at += (bytes + 255) / 256 * 256; return p; };
  if (out->sts_reason && out->sts_bitmap) { d.sts_reason = (uint8_t*)take(n); d.sts_bitmap = (uint64_t*)take(bm); }
  if (out->svc_reason && out->svc_bitmap) { d.svc_reason = (uint8_t*)take(n); d.svc_bitmap = (uint64_t*)take(bm); }
  if (out->vs_reason && out->vs_bitmap) { d.vs_reason = (uint8_t*)take(n); d.vs_bitmap = (uint64_t*)take(bm); }
  if (out->action && out->argmax && out->cull_bitmap && out->period_bitmap) {
    d.action = (uint8_t*)take(n); d.argmax = (uint32_t*)take(n * 4);
    d.cull_bitmap = (uint64_t*)take(bm); d.period_bitmap = (uint64_t*)take(bm);
  }
  if (int rc = nbg_store_step_device(st, n, vs_consts, params, &d, st->stream)) return rc;
  auto back = [&](void* host, const void* dev, size_t bytes) {
    return !dev || cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, st->stream) == cudaSuccess;
  };
  bool ok = back(out->sts_reason, d.sts_reason, n) && back(out->sts_bitmap, d.sts_bitmap, bm) &&
            back(out->svc_reason, d.svc_reason, n) && back(out->svc_bitmap, d.svc_bitmap, bm) &&
            back(out->vs_reason, d.vs_reason, n) && back(out->vs_bitmap, d.vs_bitmap, bm) &&
            back(out->action, d.action, n) && back(out->argmax, d.argmax, n * 4) &&
            back(out->cull_bitmap, d.cull_bitmap, bm) && back(out->period_bitmap, d.period_bitmap, bm);
  if (cudaStreamSynchronize(st->stream) != cudaSuccess) ok = false;
  if (!ok) { cudaGetLastError(); return NBG_E_CUDA; }
  return NBG_OK;
}

int nbg_store_check(nbg_store* st, uint64_t n, unsigned* what) {
  if (what) *what = 0;
  if (!st) return NBG_E_INVAL;
  if (n > st->capacity) return NBG_E_RANGE;
  if (n == 0) return NBG_OK;
  nbg_sts_sparse_batch sts; nbg_svc_gen_batch svc; nbg_vs_gen_batch vs; nbg_cull_batch cull;
  std::unique_lock<std::mutex> lock(st->mu);
  store_views(st, &sts, &svc, &vs, &cull);
  const Heap* h = st->heap;
  nbg_step_extents ext{};
  ext.sts_val_from = h[H_POD_F].used; ext.sts_val_to = h[H_POD_T].used;
  ext.sts_labels_from = h[H_LAB_F].used; ext.sts_labels_to = h[H_LAB_T].used;
  ext.sts_ann_from = h[H_ANN_F].used; ext.sts_ann_to = h[H_ANN_T].used;
  ext.svc_labels_to = h[H_SVC_LAB].used; ext.svc_ann_to = h[H_SVC_ANN].used; ext.svc_sel_to = h[H_SVC_SEL].used;
  ext.vs_hdr_from = h[H_VS_HDR_F].used; ext.vs_hdr_to = h[H_VS_HDR_T].used;
  ext.cull_kernels = h[H_KERNELS].used; ext.cull_terminals = h[H_TERMINALS].used;
  return nbg_check_step_device(st->ctx, st->device, &sts, &svc, &vs, &cull, n, &ext, st->stream, what);
}

int nbg_store_get_stats(const nbg_store* st, nbg_store_stats* out) {
  if (!st || !out) return NBG_E_INVAL;
  *out = nbg_store_stats{};
  out->capacity = st->capacity;
  for (const auto& f : st->fixed) out->fixed_bytes += st->capacity * f.width;
  for (const auto& h : st->heap) {
    const uint64_t per = h.elem[0] + h.elem[1];
    out->fixed_bytes += st->capacity * 8;                 // the two span words of every slot
    out->heap_bytes += h.cap * per;
    out->live_bytes += h.live * per;
    out->garbage_bytes += (h.used - h.live) * per;
  }
  out->compactions = st->compactions;
  out->puts = st->puts;
  out->rows_put = st->rows_put;
  return NBG_OK;
}

}  // extern "C"
