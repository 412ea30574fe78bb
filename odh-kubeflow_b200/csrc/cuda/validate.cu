// validate.cu — the argument checks of the host-buffer entry points, for batches that already live on the device.
//
// The nbg_*_device entry points only enqueue: they cannot look at device memory, so they trust the offsets and
// spans they are given (include/nbgraft.h).  A batch with a run that ends past its payload array, or a sparse run
// whose length disagrees with its mask, makes the kernels read outside the arrays.  nbg_check_step_device is the
// missing check as a kernel: one thread per object walks every (start, end) pair of its rows and raises a flag.
// It is meant for the places where the host-buffer calls check for free — after uploading a batch built by a new
// packer, in tests, in a debug build of the caller — not for every step.
#include <cuda_runtime.h>

#include "../../../include/nbgraft.h"

namespace {

struct Run {                 // one variable-length array: CSR (end == nullptr: off has n+1 entries) or spans
  const uint32_t* off;
  const uint32_t* end;
  unsigned long long extent; // elements in the payload array
  unsigned align;            // every run starts at a multiple of this
};
struct Sparse { Run run; const uint64_t* mask; };

constexpr int kMaxRuns = 16;
struct CheckArgs {
  Run runs[kMaxRuns];
  int n_runs;
  Sparse sparse[2];
  int n_sparse;
};

// bit 0: a run starts after it ends; bit 1: a run ends past its payload; bit 2: misaligned start;
// bit 3: a sparse run's length is not popcount(mask) rounded up to even
__global__ void __launch_bounds__(256)
check_kernel(CheckArgs a, unsigned long long n, unsigned* __restrict__ flags) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  unsigned bad = 0;
  auto look = [&](const Run& r, unsigned& len) {
    const unsigned s = r.off[i], e = r.end ? r.end[i] : r.off[i + 1];
    if (s > e) bad |= 1u;
    if ((unsigned long long)e > r.extent) bad |= 2u;
    if (s % r.align) bad |= 4u;
    len = e - s;
  };
  for (int k = 0; k < a.n_runs; ++k) { unsigned len; look(a.runs[k], len); }
  for (int k = 0; k < a.n_sparse; ++k) {
    unsigned len;
    look(a.sparse[k].run, len);
    const uint64_t* m = a.sparse[k].mask + i * 3;
    const unsigned bits = (unsigned)(__popcll(m[0]) + __popcll(m[1]) + __popcll(m[2]));
    if (len != ((bits + 1u) & ~1u)) bad |= 8u;
  }
  if (bad) atomicOr(flags, bad);
}

}  // namespace

extern "C" int nbg_device_id(const nbg_ctx* ctx, int device);

extern "C" int nbg_check_step_device(nbg_ctx* ctx, int device, const nbg_sts_sparse_batch* sts, const nbg_svc_gen_batch* svc,
                                     const nbg_vs_gen_batch* vs, const nbg_cull_batch* cull, uint64_t n,
                                     const nbg_step_extents* ext, void* stream, unsigned* what) {
  if (what) *what = 0;
  if (!ctx || !ext) return NBG_E_INVAL;
  const int id = nbg_device_id(ctx, device);
  if (id < 0) return NBG_E_INVAL;
  if (n == 0 || !(sts || svc || vs || cull)) return NBG_OK;
  CheckArgs a{};
  bool null_off = false;
  auto run = [&](const uint32_t* off, const uint32_t* end, unsigned long long extent) {
    if (!off) null_off = true;
    a.runs[a.n_runs++] = Run{off, end, extent, 1u};
  };
  auto map = [&](const nbg_kvmap& m, unsigned long long extent) { run(m.off, m.end, extent); };
  if (sts) {
    if (!sts->pod_from.mask || !sts->pod_to.mask || !sts->pod_from.off || !sts->pod_to.off) return NBG_E_INVAL;
    a.sparse[a.n_sparse++] = Sparse{Run{sts->pod_from.off, sts->pod_from.end, ext->sts_val_from, 2u}, sts->pod_from.mask};
    a.sparse[a.n_sparse++] = Sparse{Run{sts->pod_to.off, sts->pod_to.end, ext->sts_val_to, 2u}, sts->pod_to.mask};
    map(sts->labels_from, ext->sts_labels_from); map(sts->labels_to, ext->sts_labels_to);
    map(sts->ann_from, ext->sts_ann_from); map(sts->ann_to, ext->sts_ann_to);
  }
  if (svc) { map(svc->labels_to, ext->svc_labels_to); map(svc->ann_to, ext->svc_ann_to); map(svc->sel_to, ext->svc_sel_to); }
  if (vs) { map(vs->hdr_from, ext->vs_hdr_from); map(vs->hdr_to, ext->vs_hdr_to); }
  if (cull) { run(cull->k_off, cull->k_end, ext->cull_kernels); run(cull->t_off, cull->t_end, ext->cull_terminals); }
  if (null_off) return NBG_E_INVAL;
  if (cudaSetDevice(id) != cudaSuccess) { cudaGetLastError(); return NBG_E_CUDA; }
  cudaStream_t s = (cudaStream_t)stream;
  unsigned* flags = nullptr;
  if (cudaMalloc(&flags, sizeof(unsigned)) != cudaSuccess) { cudaGetLastError(); return NBG_E_NOMEM; }
  unsigned host_flags = 0;
  cudaError_t e = cudaMemsetAsync(flags, 0, sizeof(unsigned), s);
  if (e == cudaSuccess) {
    check_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(a, n, flags);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(&host_flags, flags, sizeof(unsigned), cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  cudaFree(flags);
  if (e != cudaSuccess) { cudaGetLastError(); return NBG_E_CUDA; }
  if (what) *what = host_flags;
  return host_flags ? NBG_E_RANGE : NBG_OK;
}
