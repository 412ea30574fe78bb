// join.cu — "which live StatefulSet does each Notebook control?" for a whole fleet.
//
// Reference (components/notebook-controller/controllers/notebook_controller.go:157-170):
//     r.List(ctx, namespacedStatefulSets, client.InNamespace(ss.Namespace))
//     for _, sts := range namespacedStatefulSets.Items {
//         if metav1.IsControlledBy(&sts, instance) { foundStateful = &sts; break }
//     }
// i.e. per Notebook an O(#StatefulSets in the namespace) scan for the FIRST list
// entry whose controller owner reference carries the Notebook's UID.  This is the
// step right before CopyStatefulSetFields; the batch form is a hash join.
//
// Keys are 64-bit: nbg_join_key(hash(namespace), hash(uid)), never 0; a
// StatefulSet without a controller reference has key 0 and never matches.
//   build: every StatefulSet j claims a slot for its key (atomicCAS, linear
//          probing) and records min(j) there — the first list entry wins, as the
//          `break` in the reference does;
//   probe: every Notebook looks its key up.
// The table is SCRATCH memory in HBM (two arrays, `cap` slots, cap a power of two
// >= 2 x #StatefulSets); for the fleets this path sees it fits in the 126 MB L2,
// so both phases run at L2 random-access speed while the key arrays stream.
#include "common.cuh"

namespace nbg {

__device__ __forceinline__ unsigned slot_of(unsigned long long key, unsigned mask) {
  // keys are already well mixed (nbg_combine); fold the high half in anyway
  return (unsigned)((key ^ (key >> 29)) * 0x9E3779B97F4A7C15ull >> 32) & mask;
}

__global__ void __launch_bounds__(256)
join_build_kernel(const unsigned long long* __restrict__ sts_key, unsigned long long m,
                  unsigned long long* __restrict__ t_key, unsigned* __restrict__ t_val, unsigned mask) {
  const unsigned long long j = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  const unsigned long long key = __ldg(sts_key + j);
  if (key == 0ull) return;                         // not controlled by anything
  unsigned s = slot_of(key, mask);
  while (true) {
    const unsigned long long seen = atomicCAS(t_key + s, 0ull, key);
    if (seen == 0ull || seen == key) { atomicMin(t_val + s, (unsigned)j); return; }
    s = (s + 1u) & mask;
  }
}

__global__ void __launch_bounds__(256)
join_probe_kernel(const unsigned long long* __restrict__ nb_key, unsigned long long n,
                  const unsigned long long* __restrict__ t_key, const unsigned* __restrict__ t_val,
                  unsigned mask, uint32_t* __restrict__ match) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long key = __ldg(nb_key + i);
  unsigned found = NBG_ARGMAX_NONE;
  if (key != 0ull) {
    unsigned s = slot_of(key, mask);
    while (true) {
      const unsigned long long seen = t_key[s];
      if (seen == key) { found = t_val[s]; break; }
      if (seen == 0ull) break;
      s = (s + 1u) & mask;
    }
  }
  match[i] = found;
}

// ---------------------------------------------------------------- metrics scrape
// Metrics.scrape (components/notebook-controller/pkg/metrics/metrics.go:82-98): over ALL
// StatefulSets, count per namespace those whose pod-template label "notebook-name" is present
// and equals the StatefulSet's own name — the notebook_running gauge.  A segmented count:
// one StatefulSet per lane, lanes of a warp that share a namespace elect a leader
// (__match_any_sync) which adds the whole group with one atomic.
__global__ void __launch_bounds__(256)
count_running_kernel(const uint32_t* __restrict__ ns_id, const unsigned long long* __restrict__ name_hash,
                     const unsigned long long* __restrict__ label_hash, const uint8_t* __restrict__ flags,
                     unsigned long long m, unsigned n_namespaces, unsigned* __restrict__ counts) {
  const unsigned long long j = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  bool hit = false;
  unsigned ns = 0xFFFFFFFFu;
  if (j < m) {
    ns = __ldg(ns_id + j);
    // `name, ok := labels["notebook-name"]; ok && name == v.Name`
    hit = (__ldg(flags + j) & 1u) && __ldg(label_hash + j) == __ldg(name_hash + j) && ns < n_namespaces;
  }
  const unsigned active = __ballot_sync(FULL, hit);
  if (hit) {
    const unsigned peers = __match_any_sync(active, ns);
    if ((threadIdx.x & 31u) == (unsigned)(__ffs(peers) - 1)) atomicAdd(counts + ns, (unsigned)__popc(peers));
  }
}

}  // namespace nbg

extern "C" cudaError_t nbg_launch_count_running(const uint32_t* ns_id, const uint64_t* name_hash, const uint64_t* label_hash,
                                                const uint8_t* flags, unsigned long long m, unsigned n_namespaces,
                                                uint32_t* counts, cudaStream_t stream) {
  cudaError_t e = cudaMemsetAsync(counts, 0, (size_t)n_namespaces * sizeof(uint32_t), stream);
  if (e != cudaSuccess || m == 0) return e;
  const unsigned block = 256;
  nbg::count_running_kernel<<<(unsigned)((m + block - 1) / block), block, 0, stream>>>(
      ns_id, reinterpret_cast<const unsigned long long*>(name_hash), reinterpret_cast<const unsigned long long*>(label_hash),
      flags, m, n_namespaces, counts);
  return cudaGetLastError();
}

// Table geometry shared by every caller: cap slots, keys then values.
extern "C" unsigned long long nbg_join_table_slots(unsigned long long m) {
  unsigned long long cap = 1024;
  while (cap < 2ull * m) cap <<= 1;
  return cap;
}

// `scratch` must hold nbg_join_table_slots(m) * 12 bytes, 8-byte aligned.  The
// table is cleared here (keys 0, values 0xFFFFFFFF) because every call starts
// from an empty table whatever the last call left behind.
extern "C" cudaError_t nbg_launch_owner_join(const unsigned long long* nb_key, unsigned long long n,
                                             const unsigned long long* sts_key, unsigned long long m,
                                             uint32_t* match, void* scratch, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const unsigned long long cap = nbg_join_table_slots(m);
  unsigned long long* t_key = static_cast<unsigned long long*>(scratch);
  unsigned* t_val = reinterpret_cast<unsigned*>(t_key + cap);
  cudaError_t e = cudaMemsetAsync(t_key, 0x00, cap * sizeof(unsigned long long), stream);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(t_val, 0xFF, cap * sizeof(unsigned), stream);
  if (e != cudaSuccess) return e;
  const unsigned block = 256;
  if (m) nbg::join_build_kernel<<<(unsigned)((m + block - 1) / block), block, 0, stream>>>(sts_key, m, t_key, t_val, (unsigned)(cap - 1));
  nbg::join_probe_kernel<<<(unsigned)((n + block - 1) / block), block, 0, stream>>>(nb_key, n, t_key, t_val, (unsigned)(cap - 1), match);
  return cudaGetLastError();
}
