// common.cuh — device helpers shared by the nbgraft kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../../include/nbgraft.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "nbgraft kernels are written for sm_100a (B200) only"
#endif

namespace nbg {

constexpr unsigned FULL = 0xFFFFFFFFu;

// 128-bit streaming load: read-only path, do not allocate in L1 (every tile
// byte is used exactly once by exactly one thread).
__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ unsigned diff4(const uint4& a, const uint4& b) {
  return (a.x ^ b.x) | (a.y ^ b.y) | (a.z ^ b.z) | (a.w ^ b.w);
}

// One map entry through the read-only path (cached in L1: neighbouring lanes
// own neighbouring CSR segments, so a line is reused by several lanes).
__device__ __forceinline__ nbg_kv ld_kv(const nbg_kv* p) {
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
  nbg_kv e;
  e.key = (uint64_t)v.x | ((uint64_t)v.y << 32);
  e.val = (uint64_t)v.z | ((uint64_t)v.w << 32);
  return e;
}

// The asymmetric loop of COMMON:109-113 / :116-120 / :168-180 on one object:
//   for k, v := range to { if from[k] != v { requireUpdate = true } }
// A key missing from `from` reads as "" (hash 0).  Both segments are sorted by
// key hash, so one merge pass answers every lookup.
// End of object i's run: off[i+1] for a CSR map, end[i] for a map given as spans (nbgraft.h).
// The choice depends on a kernel parameter only, so it is uniform across the grid.  The kernels of
// the reconcile step are compiled twice: SPANS = false reads off[i+1] with no test at all (the
// launcher picks it when every `end` of the batch is NULL), SPANS = true tests each map.
template <bool SPANS = true>
__device__ __forceinline__ uint32_t end_at(const nbg_kvmap& m, unsigned long long i) {
  if (!SPANS) return __ldg(m.off + i + 1);
  return __ldg((m.end ? m.end : m.off + 1) + i);
}
template <bool SPANS = true>
__device__ __forceinline__ uint32_t end_at(const nbg_podsparse& m, unsigned long long i) {
  if (!SPANS) return __ldg(m.off + i + 1);
  return __ldg((m.end ? m.end : m.off + 1) + i);
}

__device__ __forceinline__ bool live_map_differs(const nbg_kv* __restrict__ from_ent,
                                                 uint32_t f_lo, uint32_t f_hi,
                                                 const nbg_kv* __restrict__ to_ent,
                                                 uint32_t t_lo, uint32_t t_hi) {
  bool differs = false;
  uint32_t f = f_lo;
  uint64_t fkey = 0, fval = 0;
  bool fvalid = false;
  for (uint32_t t = t_lo; t < t_hi; ++t) {
    const nbg_kv live = ld_kv(to_ent + t);
    // advance the desired cursor to the first key >= live.key
    while (true) {
      if (!fvalid) {
        if (f >= f_hi) break;
        const nbg_kv d = ld_kv(from_ent + f);
        fkey = d.key; fval = d.val; fvalid = true;
      }
      if (fkey >= live.key) break;
      ++f; fvalid = false;
    }
    const uint64_t want = (fvalid && fkey == live.key) ? fval : 0ull;
    differs |= (want != live.val);
  }
  return differs;
}

// reflect.DeepEqual on two map[string]string given nil flags (COMMON:184, and
// the VS headers map): both nil or both non-nil, same length, same entries.
__device__ __forceinline__ bool map_not_deep_equal(const nbg_kv* __restrict__ a_ent,
                                                   uint32_t a_lo, uint32_t a_hi, bool a_nil,
                                                   const nbg_kv* __restrict__ b_ent,
                                                   uint32_t b_lo, uint32_t b_hi, bool b_nil) {
  if (a_nil != b_nil) return true;
  const uint32_t na = a_hi - a_lo;
  if (na != b_hi - b_lo) return true;
  bool differs = false;
  for (uint32_t i = 0; i < na; ++i) {
    const nbg_kv a = ld_kv(a_ent + a_lo + i);
    const nbg_kv b = ld_kv(b_ent + b_lo + i);
    differs |= (a.key != b.key) | (a.val != b.val);
  }
  return differs;
}

}  // namespace nbg
