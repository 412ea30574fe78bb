// diff.cu — desired-vs-live "needs update" predicates for whole fleets.
//
// Reference semantics (components/common/reconcilehelper/util.go = COMMON):
//   CopyStatefulSetFields  COMMON:107-134   (+ caller pre-step NBCTL:188-194)
//   CopyServiceFields      COMMON:166-195
//   CopyVirtualService     COMMON:199-219
//
// Shape of every kernel: a warp owns 32 consecutive objects.
//   * fixed-width tiles (PodSpec 1536 B, Service ports 128 B, VS spec 64 B per
//     side) are streamed with 128-bit loads that bypass L1, the lanes of a
//     warp covering contiguous bytes; the "any slot differs" reduction is a
//     warp vote (__any_sync / __ballot_sync);
//   * the small per-object string maps (CSR) are walked one object per lane;
//   * lane j ends up holding object j's reason byte, so reason[] is written
//     32 B at a time and the result bitmap one ballot half-word per warp.
// Nothing is reused between objects, so there is no shared-memory staging:
// HBM -> registers -> vote.  No tensor cores: there is no contraction here.
#include "common.cuh"
#include "../hash.h"

namespace nbg {

__device__ __forceinline__ void store_results(unsigned long long base, unsigned long long n,
                                              unsigned lane, unsigned reason, bool update,
                                              uint8_t* __restrict__ reason_out,
                                              uint32_t* __restrict__ bitmap_half) {
  const unsigned long long i = base + lane;
  if (i < n) reason_out[i] = (uint8_t)reason;
  const unsigned bits = __ballot_sync(FULL, (i < n) && update);
  if (lane == 0) {
    bitmap_half[base >> 5] = bits;
    if (((base >> 5) & 1) == 0 && base + 32 >= n) bitmap_half[(base >> 5) + 1] = 0u;
  }
}

// ---------------------------------------------------------------- StatefulSet
constexpr int kPodVec = NBG_POD_SLOTS * 8 / 16;       // 96 x 16 B per side
constexpr int kPodVecPerLane = kPodVec / 32;          // 3

__global__ void __launch_bounds__(128)
diff_sts_kernel(nbg_sts_batch in, unsigned long long n,
                unsigned lf_base, unsigned lt_base, unsigned af_base, unsigned at_base, unsigned deployment,
                uint8_t* __restrict__ reason_out, uint32_t* __restrict__ bitmap_half) {
  const unsigned lane = threadIdx.x & 31;
  const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned long long base = warp * 32ull;
  if (base >= n) return;
  const unsigned long long i = base + lane;
  const unsigned cnt = (unsigned)min(32ull, n - base);

  // ---- per-lane scalar work for object `i`: R, D1', L, A
  unsigned reason = 0;
  if (i < n) {
    const int rf = __ldg(in.replicas_from + i), rt = __ldg(in.replicas_to + i);
    if (deployment) {
      // CopyDeploymentSetFields compares the Replicas POINTERS (COMMON:152): two objects never share one,
      // so they are "different" unless both are nil; nothing is dereferenced, nothing panics
      if (rf != NBG_REPLICAS_NIL || rt != NBG_REPLICAS_NIL) reason |= NBG_STS_REPLICAS;
    } else if (rf == NBG_REPLICAS_NIL || rt == NBG_REPLICAS_NIL) {
      reason |= NBG_STS_NILREPLICAS;                 // Go would panic at COMMON:123
    } else if (rf != rt) {
      reason |= NBG_STS_REPLICAS;                    // COMMON:123-126
      // NBCTL:188-194: template labels are copied only on a replicas change
      if (__ldg(in.tmpl_from + i) != __ldg(in.tmpl_to + i)) reason |= NBG_STS_TMPL_LABELS;
    }
    if (live_map_differs(in.labels_from.ent, __ldg(in.labels_from.off + i) - lf_base,
                         end_at(in.labels_from, i) - lf_base,
                         in.labels_to.ent, __ldg(in.labels_to.off + i) - lt_base,
                         end_at(in.labels_to, i) - lt_base))
      reason |= NBG_STS_LABELS;                      // COMMON:109-113
    if (live_map_differs(in.ann_from.ent, __ldg(in.ann_from.off + i) - af_base,
                         end_at(in.ann_from, i) - af_base,
                         in.ann_to.ent, __ldg(in.ann_to.off + i) - at_base,
                         end_at(in.ann_to, i) - at_base))
      reason |= NBG_STS_ANNOTATIONS;                 // COMMON:116-120
  }

  // ---- warp-cooperative PodSpec compare (COMMON:128): object o's two tiles
  // are 96 x 16 B each; lane l reads vectors l, l+32, l+64 of both.
  const uint4* pf = reinterpret_cast<const uint4*>(in.pod_from) + base * kPodVec + lane;
  const uint4* pt = reinterpret_cast<const uint4*>(in.pod_to) + base * kPodVec + lane;
  unsigned pod_bits = 0;                             // bit o = object o differs
  unsigned o = 0;
  for (; o + 2 <= cnt; o += 2) {                     // two objects (6 KB) in flight per warp
    uint4 a[2 * kPodVecPerLane], b[2 * kPodVecPerLane];
#pragma unroll
    for (int v = 0; v < 2 * kPodVecPerLane; ++v) {
      a[v] = ld_stream(pf + (o * kPodVec) + v * 32);
      b[v] = ld_stream(pt + (o * kPodVec) + v * 32);
    }
    unsigned d0 = 0, d1 = 0;
#pragma unroll
    for (int v = 0; v < kPodVecPerLane; ++v) {
      d0 |= diff4(a[v], b[v]);
      d1 |= diff4(a[v + kPodVecPerLane], b[v + kPodVecPerLane]);
    }
    if (__any_sync(FULL, d0 != 0)) pod_bits |= 1u << o;
    if (__any_sync(FULL, d1 != 0)) pod_bits |= 2u << o;
  }
  if (o < cnt) {
    unsigned d = 0;
#pragma unroll
    for (int v = 0; v < kPodVecPerLane; ++v)
      d |= diff4(ld_stream(pf + (o * kPodVec) + v * 32), ld_stream(pt + (o * kPodVec) + v * 32));
    if (__any_sync(FULL, d != 0)) pod_bits |= 1u << o;
  }
  if ((pod_bits >> lane) & 1u) reason |= NBG_STS_PODSPEC;

  store_results(base, n, lane, reason, (reason & NBG_STS_UPDATE_MASK) != 0, reason_out, bitmap_half);
}

// ------------------------------------------------ StatefulSet, desired side generated on the device
// generateStatefulSet (NBCTL:433-523) as a patch of the NOTEBOOK's own PodSpec tile: the desired PodSpec is a deep
// copy of the Notebook's with at most four edits, all of which land in named slots of the tile (csrc/host/pack.cpp):
//   containers[0].workingDir   "" -> "/home/jovyan"                          slot 42          NBCTL:495-497
//   containers[0].ports        nil -> [{8888, "notebook-port", "TCP"}]       slots 52, 53; tag in 46   NBCTL:498-506
//   containers[0].env          + {NB_PREFIX, /notebook/<ns>/<name>} unless a variable of that name is there
//                                                                            slots 56+2e, 57+2e; tag in 46   NBCTL:417-431
//   securityContext            nil -> {fsGroup: 100} when ADD_FSGROUP         slot 7; presence bits in 14     NBCTL:514-521
// Vector v of a tile holds slots 2v and 2v+1, and lane l of the warp holds vectors l, l+32, l+64 of the object, so
// every edit belongs to one known lane; the two words the edits depend on (the count tags in slot 46, the presence
// bits in slot 14) travel by shuffle, "is NB_PREFIX there" is one ballot over the lanes that hold the env names.
// An object whose env list is too long for the tile to show all its names (14 or more entries) cannot be decided
// here: the host generates its tile and marks it NBG_GEN_PREPATCHED.
struct StsGenConsts { unsigned long long working_dir, port_name, port_rest, env_name; unsigned add_fsgroup; };

__device__ __forceinline__ unsigned long long lo64(const uint4& v) { return (unsigned long long)v.x | ((unsigned long long)v.y << 32); }
__device__ __forceinline__ unsigned long long hi64(const uint4& v) { return (unsigned long long)v.z | ((unsigned long long)v.w << 32); }
__device__ __forceinline__ void set_lo64(uint4& v, unsigned long long x) { v.x = (unsigned)x; v.y = (unsigned)(x >> 32); }
__device__ __forceinline__ void set_hi64(uint4& v, unsigned long long x) { v.z = (unsigned)x; v.w = (unsigned)(x >> 32); }

// a[0..2] = this lane's three vectors of ONE object's Notebook tile; on return, of its desired StatefulSet tile
__device__ __forceinline__ void generate_sts_tile(uint4* a, unsigned lane, unsigned flags, unsigned long long prefix_hash,
                                                  const StsGenConsts& k) {
  // every lane takes part in the shuffles, whatever its object's flags say
  const unsigned long long tag = __shfl_sync(FULL, lo64(a[0]), 23);       // slot 46: ports | env << 16 | mounts << 32 | envFrom << 48
  const unsigned long long presence = __shfl_sync(FULL, lo64(a[0]), 7);   // slot 14
  const unsigned ports_tag = (unsigned)(tag & 0xFFFFu), env_tag = (unsigned)((tag >> 16) & 0xFFFFu);
  const unsigned env_n = env_tag == 0xFFFFu ? 0u : env_tag;               // a nil list has no elements
  const unsigned shown = env_n < 14u ? env_n : 14u;
  bool mine = false;
  if (lane >= 28u) mine = (lane - 28u) < shown && lo64(a[0]) == k.env_name;          // env 0..3: vectors 28..31
  if (lane <= 9u) mine |= (lane + 4u) < shown && lo64(a[1]) == k.env_name;           // env 4..13: vectors 32..41
  const bool has_prefix = __ballot_sync(FULL, mine) != 0u;
  if (flags & NBG_GEN_PREPATCHED) return;
  const bool append = !has_prefix && env_n < 14u;
  if (lane == 21u && lo64(a[0]) == 0ull) set_lo64(a[0], k.working_dir);
  if (lane == 26u && ports_tag == 0xFFFFu) { set_lo64(a[0], k.port_name); set_hi64(a[0], k.port_rest); }
  if (append) {
    const unsigned v = 28u + env_n;
    if (lane == (v & 31u)) {                                             // env 0..3 in the lane's first vector, 4..13 in its second
      if (v < 32u) { set_lo64(a[0], k.env_name); set_hi64(a[0], prefix_hash); }
      else { set_lo64(a[1], k.env_name); set_hi64(a[1], prefix_hash); }
    }
  }
  if (lane == 23u) {
    const unsigned long long ports_new = ports_tag == 0xFFFFu ? 1ull : ports_tag;
    const unsigned long long env_new = append ? env_n + 1ull : env_tag;
    set_lo64(a[0], (tag & ~0xFFFFFFFFull) | ports_new | (env_new << 16));
  }
  if (k.add_fsgroup && !(presence & 2ull)) {                               // securityContext == nil
    if (lane == 3u) set_hi64(a[0], 100ull);                                // slot 7: *fsGroup
    if (lane == 4u) set_hi64(a[0], 0ull);                                  // slot 9: what the struct holds besides fsGroup/runAsUser - nothing (was: the nil marker)
    if (lane == 7u) set_lo64(a[0], presence | 6ull);                       // slot 14: securityContext and fsGroup are pointers now
  }
}

__global__ void __launch_bounds__(128)
diff_sts_gen_kernel(nbg_sts_batch in, const unsigned long long* __restrict__ prefix_hash, const uint8_t* __restrict__ gen_flags,
                    StsGenConsts k, unsigned long long n, unsigned lf_base, unsigned lt_base, unsigned at_base,
                    uint8_t* __restrict__ reason_out, uint32_t* __restrict__ bitmap_half) {
  const unsigned lane = threadIdx.x & 31;
  const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned long long base = warp * 32ull;
  if (base >= n) return;
  const unsigned long long i = base + lane;
  const unsigned cnt = (unsigned)min(32ull, n - base);
  unsigned reason = 0, my_flags = 0;
  unsigned long long my_prefix = 0;
  if (i < n) {
    my_flags = __ldg(gen_flags + i);
    my_prefix = __ldg(prefix_hash + i);
    const int rf = (my_flags & NBG_GEN_STOPPED) ? 0 : 1, rt = __ldg(in.replicas_to + i);     // NBCTL:434-437
    if (rt == NBG_REPLICAS_NIL) {
      reason |= NBG_STS_NILREPLICAS;
    } else if (rf != rt) {
      reason |= NBG_STS_REPLICAS;
      if (__ldg(in.tmpl_from + i) != __ldg(in.tmpl_to + i)) reason |= NBG_STS_TMPL_LABELS;
    }
    if (live_map_differs(in.labels_from.ent, __ldg(in.labels_from.off + i) - lf_base, end_at(in.labels_from, i) - lf_base,
                         in.labels_to.ent, __ldg(in.labels_to.off + i) - lt_base, end_at(in.labels_to, i) - lt_base))
      reason |= NBG_STS_LABELS;
    // the desired StatefulSet has no annotations (NBCTL:439-443): any live one with a non-empty value differs
    for (unsigned t = __ldg(in.ann_to.off + i) - at_base, e = end_at(in.ann_to, i) - at_base; t < e; ++t)
      if (ld_kv(in.ann_to.ent + t).val != 0ull) reason |= NBG_STS_ANNOTATIONS;
  }
  const uint4* pf = reinterpret_cast<const uint4*>(in.pod_from) + base * kPodVec + lane;
  const uint4* pt = reinterpret_cast<const uint4*>(in.pod_to) + base * kPodVec + lane;
  unsigned pod_bits = 0;
  for (unsigned o = 0; o < cnt; ++o) {
    uint4 a[kPodVecPerLane], b[kPodVecPerLane];
#pragma unroll
    for (int v = 0; v < kPodVecPerLane; ++v) {
      a[v] = ld_stream(pf + (o * kPodVec) + v * 32);
      b[v] = ld_stream(pt + (o * kPodVec) + v * 32);
    }
    generate_sts_tile(a, lane, __shfl_sync(FULL, my_flags, o), __shfl_sync(FULL, my_prefix, o), k);
    unsigned d = 0;
#pragma unroll
    for (int v = 0; v < kPodVecPerLane; ++v) d |= diff4(a[v], b[v]);
    if (__any_sync(FULL, d != 0)) pod_bits |= 1u << o;
  }
  if ((pod_bits >> lane) & 1u) reason |= NBG_STS_PODSPEC;
  store_results(base, n, lane, reason, (reason & NBG_STS_UPDATE_MASK) != 0, reason_out, bitmap_half);
}

// ------------------------------------------------ StatefulSet, sparse tiles
// Same predicate, PodSpec tiles stored as (192-bit mask, packed non-zero slots).
// Lane j compares object j's masks itself (2 x 24 B); objects whose masks agree
// have equally long value runs, which the warp then compares run by run: runs
// are padded to an even number of values and start 16-byte aligned, so a run of
// up to 64 values is ONE 128-bit streaming load per lane and side.  Four
// objects are in flight per warp.  Objects whose masks differ are decided
// without their values ever being read.
constexpr unsigned kSparseGroup = 4;

// bit o of the result = the sparse tiles of object base+o differ (warp-cooperative)
template <bool SPANS = true>
__device__ __forceinline__ unsigned sparse_tiles_differ(const nbg_podsparse& a, const nbg_podsparse& b,
                                                        unsigned long long i, bool live, unsigned cnt, unsigned lane,
                                                        unsigned a_base, unsigned b_base) {
  unsigned va0 = 0, vb0 = 0, my_vecs = 0;               // run starts and length, in 16-byte vectors
  bool mask_ne = false;
  if (live) {
    const unsigned long long* ma = reinterpret_cast<const unsigned long long*>(a.mask) + i * 3;
    const unsigned long long* mb = reinterpret_cast<const unsigned long long*>(b.mask) + i * 3;
    mask_ne = (__ldg(ma) != __ldg(mb)) | (__ldg(ma + 1) != __ldg(mb + 1)) | (__ldg(ma + 2) != __ldg(mb + 2));
    const unsigned f0 = __ldg(a.off + i) - a_base, f1 = end_at<SPANS>(a, i) - a_base;
    va0 = f0 >> 1;
    vb0 = (__ldg(b.off + i) - b_base) >> 1;
    my_vecs = mask_ne ? 0u : (f1 - f0) >> 1;
  }
  unsigned bits = __ballot_sync(FULL, mask_ne);
  const uint4* av = reinterpret_cast<const uint4*>(a.val);
  const uint4* bv = reinterpret_cast<const uint4*>(b.val);
  for (unsigned o = 0; o < cnt; o += kSparseGroup) {
    unsigned d[kSparseGroup];
    uint4 x[kSparseGroup], y[kSparseGroup];
    unsigned more = 0;                                   // some run of this group is longer than 32 vectors
#pragma unroll
    for (unsigned g = 0; g < kSparseGroup; ++g) {
      const unsigned src = min(o + g, 31u);
      const unsigned vecs = __shfl_sync(FULL, my_vecs, src);   // 0 past the end (my_vecs == 0 there)
      const unsigned ao = __shfl_sync(FULL, va0, src), bo = __shfl_sync(FULL, vb0, src);
      x[g] = make_uint4(0, 0, 0, 0); y[g] = x[g];
      if (lane < vecs) { x[g] = ld_stream(av + ao + lane); y[g] = ld_stream(bv + bo + lane); }
      more |= vecs > 32u;
    }
#pragma unroll
    for (unsigned g = 0; g < kSparseGroup; ++g) d[g] = diff4(x[g], y[g]);
    if (more) {                                          // warp-uniform; runs of 65..192 values
#pragma unroll
      for (unsigned g = 0; g < kSparseGroup; ++g) {
        const unsigned src = min(o + g, 31u);
        const unsigned vecs = __shfl_sync(FULL, my_vecs, src);
        const unsigned ao = __shfl_sync(FULL, va0, src), bo = __shfl_sync(FULL, vb0, src);
        for (unsigned k = lane + 32u; k < vecs; k += 32u) d[g] |= diff4(ld_stream(av + ao + k), ld_stream(bv + bo + k));
      }
    }
#pragma unroll
    for (unsigned g = 0; g < kSparseGroup; ++g)
      if (__any_sync(FULL, d[g] != 0)) bits |= 1u << min(o + g, 31u);
  }
  return bits;
}

// CSR form only.  There the value runs of a warp's 32 objects are ONE contiguous block per side, so the
// warp sweeps the desired side's block 32 vectors per round, fully coalesced, instead of going object by
// object: each lane finds the object its vector belongs to (binary search over the lanes' run starts, five
// shuffles), reads the matching vector of the live side and records a difference under that object's bit.
// Rounds do not depend on one another, so several are in flight at once.  Objects whose masks differ are
// decided without their values being read, as above.
#ifndef NBG_CSR_UNROLL
#define NBG_CSR_UNROLL 2
#endif
constexpr int kCsrUnroll = NBG_CSR_UNROLL;
__device__ __forceinline__ unsigned sparse_tiles_differ_csr(const nbg_podsparse& a, const nbg_podsparse& b,
                                                            unsigned long long i, bool live, unsigned cnt, unsigned lane,
                                                            unsigned a_base, unsigned b_base) {
  unsigned fv0 = 0, fv1 = 0, tv0 = 0;                    // run start/end (desired) and start (live), in 16-byte vectors
  bool mask_ne = false;
  if (live) {
    const unsigned long long* ma = reinterpret_cast<const unsigned long long*>(a.mask) + i * 3;
    const unsigned long long* mb = reinterpret_cast<const unsigned long long*>(b.mask) + i * 3;
    mask_ne = (__ldg(ma) != __ldg(mb)) | (__ldg(ma + 1) != __ldg(mb + 1)) | (__ldg(ma + 2) != __ldg(mb + 2));
    fv0 = (__ldg(a.off + i) - a_base) >> 1;
    fv1 = (__ldg(a.off + i + 1) - a_base) >> 1;
    tv0 = (__ldg(b.off + i) - b_base) >> 1;
  }
  const unsigned ne_bits = __ballot_sync(FULL, mask_ne);
  const unsigned first = __shfl_sync(FULL, fv0, 0), last = __shfl_sync(FULL, fv1, cnt - 1);
  if (!live) fv0 = last;                                 // lanes past the end of the batch never own a vector
  const uint4* av = reinterpret_cast<const uint4*>(a.val);
  const uint4* bv = reinterpret_cast<const uint4*>(b.val);
  unsigned mine = 0;
  for (unsigned vb = first; vb < last; vb += 32u * kCsrUnroll) {   // warp-uniform trip count; kCsrUnroll rounds per trip
    unsigned obj[kCsrUnroll];
    uint4 x[kCsrUnroll], y[kCsrUnroll];
#pragma unroll
    for (int k = 0; k < kCsrUnroll; ++k) {               // all the rounds' loads are issued before any is compared
      const unsigned v = vb + 32u * k + lane;
      unsigned o = 0;                                    // largest lane whose run starts at or before v
#pragma unroll
      for (unsigned step = 16u; step; step >>= 1) {
        const unsigned cand = __shfl_sync(FULL, fv0, o + step);
        if (cand <= v) o += step;
      }
      const unsigned f_start = __shfl_sync(FULL, fv0, o), t_start = __shfl_sync(FULL, tv0, o);
      obj[k] = o;
      x[k] = make_uint4(0, 0, 0, 0);
      y[k] = x[k];
      if (v < last && !((ne_bits >> o) & 1u)) {
        x[k] = ld_stream(av + v);
        y[k] = ld_stream(bv + t_start + (v - f_start));
      }
    }
#pragma unroll
    for (int k = 0; k < kCsrUnroll; ++k)
      if (diff4(x[k], y[k])) mine |= 1u << obj[k];
  }
  return ne_bits | __reduce_or_sync(FULL, mine);
}

#ifdef NBG_STS_BULK
// ---- experiment (tools/build_variant.sh ... -DNBG_STS_BULK): the same sweep with each warp's two value blocks
// brought into shared memory by the bulk-copy engine (cp.async.bulk global -> shared, completion on an mbarrier)
// instead of per-lane 128-bit loads.  One elected lane issues two copies per tile; address generation for the
// bulk of the bytes leaves the issue slots.  A tile whose blocks exceed the budget, or whose value arrays are not
// 16-byte aligned, takes the register path.  Measured against the register path in DESIGN.md section 3.
#ifndef NBG_BULK_VECS
#define NBG_BULK_VECS 640                      // 16-byte vectors per side and warp (10 KB)
#endif
constexpr unsigned kBulkVecs = NBG_BULK_VECS;
constexpr unsigned kBulkWarps = 4;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ unsigned sparse_tiles_differ_bulk(const nbg_podsparse& a, const nbg_podsparse& b,
                                                             unsigned long long i, bool live, unsigned cnt, unsigned lane,
                                                             unsigned a_base, unsigned b_base, uint4* sa, uint4* sb,
                                                             unsigned long long* bar, bool& staged) {
  unsigned fv0 = 0, fv1 = 0, tv0 = 0, tv1 = 0;
  bool mask_ne = false;
  if (live) {
    const unsigned long long* ma = reinterpret_cast<const unsigned long long*>(a.mask) + i * 3;
    const unsigned long long* mb = reinterpret_cast<const unsigned long long*>(b.mask) + i * 3;
    mask_ne = (__ldg(ma) != __ldg(mb)) | (__ldg(ma + 1) != __ldg(mb + 1)) | (__ldg(ma + 2) != __ldg(mb + 2));
    fv0 = (__ldg(a.off + i) - a_base) >> 1;
    fv1 = (__ldg(a.off + i + 1) - a_base) >> 1;
    tv0 = (__ldg(b.off + i) - b_base) >> 1;
    tv1 = (__ldg(b.off + i + 1) - b_base) >> 1;
  }
  const unsigned ne_bits = __ballot_sync(FULL, mask_ne);
  const unsigned first = __shfl_sync(FULL, fv0, 0), last = __shfl_sync(FULL, fv1, cnt - 1);
  const unsigned tfirst = __shfl_sync(FULL, tv0, 0), tlast = __shfl_sync(FULL, tv1, cnt - 1);
  if (!live) fv0 = last;
  const uint4* av = reinterpret_cast<const uint4*>(a.val);
  const uint4* bv = reinterpret_cast<const uint4*>(b.val);
  staged = (last - first) <= kBulkVecs && (tlast - tfirst) <= kBulkVecs &&
           ((reinterpret_cast<unsigned long long>(av) | reinterpret_cast<unsigned long long>(bv)) & 15ull) == 0;
  if (!staged) return ~0u;                               // caller takes the register path
  const unsigned bytes_a = (last - first) * 16u, bytes_b = (tlast - tfirst) * 16u;
  if (lane == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes_a + bytes_b) : "memory");
    if (bytes_a)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   :: "r"(smem_u32(sa)), "l"(av + first), "r"(bytes_a), "r"(smem_u32(bar)) : "memory");
    if (bytes_b)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   :: "r"(smem_u32(sb)), "l"(bv + tfirst), "r"(bytes_b), "r"(smem_u32(bar)) : "memory");
  }
  return ne_bits;                                        // the copies are in flight; finish with sparse_bulk_compare
}

__device__ __forceinline__ unsigned sparse_bulk_compare(const nbg_podsparse& a, const nbg_podsparse& b, unsigned long long i,
                                                        bool live, unsigned cnt, unsigned lane, unsigned a_base, unsigned b_base,
                                                        const uint4* sa, const uint4* sb, unsigned long long* bar, unsigned ne_bits) {
  unsigned fv0 = 0, fv1 = 0, tv0 = 0;
  if (live) {                                            // the offsets again (L1 hits): cheaper than keeping them alive
    fv0 = (__ldg(a.off + i) - a_base) >> 1;
    fv1 = (__ldg(a.off + i + 1) - a_base) >> 1;
    tv0 = (__ldg(b.off + i) - b_base) >> 1;
  }
  const unsigned first = __shfl_sync(FULL, fv0, 0), last = __shfl_sync(FULL, fv1, cnt - 1);
  const unsigned tfirst = __shfl_sync(FULL, tv0, 0);
  if (!live) fv0 = last;
  unsigned done = 0;
  while (!done)                                          // phase 0 of a barrier used once per kernel
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(smem_u32(bar)) : "memory");
  unsigned mine = 0;
  for (unsigned vb = first; vb < last; vb += 32u) {
    const unsigned v = vb + lane;
    unsigned o = 0;
#pragma unroll
    for (unsigned step = 16u; step; step >>= 1) {
      const unsigned cand = __shfl_sync(FULL, fv0, o + step);
      if (cand <= v) o += step;
    }
    const unsigned f_start = __shfl_sync(FULL, fv0, o), t_start = __shfl_sync(FULL, tv0, o);
    if (v < last && !((ne_bits >> o) & 1u)) {
      const uint4 x = sa[v - first];
      const uint4 y = sb[t_start + (v - f_start) - tfirst];
      if (diff4(x, y)) mine |= 1u << o;
    }
  }
  return ne_bits | __reduce_or_sync(FULL, mine);
}

__global__ void __launch_bounds__(32 * kBulkWarps)
diff_sts_sparse_bulk_kernel(nbg_sts_sparse_batch in, unsigned long long n,
                            unsigned lf_base, unsigned lt_base, unsigned af_base, unsigned at_base,
                            unsigned vf_base, unsigned vt_base,
                            uint8_t* __restrict__ reason_out, uint32_t* __restrict__ bitmap_half) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint4* sa = reinterpret_cast<uint4*>(smem_raw) + (size_t)wid * 2 * kBulkVecs;
  uint4* sb = sa + kBulkVecs;
  unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem_raw + (size_t)kBulkWarps * 2 * kBulkVecs * 16) + wid;
  if (lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned long long base = warp * 32ull;
  if (base >= n) return;
  const unsigned long long i = base + lane;
  const unsigned cnt = (unsigned)min(32ull, n - base);
  bool staged;
  const unsigned ne_bits = sparse_tiles_differ_bulk(in.pod_from, in.pod_to, i, i < n, cnt, lane, vf_base, vt_base, sa, sb, bar, staged);
  // the scalar work runs while the copies are in flight
  unsigned reason = 0;
  if (i < n) {
    const int rf = __ldg(in.replicas_from + i), rt = __ldg(in.replicas_to + i);
    if (rf == NBG_REPLICAS_NIL || rt == NBG_REPLICAS_NIL) {
      reason |= NBG_STS_NILREPLICAS;
    } else if (rf != rt) {
      reason |= NBG_STS_REPLICAS;
      if (__ldg(in.tmpl_from + i) != __ldg(in.tmpl_to + i)) reason |= NBG_STS_TMPL_LABELS;
    }
    if (live_map_differs(in.labels_from.ent, __ldg(in.labels_from.off + i) - lf_base, end_at<false>(in.labels_from, i) - lf_base,
                         in.labels_to.ent, __ldg(in.labels_to.off + i) - lt_base, end_at<false>(in.labels_to, i) - lt_base))
      reason |= NBG_STS_LABELS;
    if (live_map_differs(in.ann_from.ent, __ldg(in.ann_from.off + i) - af_base, end_at<false>(in.ann_from, i) - af_base,
                         in.ann_to.ent, __ldg(in.ann_to.off + i) - at_base, end_at<false>(in.ann_to, i) - at_base))
      reason |= NBG_STS_ANNOTATIONS;
  }
  __syncwarp();
  const unsigned pod_bits = staged ? sparse_bulk_compare(in.pod_from, in.pod_to, i, i < n, cnt, lane, vf_base, vt_base, sa, sb, bar, ne_bits)
                                   : sparse_tiles_differ_csr(in.pod_from, in.pod_to, i, i < n, cnt, lane, vf_base, vt_base);
  if ((pod_bits >> lane) & 1u) reason |= NBG_STS_PODSPEC;
  store_results(base, n, lane, reason, (reason & NBG_STS_UPDATE_MASK) != 0, reason_out, bitmap_half);
}
#endif  // NBG_STS_BULK

template <bool SPANS>
__global__ void __launch_bounds__(128)
diff_sts_sparse_kernel(nbg_sts_sparse_batch in, unsigned long long n,
                       unsigned lf_base, unsigned lt_base, unsigned af_base, unsigned at_base,
                       unsigned vf_base, unsigned vt_base,
                       uint8_t* __restrict__ reason_out, uint32_t* __restrict__ bitmap_half) {
  const unsigned lane = threadIdx.x & 31;
  const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned long long base = warp * 32ull;
  if (base >= n) return;
  const unsigned long long i = base + lane;
  const unsigned cnt = (unsigned)min(32ull, n - base);

  // tiles first (the streaming part), then the per-lane scalar work: keeps the two register sets apart
  unsigned pod_bits;
  if constexpr (SPANS) pod_bits = sparse_tiles_differ<true>(in.pod_from, in.pod_to, i, i < n, cnt, lane, vf_base, vt_base);
  else pod_bits = sparse_tiles_differ_csr(in.pod_from, in.pod_to, i, i < n, cnt, lane, vf_base, vt_base);
  unsigned reason = ((pod_bits >> lane) & 1u) ? NBG_STS_PODSPEC : 0u;
  if (i < n) {
    const int rf = __ldg(in.replicas_from + i), rt = __ldg(in.replicas_to + i);
    if (rf == NBG_REPLICAS_NIL || rt == NBG_REPLICAS_NIL) {
      reason |= NBG_STS_NILREPLICAS;
    } else if (rf != rt) {
      reason |= NBG_STS_REPLICAS;
      if (__ldg(in.tmpl_from + i) != __ldg(in.tmpl_to + i)) reason |= NBG_STS_TMPL_LABELS;
    }
    if (live_map_differs(in.labels_from.ent, __ldg(in.labels_from.off + i) - lf_base,
                         end_at<SPANS>(in.labels_from, i) - lf_base,
                         in.labels_to.ent, __ldg(in.labels_to.off + i) - lt_base,
                         end_at<SPANS>(in.labels_to, i) - lt_base))
      reason |= NBG_STS_LABELS;
    if (live_map_differs(in.ann_from.ent, __ldg(in.ann_from.off + i) - af_base,
                         end_at<SPANS>(in.ann_from, i) - af_base,
                         in.ann_to.ent, __ldg(in.ann_to.off + i) - at_base,
                         end_at<SPANS>(in.ann_to, i) - at_base))
      reason |= NBG_STS_ANNOTATIONS;
  }
  store_results(base, n, lane, reason, (reason & NBG_STS_UPDATE_MASK) != 0, reason_out, bitmap_half);
}

// ------------------------------------------------ webhook restart-block predicate
// maybeRestartRunningNotebook (odh notebook_webhook.go:505-564) for a fleet: three PodSpecs per
// Notebook — old, updated (what the user sent) and mutated (after the webhook's injections) —
// as SEMANTIC tiles (pack_podspec(semantic_canon(x)): equal tiles <=> equality.Semantic.DeepEqual).
//   create / stopped / restarting           -> let the update through           (:515-531)
//   old != updated                          -> the user already restarts the pod (:548-551)
//   old == updated == mutated               -> nothing changes                   (:554-557)
//   old == updated != mutated               -> BLOCK: keep the updated spec      (:559-563)
// Three sparse tiles a, b, c per object: which objects have b != a, which have c != a.  Same
// scheme as sparse_tiles_differ, with a's run loaded ONCE and compared against both others; a side
// whose mask already differs from a's is decided without its values being read.
#ifndef NBG_RB_GROUP
#define NBG_RB_GROUP 2
#endif
__device__ __forceinline__ void sparse_tiles_differ3(const nbg_podsparse& a, const nbg_podsparse& b, const nbg_podsparse& c,
                                                     unsigned long long i, bool live, unsigned cnt, unsigned lane,
                                                     unsigned a_base, unsigned b_base, unsigned c_base,
                                                     unsigned& ab_bits, unsigned& ac_bits) {
  constexpr unsigned G = NBG_RB_GROUP;
  unsigned va0 = 0, vb0 = 0, vc0 = 0, vecs_b = 0, vecs_c = 0;
  bool b_ne = false, c_ne = false;
  if (live) {
    const unsigned long long* ma = reinterpret_cast<const unsigned long long*>(a.mask) + i * 3;
    const unsigned long long* mb = reinterpret_cast<const unsigned long long*>(b.mask) + i * 3;
    const unsigned long long* mc = reinterpret_cast<const unsigned long long*>(c.mask) + i * 3;
    const unsigned long long a0 = __ldg(ma), a1 = __ldg(ma + 1), a2 = __ldg(ma + 2);
    b_ne = (a0 != __ldg(mb)) | (a1 != __ldg(mb + 1)) | (a2 != __ldg(mb + 2));
    c_ne = (a0 != __ldg(mc)) | (a1 != __ldg(mc + 1)) | (a2 != __ldg(mc + 2));
    const unsigned f0 = __ldg(a.off + i) - a_base, f1 = end_at(a, i) - a_base;
    va0 = f0 >> 1;
    vb0 = (__ldg(b.off + i) - b_base) >> 1;
    vc0 = (__ldg(c.off + i) - c_base) >> 1;
    vecs_b = b_ne ? 0u : (f1 - f0) >> 1;
    vecs_c = c_ne ? 0u : (f1 - f0) >> 1;
  }
  ab_bits = __ballot_sync(FULL, b_ne);
  ac_bits = __ballot_sync(FULL, c_ne);
  const uint4* av = reinterpret_cast<const uint4*>(a.val);
  const uint4* bv = reinterpret_cast<const uint4*>(b.val);
  const uint4* cv = reinterpret_cast<const uint4*>(c.val);
  for (unsigned o = 0; o < cnt; o += G) {
    unsigned db[G], dc[G];
    uint4 x[G], y[G], z[G];
    unsigned more = 0;
#pragma unroll
    for (unsigned g = 0; g < G; ++g) {
      const unsigned src = min(o + g, 31u);
      const unsigned nb = __shfl_sync(FULL, vecs_b, src), nc = __shfl_sync(FULL, vecs_c, src);
      const unsigned ao = __shfl_sync(FULL, va0, src), bo = __shfl_sync(FULL, vb0, src), co = __shfl_sync(FULL, vc0, src);
      x[g] = make_uint4(0, 0, 0, 0);
      if (lane < max(nb, nc)) x[g] = ld_stream(av + ao + lane);
      y[g] = x[g]; z[g] = x[g];                          // a side that is not read compares equal
      if (lane < nb) y[g] = ld_stream(bv + bo + lane);
      if (lane < nc) z[g] = ld_stream(cv + co + lane);
      more |= max(nb, nc) > 32u;
    }
#pragma unroll
    for (unsigned g = 0; g < G; ++g) { db[g] = diff4(x[g], y[g]); dc[g] = diff4(x[g], z[g]); }
    if (more) {                                          // warp-uniform; runs of 65..192 values
#pragma unroll
      for (unsigned g = 0; g < G; ++g) {
        const unsigned src = min(o + g, 31u);
        const unsigned nb = __shfl_sync(FULL, vecs_b, src), nc = __shfl_sync(FULL, vecs_c, src);
        const unsigned ao = __shfl_sync(FULL, va0, src), bo = __shfl_sync(FULL, vb0, src), co = __shfl_sync(FULL, vc0, src);
        for (unsigned k = lane + 32u; k < max(nb, nc); k += 32u) {
          const uint4 xa = ld_stream(av + ao + k);
          if (k < nb) db[g] |= diff4(xa, ld_stream(bv + bo + k));
          if (k < nc) dc[g] |= diff4(xa, ld_stream(cv + co + k));
        }
      }
    }
#pragma unroll
    for (unsigned g = 0; g < G; ++g) {
      if (__any_sync(FULL, db[g] != 0)) ab_bits |= 1u << min(o + g, 31u);
      if (__any_sync(FULL, dc[g] != 0)) ac_bits |= 1u << min(o + g, 31u);
    }
  }
}

// The same comparison for CSR batches, as a sweep over a's contiguous block (see sparse_tiles_differ_csr).
__device__ __forceinline__ void sparse_tiles_differ3_csr(const nbg_podsparse& a, const nbg_podsparse& b, const nbg_podsparse& c,
                                                         unsigned long long i, bool live, unsigned cnt, unsigned lane,
                                                         unsigned a_base, unsigned b_base, unsigned c_base,
                                                         unsigned& ab_bits, unsigned& ac_bits) {
  unsigned av0 = 0, av1 = 0, bv0 = 0, cv0 = 0;
  bool b_ne = false, c_ne = false;
  if (live) {
    const unsigned long long* ma = reinterpret_cast<const unsigned long long*>(a.mask) + i * 3;
    const unsigned long long* mb = reinterpret_cast<const unsigned long long*>(b.mask) + i * 3;
    const unsigned long long* mc = reinterpret_cast<const unsigned long long*>(c.mask) + i * 3;
    const unsigned long long a0 = __ldg(ma), a1 = __ldg(ma + 1), a2 = __ldg(ma + 2);
    b_ne = (a0 != __ldg(mb)) | (a1 != __ldg(mb + 1)) | (a2 != __ldg(mb + 2));
    c_ne = (a0 != __ldg(mc)) | (a1 != __ldg(mc + 1)) | (a2 != __ldg(mc + 2));
    av0 = (__ldg(a.off + i) - a_base) >> 1;
    av1 = (__ldg(a.off + i + 1) - a_base) >> 1;
    bv0 = (__ldg(b.off + i) - b_base) >> 1;
    cv0 = (__ldg(c.off + i) - c_base) >> 1;
  }
  const unsigned b_skip = __ballot_sync(FULL, b_ne), c_skip = __ballot_sync(FULL, c_ne);
  const unsigned first = __shfl_sync(FULL, av0, 0), last = __shfl_sync(FULL, av1, cnt - 1);
  if (!live) av0 = last;
  const uint4* av = reinterpret_cast<const uint4*>(a.val);
  const uint4* bv = reinterpret_cast<const uint4*>(b.val);
  const uint4* cv = reinterpret_cast<const uint4*>(c.val);
  unsigned mine_b = 0, mine_c = 0;
  for (unsigned vb = first; vb < last; vb += 32u) {      // warp-uniform trip count
    const unsigned v = vb + lane;
    unsigned o = 0;
#pragma unroll
    for (unsigned step = 16u; step; step >>= 1) {
      const unsigned cand = __shfl_sync(FULL, av0, o + step);
      if (cand <= v) o += step;
    }
    const unsigned a_start = __shfl_sync(FULL, av0, o), b_start = __shfl_sync(FULL, bv0, o), c_start = __shfl_sync(FULL, cv0, o);
    const bool need_b = v < last && !((b_skip >> o) & 1u), need_c = v < last && !((c_skip >> o) & 1u);
    if (need_b || need_c) {
      const uint4 x = ld_stream(av + v);
      uint4 y = x, z = x;
      if (need_b) y = ld_stream(bv + b_start + (v - a_start));
      if (need_c) z = ld_stream(cv + c_start + (v - a_start));
      if (diff4(x, y)) mine_b |= 1u << o;
      if (diff4(x, z)) mine_c |= 1u << o;
    }
  }
  ab_bits = b_skip | __reduce_or_sync(FULL, mine_b);
  ac_bits = c_skip | __reduce_or_sync(FULL, mine_c);
}

template <bool SPANS>
__global__ void __launch_bounds__(128)
restart_block_kernel(nbg_restart_batch in, unsigned long long n, unsigned o_base, unsigned u_base, unsigned m_base,
                     uint8_t* __restrict__ decision_out, uint32_t* __restrict__ bitmap_half) {
  const unsigned lane = threadIdx.x & 31;
  const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned long long base = warp * 32ull;
  if (base >= n) return;
  const unsigned long long i = base + lane;
  const unsigned cnt = (unsigned)min(32ull, n - base);
  const bool live = i < n;
  unsigned upd_bits, mut_bits;
  if constexpr (SPANS)
    sparse_tiles_differ3(in.old_spec, in.updated_spec, in.mutated_spec, i, live, cnt, lane, o_base, u_base, m_base, upd_bits, mut_bits);
  else
    sparse_tiles_differ3_csr(in.old_spec, in.updated_spec, in.mutated_spec, i, live, cnt, lane, o_base, u_base, m_base, upd_bits, mut_bits);
  const unsigned fl = live ? __ldg(in.flags + i) : 0u;
  unsigned decision = 0;
  if (live && !(fl & (NBG_RB_CREATE | NBG_RB_STOPPED | NBG_RB_RESTARTING))) {
    if ((upd_bits >> lane) & 1u) decision = NBG_RB_EXTERNAL_CHANGE;
    else if (!((mut_bits >> lane) & 1u)) decision = NBG_RB_NO_CHANGE;
    else decision = NBG_RB_BLOCKED;
  }
  store_results(base, n, lane, decision, (decision & NBG_RB_BLOCKED) != 0, decision_out, bitmap_half);
}

// ------------------------------------------------- small fixed tiles (Svc, VS)
// VEC = 16-byte vectors per side per object (8 for Service ports, 4 for the VS
// spec).  32/VEC objects are compared per load instruction; the warp covers its
// 32 objects in VEC rounds.  Returns a mask: bit o = object o's tiles differ.
template <int VEC>
__device__ __forceinline__ unsigned tiles_differ(const uint64_t* __restrict__ from,
                                                 const uint64_t* __restrict__ to,
                                                 unsigned long long base, unsigned long long n,
                                                 unsigned lane) {
  constexpr int OBJ_PER_ROUND = 32 / VEC;
  const uint4* pf = reinterpret_cast<const uint4*>(from) + base * VEC;
  const uint4* pt = reinterpret_cast<const uint4*>(to) + base * VEC;
  const unsigned long long limit = (n - base) * VEC;   // vectors that exist past `base`
  unsigned mask = 0;
#pragma unroll
  for (int r = 0; r < VEC; ++r) {
    const unsigned idx = r * 32 + lane;                // vector index within the warp's run
    unsigned d = 0;
    if (idx < limit) d = diff4(ld_stream(pf + idx), ld_stream(pt + idx));
    const unsigned ballot = __ballot_sync(FULL, d != 0);
#pragma unroll
    for (int g = 0; g < OBJ_PER_ROUND; ++g) {
      const unsigned group = (ballot >> (g * VEC)) & ((VEC == 32) ? FULL : ((1u << VEC) - 1u));
      if (group) mask |= 1u << (r * OBJ_PER_ROUND + g);
    }
  }
  return mask;
}

__global__ void __launch_bounds__(128)
diff_svc_kernel(nbg_svc_batch in, unsigned long long n,
                unsigned lf_base, unsigned lt_base, unsigned af_base, unsigned at_base,
                unsigned sf_base, unsigned st_base,
                uint8_t* __restrict__ reason_out, uint32_t* __restrict__ bitmap_half) {
  const unsigned lane = threadIdx.x & 31;
  const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned long long base = warp * 32ull;
  if (base >= n) return;
  const unsigned long long i = base + lane;

  const unsigned port_bits = tiles_differ<NBG_SVC_SLOTS / 2>(in.ports_from, in.ports_to, base, n, lane);

  unsigned reason = 0;
  if (i < n) {
    const unsigned fl = __ldg(in.flags + i);
    if (live_map_differs(in.labels_from.ent, __ldg(in.labels_from.off + i) - lf_base,
                         end_at(in.labels_from, i) - lf_base,
                         in.labels_to.ent, __ldg(in.labels_to.off + i) - lt_base,
                         end_at(in.labels_to, i) - lt_base))
      reason |= NBG_SVC_LABELS;                      // COMMON:168-172
    if (live_map_differs(in.ann_from.ent, __ldg(in.ann_from.off + i) - af_base,
                         end_at(in.ann_from, i) - af_base,
                         in.ann_to.ent, __ldg(in.ann_to.off + i) - at_base,
                         end_at(in.ann_to, i) - at_base))
      reason |= NBG_SVC_ANNOTATIONS;                 // COMMON:175-179
    if (map_not_deep_equal(in.sel_from.ent, __ldg(in.sel_from.off + i) - sf_base,
                           end_at(in.sel_from, i) - sf_base, (fl & NBG_F_FROM_SEL_NIL) != 0,
                           in.sel_to.ent, __ldg(in.sel_to.off + i) - st_base,
                           end_at(in.sel_to, i) - st_base, (fl & NBG_F_TO_SEL_NIL) != 0))
      reason |= NBG_SVC_SELECTOR;                    // COMMON:184
    if ((port_bits >> lane) & 1u) reason |= NBG_SVC_PORTS;  // COMMON:189
  }
  store_results(base, n, lane, reason, (reason & NBG_SVC_UPDATE_MASK) != 0, reason_out, bitmap_half);
}

__global__ void __launch_bounds__(128)
diff_vs_kernel(nbg_vs_batch in, unsigned long long n, unsigned hf_base, unsigned ht_base,
               uint8_t* __restrict__ reason_out, uint32_t* __restrict__ bitmap_half) {
  const unsigned lane = threadIdx.x & 31;
  const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned long long base = warp * 32ull;
  if (base >= n) return;
  const unsigned long long i = base + lane;

  const unsigned spec_bits = tiles_differ<NBG_VS_SLOTS / 2>(in.spec_from, in.spec_to, base, n, lane);

  unsigned reason = 0;
  bool update = false;
  if (i < n) {
    const unsigned fl = __ldg(in.flags + i);
    if (fl & NBG_F_FROM_SPEC_MISS) {                 // COMMON:200-206: return false
      reason = NBG_VS_FROM_MISSING;
    } else if (fl & NBG_F_TO_SPEC_MISS) {            // COMMON:208-212: copy, return true
      reason = NBG_VS_TO_MISSING; update = true;
    } else {                                         // COMMON:214
      bool differs = (spec_bits >> lane) & 1u;
      differs |= map_not_deep_equal(in.hdr_from.ent, __ldg(in.hdr_from.off + i) - hf_base,
                                    end_at(in.hdr_from, i) - hf_base,
                                    (fl & NBG_F_FROM_HDR_NIL) != 0,
                                    in.hdr_to.ent, __ldg(in.hdr_to.off + i) - ht_base,
                                    end_at(in.hdr_to, i) - ht_base,
                                    (fl & NBG_F_TO_HDR_NIL) != 0);
      if (differs) { reason = NBG_VS_SPEC_DIFFERS; update = true; }
    }
  }
  store_results(base, n, lane, reason, update, reason_out, bitmap_half);
}

// ------------------------------- desired Service / VirtualService made on the device
// Same votes as tiles_differ, but the desired vector of every lane is synthesised from a few
// per-notebook words and constants instead of being loaded (generateService NBCTL:525-552,
// generateVirtualService NBCTL:558-658; slot layout = csrc/host/pack.cpp).
struct SvcGenConsts { unsigned long long port_name, proto_tcp, sel_key; };

#ifndef NBG_GEN_MINBLOCKS
#define NBG_GEN_MINBLOCKS 16
#endif
template <bool SPANS>
__global__ void __launch_bounds__(128, NBG_GEN_MINBLOCKS)
diff_svc_gen_kernel(nbg_svc_gen_batch in, SvcGenConsts k, unsigned long long n,
                    unsigned lt_base, unsigned at_base, unsigned st_base,
                    uint8_t* __restrict__ reason_out, uint32_t* __restrict__ bitmap_half) {
  constexpr int VEC = NBG_SVC_SLOTS / 2;               // 4 lanes per object
  static_assert(VEC == 4, "the desired tile below is written for the 8-slot layout of csrc/host/pack.cpp");
  const unsigned lane = threadIdx.x & 31;
  const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned long long base = warp * 32ull;
  if (base >= n) return;
  const unsigned long long i = base + lane;
  const uint4* pt = reinterpret_cast<const uint4*>(in.ports_to) + base * VEC;
  const unsigned long long limit = (n - base) * VEC;
  unsigned port_bits = 0;
#pragma unroll
  for (int r = 0; r < VEC; ++r) {
    const unsigned idx = r * 32 + lane, part = lane & 3u;
    const unsigned long long obj = base + (idx >> 2);
    unsigned d = 0;
    if (idx < limit) {
      unsigned long long a = 0, b = 0;                 // desired slots 2*part, 2*part+1
      if (part == 0) { a = k.port_name; b = k.proto_tcp; }                    // Name, Protocol
      else if (part == 1) { b = 80ull; }                                      // AppProtocol nil; Port 80, NodePort 0
      else if (part == 2) { a = (unsigned long long)(unsigned)__ldg(in.target_port + obj); }   // TargetPort Int
      else if (part == 3) { a = 1ull; }                                       // one port, nothing after it
      const uint4 want = make_uint4((unsigned)a, (unsigned)(a >> 32), (unsigned)b, (unsigned)(b >> 32));
      d = diff4(want, ld_stream(pt + idx));
    }
    const unsigned ballot = __ballot_sync(FULL, d != 0);
#pragma unroll
    for (int g = 0; g < 8; ++g)
      if ((ballot >> (g * 4)) & 0xFu) port_bits |= 1u << (r * 8 + g);
  }
  unsigned reason = 0;
  if (i < n) {
    const unsigned fl = __ldg(in.flags + i);
    const unsigned l0 = __ldg(in.labels_to.off + i) - lt_base, l1 = end_at<SPANS>(in.labels_to, i) - lt_base;
    const unsigned a0 = __ldg(in.ann_to.off + i) - at_base, a1 = end_at<SPANS>(in.ann_to, i) - at_base;
    const unsigned s0 = __ldg(in.sel_to.off + i) - st_base, s1 = end_at<SPANS>(in.sel_to, i) - st_base;
    // desired has no labels and no annotations: any live entry with a non-empty value differs
    for (unsigned t = l0; t < l1; ++t)
      if (ld_kv(in.labels_to.ent + t).val != 0ull) reason |= NBG_SVC_LABELS;
    for (unsigned t = a0; t < a1; ++t)
      if (ld_kv(in.ann_to.ent + t).val != 0ull) reason |= NBG_SVC_ANNOTATIONS;
    // desired selector = {"statefulset": name}, non-nil
    bool sel_ne = (fl & NBG_F_TO_SEL_NIL) != 0 || (s1 - s0) != 1u;
    if (!sel_ne) {
      const nbg_kv e = ld_kv(in.sel_to.ent + s0);
      sel_ne = e.key != k.sel_key || e.val != __ldg(reinterpret_cast<const unsigned long long*>(in.name_hash) + i);
    }
    if (sel_ne) reason |= NBG_SVC_SELECTOR;
    if ((port_bits >> lane) & 1u) reason |= NBG_SVC_PORTS;
  }
  store_results(base, n, lane, reason, (reason & NBG_SVC_UPDATE_MASK) != 0, reason_out, bitmap_half);
}

template <bool SPANS>
__global__ void __launch_bounds__(128, NBG_GEN_MINBLOCKS)
diff_vs_gen_kernel(nbg_vs_gen_batch in, nbg_vs_gen_consts k, unsigned long long n, unsigned hf_base, unsigned ht_base,
                   uint8_t* __restrict__ reason_out, uint32_t* __restrict__ bitmap_half) {
  constexpr int VEC = NBG_VS_SLOTS / 2;                // 4 lanes per object
  const unsigned lane = threadIdx.x & 31;
  const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned long long base = warp * 32ull;
  if (base >= n) return;
  const unsigned long long i = base + lane;
  const uint4* pt = reinterpret_cast<const uint4*>(in.spec_to) + base * VEC;
  const unsigned long long limit = (n - base) * VEC;
  const unsigned long long* prefix = reinterpret_cast<const unsigned long long*>(in.prefix_hash);
  const unsigned long long* rewrite = reinterpret_cast<const unsigned long long*>(in.rewrite_hash);
  const unsigned long long* host = reinterpret_cast<const unsigned long long*>(in.host_hash);
  unsigned spec_bits = 0;
#pragma unroll
  for (int r = 0; r < VEC; ++r) {
    const unsigned idx = r * 32 + lane, part = lane & 3u;
    const unsigned long long obj = base + (idx >> 2);
    unsigned d = 0;
    if (idx < limit) {
      unsigned long long a, b;
      if (part == 0) { a = k.istio_host_hash; b = k.gateway_hash; }            // hosts[0], gateways[0]
      else if (part == 1) { a = __ldg(prefix + obj); b = __ldg(rewrite + obj); } // match prefix, rewrite uri
      else if (part == 2) { a = __ldg(host + obj); b = 80ull; }                // destination host, port number
      else { a = 1ull; b = k.residual; }                                       // number is an int64; shape digest
      const uint4 want = make_uint4((unsigned)a, (unsigned)(a >> 32), (unsigned)b, (unsigned)(b >> 32));
      d = diff4(want, ld_stream(pt + idx));
    }
    const unsigned ballot = __ballot_sync(FULL, d != 0);
#pragma unroll
    for (int g = 0; g < 8; ++g)
      if ((ballot >> (g * 4)) & 0xFu) spec_bits |= 1u << (r * 8 + g);
  }
  unsigned reason = 0;
  bool update = false;
  if (i < n) {
    const unsigned fl = __ldg(in.flags + i);
    if (fl & NBG_F_TO_SPEC_MISS) {                     // from.spec always exists here
      reason = NBG_VS_TO_MISSING; update = true;
    } else {
      bool differs = (spec_bits >> lane) & 1u;
      differs |= map_not_deep_equal(in.hdr_from.ent, __ldg(in.hdr_from.off + i) - hf_base,
                                    end_at<SPANS>(in.hdr_from, i) - hf_base, false,
                                    in.hdr_to.ent, __ldg(in.hdr_to.off + i) - ht_base,
                                    end_at<SPANS>(in.hdr_to, i) - ht_base, (fl & NBG_F_TO_HDR_NIL) != 0);
      if (differs) { reason = NBG_VS_SPEC_DIFFERS; update = true; }
    }
  }
  store_results(base, n, lane, reason, update, reason_out, bitmap_half);
}

// ------------------------------------------------ ODH comparator family
// One object pair per lane; both tests are short loops over small CSR vectors.
__global__ void __launch_bounds__(128)
diff_generic_kernel(nbg_generic_batch in, unsigned long long n, int mode,
                    unsigned la_base, unsigned lb_base, unsigned sa_base, unsigned sb_base,
                    uint8_t* __restrict__ reason_out, uint32_t* __restrict__ bitmap_half) {
  const unsigned lane = threadIdx.x & 31;
  const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned long long base = warp * 32ull;
  if (base >= n) return;
  const unsigned long long i = base + lane;
  unsigned reason = 0;
  if (i < n) {
    const unsigned la0 = __ldg(in.labels_a.off + i) - la_base, la1 = end_at(in.labels_a, i) - la_base;
    const unsigned lb0 = __ldg(in.labels_b.off + i) - lb_base, lb1 = end_at(in.labels_b, i) - lb_base;
    const unsigned sa0 = __ldg(in.spec_a.off + i) - sa_base, sa1 = end_at(in.spec_a, i) - sa_base;
    const unsigned sb0 = __ldg(in.spec_b.off + i) - sb_base, sb1 = end_at(in.spec_b, i) - sb_base;
    if (mode == NBG_GEN_DEEPEQUAL) {
      // flattened trees: nil-ness, lengths and kinds are entries of the vectors themselves
      if (map_not_deep_equal(in.labels_a.ent, la0, la1, false, in.labels_b.ent, lb0, lb1, false)) reason |= NBG_GEN_LABELS;
      if (map_not_deep_equal(in.spec_a.ent, sa0, sa1, false, in.spec_b.ent, sb0, sb1, false)) reason |= NBG_GEN_SPEC;
    } else {
      // len(found) != len(desired), or a desired entry whose found value differs ("" when missing)
      if ((la1 - la0) != (lb1 - lb0) || live_map_differs(in.labels_b.ent, lb0, lb1, in.labels_a.ent, la0, la1)) reason |= NBG_GEN_LABELS;
      if ((sa1 - sa0) != (sb1 - sb0) || live_map_differs(in.spec_b.ent, sb0, sb1, in.spec_a.ent, sa0, sa1)) reason |= NBG_GEN_SPEC;
    }
  }
  store_results(base, n, lane, reason, reason != 0, reason_out, bitmap_half);
}

}  // namespace nbg

extern "C" cudaError_t nbg_launch_diff_generic(const nbg_generic_batch* in, unsigned long long n, int mode,
                                               const unsigned* bases /*4*/, uint8_t* reason, uint64_t* bitmap,
                                               cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const unsigned block = 128;
  nbg::diff_generic_kernel<<<(unsigned)((n + block - 1) / block), block, 0, stream>>>(
      *in, n, mode, bases[0], bases[1], bases[2], bases[3], reason, reinterpret_cast<uint32_t*>(bitmap));
  return cudaGetLastError();
}

// *_base: first CSR offset of the slice when the entry pointers already point
// at the slice's first entry (host-buffer chunking); 0 for whole arrays.
extern "C" cudaError_t nbg_launch_diff_sts(const nbg_sts_batch* in, unsigned long long n,
                                           const unsigned* bases /*4*/, int deployment, uint8_t* reason,
                                           uint64_t* bitmap, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const unsigned block = 128;                          // 4 warps = 128 objects per CTA
  const unsigned long long grid = (n + block - 1) / block;
  nbg::diff_sts_kernel<<<(unsigned)grid, block, 0, stream>>>(
      *in, n, bases[0], bases[1], bases[2], bases[3], deployment ? 1u : 0u, reason, reinterpret_cast<uint32_t*>(bitmap));
  return cudaGetLastError();
}

extern "C" cudaError_t nbg_launch_diff_sts_gen(const nbg_sts_batch* in, const uint64_t* prefix_hash, const uint8_t* gen_flags,
                                               const nbg_sts_gen_consts* consts, unsigned long long n,
                                               const unsigned* bases /*3: labels_from, labels_to, ann_to*/, uint8_t* reason,
                                               uint64_t* bitmap, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  nbg::StsGenConsts k;
  k.working_dir = consts->working_dir; k.port_name = consts->port_name; k.port_rest = consts->port_rest;
  k.env_name = consts->env_name; k.add_fsgroup = consts->add_fsgroup;
  const unsigned block = 128;
  nbg::diff_sts_gen_kernel<<<(unsigned)((n + block - 1) / block), block, 0, stream>>>(
      *in, reinterpret_cast<const unsigned long long*>(prefix_hash), gen_flags, k, n, bases[0], bases[1], bases[2], reason,
      reinterpret_cast<uint32_t*>(bitmap));
  return cudaGetLastError();
}

extern "C" cudaError_t nbg_launch_diff_sts_sparse(const nbg_sts_sparse_batch* in, unsigned long long n,
                                                  const unsigned* bases /*6*/, uint8_t* reason,
                                                  uint64_t* bitmap, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const unsigned block = 128;
  const unsigned long long grid = (n + block - 1) / block;
  const bool spans = in->pod_from.end || in->pod_to.end || in->labels_from.end || in->labels_to.end || in->ann_from.end || in->ann_to.end;
#ifdef NBG_STS_BULK
  if (!spans) {
    const size_t smem = (size_t)nbg::kBulkWarps * 2 * nbg::kBulkVecs * 16 + nbg::kBulkWarps * 8;
    static bool attr_set = false;
    if (!attr_set) {
      cudaFuncSetAttribute(nbg::diff_sts_sparse_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      attr_set = true;
    }
    const unsigned bblock = 32 * nbg::kBulkWarps;
    nbg::diff_sts_sparse_bulk_kernel<<<(unsigned)((n + bblock - 1) / bblock), bblock, smem, stream>>>(
        *in, n, bases[0], bases[1], bases[2], bases[3], bases[4], bases[5], reason, reinterpret_cast<uint32_t*>(bitmap));
    return cudaGetLastError();
  }
#endif
  (spans ? nbg::diff_sts_sparse_kernel<true> : nbg::diff_sts_sparse_kernel<false>)<<<(unsigned)grid, block, 0, stream>>>(
      *in, n, bases[0], bases[1], bases[2], bases[3], bases[4], bases[5], reason,
      reinterpret_cast<uint32_t*>(bitmap));
  return cudaGetLastError();
}

extern "C" cudaError_t nbg_launch_diff_svc_gen(const nbg_svc_gen_batch* in, unsigned long long n, const unsigned* bases /*3*/,
                                               uint8_t* reason, uint64_t* bitmap, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  nbg::SvcGenConsts k;
  k.port_name = nbg_hash_bytes("http-notebook", 13);   // NBCTL:541
  k.proto_tcp = nbg_hash_bytes("TCP", 3);              // NBCTL:544
  k.sel_key = nbg_hash_bytes("statefulset", 11);       // NBCTL:537
  const unsigned block = 128;
  const bool spans = in->labels_to.end || in->ann_to.end || in->sel_to.end;
  (spans ? nbg::diff_svc_gen_kernel<true> : nbg::diff_svc_gen_kernel<false>)<<<(unsigned)((n + block - 1) / block), block, 0, stream>>>(
      *in, k, n, bases[0], bases[1], bases[2], reason, reinterpret_cast<uint32_t*>(bitmap));
  return cudaGetLastError();
}
extern "C" cudaError_t nbg_launch_diff_vs_gen(const nbg_vs_gen_batch* in, const nbg_vs_gen_consts* consts, unsigned long long n,
                                              const unsigned* bases /*2*/, uint8_t* reason, uint64_t* bitmap, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const unsigned block = 128;
  const bool spans = in->hdr_from.end || in->hdr_to.end;
  (spans ? nbg::diff_vs_gen_kernel<true> : nbg::diff_vs_gen_kernel<false>)<<<(unsigned)((n + block - 1) / block), block, 0, stream>>>(
      *in, *consts, n, bases[0], bases[1], reason, reinterpret_cast<uint32_t*>(bitmap));
  return cudaGetLastError();
}

extern "C" cudaError_t nbg_launch_restart_block(const nbg_restart_batch* in, unsigned long long n, const unsigned* bases /*3*/,
                                                uint8_t* decision, uint64_t* bitmap, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const unsigned block = 128;
  const bool spans = in->old_spec.end || in->updated_spec.end || in->mutated_spec.end;
  (spans ? nbg::restart_block_kernel<true> : nbg::restart_block_kernel<false>)<<<(unsigned)((n + block - 1) / block), block, 0, stream>>>(
      *in, n, bases[0], bases[1], bases[2], decision, reinterpret_cast<uint32_t*>(bitmap));
  return cudaGetLastError();
}

extern "C" cudaError_t nbg_launch_diff_svc(const nbg_svc_batch* in, unsigned long long n,
                                           const unsigned* bases /*6*/, uint8_t* reason,
                                           uint64_t* bitmap, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const unsigned block = 128;
  const unsigned long long grid = (n + block - 1) / block;
  nbg::diff_svc_kernel<<<(unsigned)grid, block, 0, stream>>>(
      *in, n, bases[0], bases[1], bases[2], bases[3], bases[4], bases[5], reason,
      reinterpret_cast<uint32_t*>(bitmap));
  return cudaGetLastError();
}

extern "C" cudaError_t nbg_launch_diff_vs(const nbg_vs_batch* in, unsigned long long n,
                                          const unsigned* bases /*2*/, uint8_t* reason,
                                          uint64_t* bitmap, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const unsigned block = 128;
  const unsigned long long grid = (n + block - 1) / block;
  nbg::diff_vs_kernel<<<(unsigned)grid, block, 0, stream>>>(
      *in, n, bases[0], bases[1], reason, reinterpret_cast<uint32_t*>(bitmap));
  return cudaGetLastError();
}
