// cull.cu — the culler's idleness decision for a whole fleet in one launch.
//
// A warp owns 32 consecutive notebooks, so the kernel/terminal records it walks
// are one contiguous run of the record arrays (CSR order) and every DRAM sector
// fetched is consumed by the warp that asked for it.  Gates, scans and decisions
// are scalar work, one notebook per lane.  Per-notebook words and outputs are
// coalesced across the warp; the two result bitmaps are assembled with
// __ballot_sync, one 32-bit half-word per warp.
//
// Reference semantics, all in components/notebook-controller/controllers/
// culling_controller.go (CULL):
//   C1 allKernelsAreIdle                    CULL:315-327
//   C2 getNotebookRecentTime                CULL:329-349 (strict After => first max wins;
//                                                      Format(RFC3339) drops the fraction)
//   C3 compareAnnotationTimeToResource      CULL:351-369 (annotation.After(t) => no update)
//   C4 updateTimestampFromKernelsActivity   CULL:371-402
//   C5 updateTimestampFromTerminalsActivity CULL:404-428
//   C7 notebookIsIdle                       CULL:220-241 (strict now.After(last+idle))
//   C8 cullingCheckPeriodHasPassed          CULL:207-217 (parse error ignored => zero Time)
//   flow of CullingReconciler.Reconcile     CULL:86-203  (mode NBG_CULL_RECONCILE)
#include "common.cuh"

namespace nbg {

struct Instant {
  long long sec;
  unsigned nsec;
};

__device__ __forceinline__ bool after(const Instant& a, const Instant& b) {  // Go: a.After(b)
  return a.sec > b.sec || (a.sec == b.sec && a.nsec > b.nsec);
}

// Go's zero time.Time (0001-01-01T00:00:00Z) as a Unix second count: what
// time.Parse returns alongside an error (CULL:212 ignores that error).
constexpr long long kZeroTimeUnix = -62135596800ll;

// Partial scan result over some records of one notebook's list.
struct Part {
  long long sec;
  unsigned nsec;
  unsigned idx;  // index within the segment of the best record; kNone = no record seen
  unsigned acc;  // AND of the records' nsf words: bit 30 = all parsed, bit 31 = all idle
};
constexpr unsigned kNone = 0xFFFFFFFFu;

__device__ __forceinline__ void take(Part& p, long long sec, unsigned w, unsigned idx) {
  p.acc &= w;
  const Instant t = {sec, w & NBG_NSEC_MASK}, cur = {p.sec, p.nsec};
  if (p.idx == kNone || after(t, cur)) { p.sec = t.sec; p.nsec = t.nsec; p.idx = idx; }
}

__device__ __forceinline__ Part scan_serial(const long long* __restrict__ sec, const unsigned* __restrict__ nsf,
                                            unsigned lo, unsigned hi) {
  Part p = {0, 0u, kNone, ~0u};
  for (unsigned r = lo; r < hi; ++r) take(p, __ldg(sec + r), __ldg(nsf + r), r - lo);
  return p;
}

#ifndef NBG_CULL_BLOCK
#define NBG_CULL_BLOCK 256
#endif
#ifndef NBG_CULL_MINBLOCKS
#define NBG_CULL_MINBLOCKS 8   // 32 registers/thread: 64 resident warps per SM
#endif

template <bool SPANS>
__global__ void __launch_bounds__(NBG_CULL_BLOCK, NBG_CULL_MINBLOCKS)
cull_kernel(nbg_cull_batch in, unsigned long long n, nbg_cull_params p,
            unsigned k_base, unsigned t_base,
            uint8_t* __restrict__ action, uint32_t* __restrict__ argmax,
            uint32_t* __restrict__ cull_half, uint32_t* __restrict__ period_half) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned lane = threadIdx.x & 31u;
  if (i - lane >= n) return;                             // whole warp past the end
  const bool live = i < n;
  unsigned act = 0, amax = NBG_ARGMAX_NONE;

  // ---- per-notebook gates (scalar work, one notebook per lane)
  Instant ann = {0, 0u}, chk = {0, 0u};
  bool ann_ok = false, stop = false, chk_present = false, chk_ok = false, evaluate = false;
  const Instant now = {p.now_sec, (unsigned)p.now_nsec};
  const Instant now_trunc = {p.now_sec, 0u};             // createTimestamp(): Format(RFC3339), CULL:512-515
  unsigned k_lo = 0, k_hi = 0, t_lo = 0, t_hi = 0;
  if (live) {
    // every per-notebook word, offsets included, is requested up front so the
    // loads overlap; the gates below only decide whether the records are read
    const unsigned annw = __ldg(in.ann_nsf + i);
    const unsigned chkw = __ldg(in.chk_nsf + i);
    const unsigned nbf = __ldg(in.nb_flags + i);
    const unsigned k0 = __ldg(in.k_off + i), k1 = SPANS ? __ldg((in.k_end ? in.k_end : in.k_off + 1) + i) : __ldg(in.k_off + i + 1);
    const unsigned t0 = __ldg(in.t_off + i), t1 = SPANS ? __ldg((in.t_end ? in.t_end : in.t_off + 1) + i) : __ldg(in.t_off + i + 1);
    ann.sec = __ldg(in.ann_sec + i); ann.nsec = annw & NBG_NSEC_MASK;
    chk.sec = __ldg(in.chk_sec + i); chk.nsec = chkw & NBG_NSEC_MASK;
    ann_ok = (annw & NBG_ANN_PARSE_OK) != 0;
    stop = (annw & NBG_ANN_STOP_SET) != 0;
    chk_present = (chkw & NBG_CHK_PRESENT) != 0;
    chk_ok = (chkw & NBG_CHK_PARSE_OK) != 0;
    bool proceed = true;
    if (p.mode == NBG_CULL_RECONCILE) {
      if (stop) {                                        // CULL:104-117
        act = NBG_ACT_STOPPED; proceed = false;
      } else if (nbf & NBG_NB_POD_MISSING) {             // CULL:121-135
        act = NBG_ACT_NOPOD; proceed = false;
      } else if (!((nbf & NBG_NB_LA_PRESENT) && chk_present)) {  // CULL:141-153, :440-456
        act |= NBG_ACT_INIT;
        ann = now_trunc; ann_ok = true;
        chk = now_trunc; chk_ok = true; chk_present = true;
      }
    }
    if (proceed) {
      bool period = false;                               // C8 — CULL:207-217
      if (chk_present) {
        Instant next;
        next.sec = (chk_ok ? chk.sec : kZeroTimeUnix) + (long long)p.check_period_min * 60ll;
        next.nsec = chk_ok ? chk.nsec : 0u;
        period = after(now, next);                       // next.Before(now)
      }
      if (period) act |= NBG_ACT_PERIOD;
      evaluate = (p.mode == NBG_CULL_PREDICATES) || period;
    }
    if (evaluate) {                                      // gated notebooks never touch their records
      k_lo = k0 - k_base; k_hi = k1 - k_base;
      t_lo = t0 - t_base; t_hi = t1 - t_base;
    }
  }

  // ---- record scans, one notebook per lane.  (Tried and measured slower on B200 at 1.25 M
  // notebooks: four-lane groups with butterfly merges, 85 us; staging each warp's record run in
  // shared memory first, 54-65 us; the whole warp in lockstep over the longest list with predicated
  // loads and selects, 16 instructions per slot and no branch, 45-49 us; the same with kernel and
  // terminal records in one loop, 56-66 us; this plain loop with every per-notebook load hoisted, 43 us
  // (unchanged by 64- or 128-thread blocks and by prefetching the terminal records' lines before the kernel scan;
  // a branch-free take() — selects instead of the predicated update — 258 us against 247 us at 10 M notebooks).
  // ncu source counters: 404 of the 687 warp instructions per 32 notebooks are these two loops, but the
  // warps wait on three dependent memory round trips (offsets -> kernel records -> terminal records)
  // whatever the loops cost: DESIGN.md section 3.)
  const Part ks = scan_serial((const long long*)in.k_sec, in.k_nsf, k_lo, k_hi);
  const Part ts = scan_serial((const long long*)in.t_sec, in.t_nsf, t_lo, t_hi);

  if (live) {
    if (evaluate) {
      unsigned src = 0;
      if (ks.idx != kNone) {                             // C4 — len(kernels) > 0
        if (!(ks.acc & NBG_REC_IDLE)) {                  // CULL:378-387
          ann = now_trunc; ann_ok = true; src = 1;
        } else if (ks.acc & NBG_REC_PARSE_OK) {          // CULL:389-399
          const Instant t = {ks.sec, 0u};
          if (ann_ok && !after(ann, t)) { ann = t; src = 2; amax = ks.idx; }
        }
      }
      if (ts.idx != kNone && (ts.acc & NBG_REC_PARSE_OK)) {  // C5, on the annotation C4 may have written
        const Instant t = {ts.sec, 0u};
        if (ann_ok && !after(ann, t)) { ann = t; src = 3; amax = ts.idx; }
      }
      act |= src;
      if (!stop && ann_ok) {                             // C7 — CULL:220-241, after C4/C5
        Instant cap;
        cap.sec = ann.sec + (long long)p.cull_idle_min * 60ll;
        cap.nsec = ann.nsec;
        if (after(now, cap)) act |= NBG_ACT_CULL;
      }
    }
    action[i] = (uint8_t)act;
    argmax[i] = amax;
  }

  const unsigned cull_bits = __ballot_sync(FULL, (act & NBG_ACT_CULL) != 0);
  const unsigned period_bits = __ballot_sync(FULL, (act & NBG_ACT_PERIOD) != 0);
  if (lane == 0) {
    cull_half[i >> 5] = cull_bits;
    period_half[i >> 5] = period_bits;
    // the bitmaps are arrays of 64-bit words: when n ends inside the low half
    // of the last word, nobody owns its high half, so clear it here.
    if (((i >> 5) & 1) == 0 && i + 32 >= n) {
      cull_half[(i >> 5) + 1] = 0u;
      period_half[(i >> 5) + 1] = 0u;
    }
  }
}

// ---- glue for nbg_cull_text: device-parsed annotation words + presence flags -> cull inputs
__global__ void __launch_bounds__(256)
cull_text_fixup_kernel(const uint8_t* __restrict__ flags_in, unsigned long long n, uint32_t* __restrict__ ann_nsf,
                       uint32_t* __restrict__ chk_nsf, uint8_t* __restrict__ nb_flags) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned f = flags_in[i];
  // a key that is absent cannot have parsed: keep only the words of present keys
  unsigned a = (f & NBG_NB_LA_PRESENT) ? (ann_nsf[i] & (NBG_NSEC_MASK | NBG_ANN_PARSE_OK)) : 0u;
  if (f & NBG_NBT_STOP_SET) a |= NBG_ANN_STOP_SET;
  unsigned c = (f & NBG_NBT_CHK_PRESENT) ? ((chk_nsf[i] & (NBG_NSEC_MASK | NBG_CHK_PARSE_OK)) | NBG_CHK_PRESENT) : 0u;
  ann_nsf[i] = a;
  chk_nsf[i] = c;
  nb_flags[i] = (uint8_t)(f & (NBG_NB_LA_PRESENT | NBG_NB_POD_MISSING));
}

// zone offset of the record that supplies the new last-activity (what Format(RFC3339) needs)
__global__ void __launch_bounds__(256)
cull_text_zone_kernel(const uint8_t* __restrict__ action, const uint32_t* __restrict__ argmax,
                      const uint32_t* __restrict__ k_off, const int32_t* __restrict__ k_zone, unsigned k_base,
                      const uint32_t* __restrict__ t_off, const int32_t* __restrict__ t_zone, unsigned t_base,
                      unsigned long long n, int32_t* __restrict__ out) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned src = action[i] & NBG_ACT_SRC_MASK;
  int z = 0;
  if (src == 2u) z = k_zone[k_off[i] - k_base + argmax[i]];
  else if (src == 3u) z = t_zone[t_off[i] - t_base + argmax[i]];
  out[i] = z;
}

}  // namespace nbg

extern "C" cudaError_t nbg_launch_cull_text_fixup(const uint8_t* flags_in, unsigned long long n, uint32_t* ann_nsf,
                                                  uint32_t* chk_nsf, uint8_t* nb_flags, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  nbg::cull_text_fixup_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(flags_in, n, ann_nsf, chk_nsf, nb_flags);
  return cudaGetLastError();
}
extern "C" cudaError_t nbg_launch_cull_text_zone(const uint8_t* action, const uint32_t* argmax, const uint32_t* k_off,
                                                 const int32_t* k_zone, unsigned k_base, const uint32_t* t_off,
                                                 const int32_t* t_zone, unsigned t_base, unsigned long long n, int32_t* out,
                                                 cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  nbg::cull_text_zone_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(action, argmax, k_off, k_zone, k_base, t_off,
                                                                               t_zone, t_base, n, out);
  return cudaGetLastError();
}

// k_base / t_base: value of k_off[0] / t_off[0] when the record pointers in
// `in` already point at the first record of this slice (host-buffer chunking);
// 0 when offsets index the arrays directly.
extern "C" cudaError_t nbg_launch_cull(const nbg_cull_batch* in, unsigned long long n,
                                       const nbg_cull_params* p, unsigned k_base, unsigned t_base,
                                       uint8_t* action, uint32_t* argmax, uint64_t* cull_bitmap,
                                       uint64_t* period_bitmap, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const unsigned block = NBG_CULL_BLOCK;
  const unsigned long long grid = (n + block - 1) / block;
  ((in->k_end || in->t_end) ? nbg::cull_kernel<true> : nbg::cull_kernel<false>)<<<(unsigned)grid, block, 0, stream>>>(
      *in, n, *p, k_base, t_base, action, argmax, reinterpret_cast<uint32_t*>(cull_bitmap),
      reinterpret_cast<uint32_t*>(period_bitmap));
  return cudaGetLastError();
}
