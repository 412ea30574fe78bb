// cull.cu — the culler's idleness decision for a whole fleet in one launch.
//
// One thread owns one notebook; a warp owns 32 consecutive notebooks, so the
// kernel/terminal records it walks are one contiguous run of the record arrays
// (CSR order) and every DRAM sector fetched is consumed by the warp that asked
// for it.  Per-notebook words and outputs are coalesced across the warp; the
// two result bitmaps are assembled with __ballot_sync, one 32-bit half-word
// per warp.
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

struct Scan {
  bool all_ok;    // every last_activity parsed
  bool all_idle;  // every execution_state == "idle"
  Instant best;   // maximum instant, first occurrence
  unsigned idx;   // its index within the segment
};

__device__ __forceinline__ Scan scan_records(const long long* __restrict__ sec,
                                             const unsigned* __restrict__ nsf,
                                             unsigned lo, unsigned hi) {
  Scan s;
  s.all_ok = true; s.all_idle = true;
  s.best.sec = 0; s.best.nsec = 0; s.idx = 0;
  for (unsigned r = lo; r < hi; ++r) {
    const unsigned w = __ldg(nsf + r);
    Instant t;
    t.sec = __ldg(sec + r);
    t.nsec = w & NBG_NSEC_MASK;
    s.all_ok &= (w & NBG_REC_PARSE_OK) != 0;
    s.all_idle &= (w & NBG_REC_IDLE) != 0;
    if (r == lo || after(t, s.best)) { s.best = t; s.idx = r - lo; }
  }
  return s;
}

__global__ void __launch_bounds__(256)
cull_kernel(nbg_cull_batch in, unsigned long long n, nbg_cull_params p,
            unsigned k_base, unsigned t_base,
            uint8_t* __restrict__ action, uint32_t* __restrict__ argmax,
            uint32_t* __restrict__ cull_half, uint32_t* __restrict__ period_half) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < n;
  unsigned act = 0, amax = NBG_ARGMAX_NONE;

  if (live) {
    const unsigned annw = __ldg(in.ann_nsf + i);
    const unsigned chkw = __ldg(in.chk_nsf + i);
    const unsigned nbf = __ldg(in.nb_flags + i);
    Instant ann; ann.sec = __ldg(in.ann_sec + i); ann.nsec = annw & NBG_NSEC_MASK;
    Instant chk; chk.sec = __ldg(in.chk_sec + i); chk.nsec = chkw & NBG_NSEC_MASK;
    bool ann_ok = (annw & NBG_ANN_PARSE_OK) != 0;
    const bool stop = (annw & NBG_ANN_STOP_SET) != 0;
    bool chk_present = (chkw & NBG_CHK_PRESENT) != 0;
    bool chk_ok = (chkw & NBG_CHK_PARSE_OK) != 0;
    const Instant now = {p.now_sec, (unsigned)p.now_nsec};
    const Instant now_trunc = {p.now_sec, 0u};  // createTimestamp(): Format(RFC3339), CULL:512-515
    bool proceed = true;

    if (p.mode == NBG_CULL_RECONCILE) {
      if (stop) {                                   // CULL:104-117
        act = NBG_ACT_STOPPED; proceed = false;
      } else if (nbf & NBG_NB_POD_MISSING) {        // CULL:121-135
        act = NBG_ACT_NOPOD; proceed = false;
      } else if (!((nbf & NBG_NB_LA_PRESENT) && chk_present)) {  // CULL:141-153, :440-456
        act |= NBG_ACT_INIT;
        ann = now_trunc; ann_ok = true;
        chk = now_trunc; chk_ok = true; chk_present = true;
      }
    }

    if (proceed) {
      // C8 — CULL:207-217
      bool period = false;
      if (chk_present) {
        Instant next;
        next.sec = (chk_ok ? chk.sec : kZeroTimeUnix) + (long long)p.check_period_min * 60ll;
        next.nsec = chk_ok ? chk.nsec : 0u;
        period = after(now, next);                  // next.Before(now)
      }
      if (period) act |= NBG_ACT_PERIOD;

      if (p.mode == NBG_CULL_PREDICATES || period) {
        unsigned src = 0;
        // C4 — kernels
        const unsigned k_lo = __ldg(in.k_off + i) - k_base, k_hi = __ldg(in.k_off + i + 1) - k_base;
        if (k_hi > k_lo) {
          const Scan s = scan_records((const long long*)in.k_sec, in.k_nsf, k_lo, k_hi);
          if (!s.all_idle) {                        // CULL:378-387
            ann = now_trunc; ann_ok = true; src = 1;
          } else if (s.all_ok) {                    // CULL:389-399
            const Instant t = {s.best.sec, 0u};
            if (ann_ok && !after(ann, t)) { ann = t; src = 2; amax = s.idx; }
          }
        }
        // C5 — terminals, on the annotation C4 may just have written
        const unsigned t_lo = __ldg(in.t_off + i) - t_base, t_hi = __ldg(in.t_off + i + 1) - t_base;
        if (t_hi > t_lo) {
          const Scan s = scan_records((const long long*)in.t_sec, in.t_nsf, t_lo, t_hi);
          if (s.all_ok) {
            const Instant t = {s.best.sec, 0u};
            if (ann_ok && !after(ann, t)) { ann = t; src = 3; amax = s.idx; }
          }
        }
        act |= src;
        // C7 — CULL:220-241, on the annotation after C4/C5
        if (!stop && ann_ok) {
          Instant cap;
          cap.sec = ann.sec + (long long)p.cull_idle_min * 60ll;
          cap.nsec = ann.nsec;
          if (after(now, cap)) act |= NBG_ACT_CULL;
        }
      }
    }
    action[i] = (uint8_t)act;
    argmax[i] = amax;
  }

  const unsigned cull_bits = __ballot_sync(FULL, (act & NBG_ACT_CULL) != 0);
  const unsigned period_bits = __ballot_sync(FULL, (act & NBG_ACT_PERIOD) != 0);
  if ((threadIdx.x & 31) == 0 && live) {
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

}  // namespace nbg

// k_base / t_base: value of k_off[0] / t_off[0] when the record pointers in
// `in` already point at the first record of this slice (host-buffer chunking);
// 0 when offsets index the arrays directly.
extern "C" cudaError_t nbg_launch_cull(const nbg_cull_batch* in, unsigned long long n,
                                       const nbg_cull_params* p, unsigned k_base, unsigned t_base,
                                       uint8_t* action, uint32_t* argmax, uint64_t* cull_bitmap,
                                       uint64_t* period_bitmap, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const unsigned block = 256;
  const unsigned long long grid = (n + block - 1) / block;
  nbg::cull_kernel<<<(unsigned)grid, block, 0, stream>>>(
      *in, n, *p, k_base, t_base, action, argmax, reinterpret_cast<uint32_t*>(cull_bitmap),
      reinterpret_cast<uint32_t*>(period_bitmap));
  return cudaGetLastError();
}
