// text.cu — the two string steps that sit right in front of the predicates,
// moved to the device: on the host they dominate the
// end-to-end cost of a batch (parsing ~6 timestamps and hashing ~20 strings per
// notebook is 100-1000x slower than the compare kernels).
//
//   parse_records_kernel  time.Parse(time.RFC3339, last_activity) for every kernel /
//                         terminal record (CULL:332,339) and execution_state == "idle"
//                         (CULL:320), producing the (sec, nsf) words cull_kernel reads
//   hash_strings_kernel   nbg_hash64 of every string of a CSR string table (label keys
//                         and values, env values, ...), bit-identical to the host hash
//
// Strings arrive as CSR: off u32[R+1] + bytes.  A warp's 32 strings are one
// contiguous byte run; the warp copies that run into shared memory with
// coalesced 16-byte loads and every lane then walks its own string out of
// shared memory (runs longer than the staging buffer are read from global
// memory directly, same code path through the byte accessor).
//
// The RFC3339 rules are those of Go's layout parser for "2006-01-02T15:04:05Z07:00"
// (see csrc/host/rfc3339.hpp for the list and its provenance); the kernel is
// checked byte for byte against both CPU parsers, including on mutated strings.
#include "common.cuh"
#include "../hash.h"

namespace nbg {

constexpr unsigned kStageBytes = 2048;   // per warp: 32 strings of up to 64 bytes on average

struct Bytes {                           // one string, read from shared or global memory
  const unsigned char* p;
  unsigned len;
  __device__ __forceinline__ unsigned at(unsigned k) const { return k < len ? p[k] : 0u; }
  __device__ __forceinline__ bool digit(unsigned k) const { const unsigned c = at(k); return c >= '0' && c <= '9'; }
};

// Stage the warp's byte run [run_lo, run_hi) of `text`; returns this lane's string.
__device__ __forceinline__ Bytes stage_strings(const unsigned char* __restrict__ text, unsigned lo, unsigned hi,
                                               unsigned run_lo, unsigned run_hi, unsigned char* stage, unsigned lane) {
  Bytes b;
  b.len = hi - lo;
  const unsigned run = run_hi - run_lo;
  const unsigned mis = (unsigned)((uintptr_t)(text + run_lo) & 15u);      // align the copy down to 16 bytes
  if (run + mis <= kStageBytes) {
    const uint4* src = reinterpret_cast<const uint4*>(text + run_lo - mis);
    uint4* dst = reinterpret_cast<uint4*>(stage);
    const unsigned vecs = (run + mis + 15u) >> 4;   // may read up to 15 bytes past the run: stays inside the
    for (unsigned v = lane; v < vecs; v += 32u) dst[v] = ld_stream(src + v);   // 16-byte-padded array (see nbgraft.h)
    __syncwarp();
    b.p = stage + mis + (lo - run_lo);
  } else {
    b.p = text + lo;
  }
  return b;
}

__device__ __forceinline__ bool leap(int y) { return (y % 4 == 0) && (y % 100 != 0 || y % 400 == 0); }
__device__ __forceinline__ int days_in(int m, int y) {
  return m == 2 ? (leap(y) ? 29 : 28) : ((m == 4 || m == 6 || m == 9 || m == 11) ? 30 : 31);
}
__device__ __forceinline__ long long days_from_civil(long long y, int m, int d) {
  y -= m <= 2;
  const long long era = (y >= 0 ? y : y - 399) / 400;
  const long long yoe = y - era * 400;
  const long long doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
  const long long doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  return era * 146097 + doe - 719468;
}

// time.Parse(time.RFC3339, s)
__device__ bool parse_rfc3339(const Bytes& s, long long* sec_out, unsigned* nsec_out, int* off_out) {
  unsigned p = 0;
  const unsigned n = s.len;
  if (n < 4 || !s.digit(0) || !s.digit(1) || !s.digit(2) || !s.digit(3)) return false;
  const int year = (int)((s.at(0) - '0') * 1000 + (s.at(1) - '0') * 100 + (s.at(2) - '0') * 10 + (s.at(3) - '0'));
  p = 4;
  auto num2 = [&](int* v) {
    if (!s.digit(p) || !s.digit(p + 1)) return false;
    *v = (int)((s.at(p) - '0') * 10 + (s.at(p + 1) - '0'));
    p += 2;
    return true;
  };
  auto lit = [&](unsigned c) { if (p < n && s.at(p) == c) { ++p; return true; } return false; };
  int month, day, hour, minute, second;
  if (!lit('-') || !num2(&month) || month < 1 || month > 12) return false;
  if (!lit('-') || !num2(&day)) return false;
  if (!lit('T')) return false;
  if (!s.digit(p)) return false;                       // hour: one or two digits
  hour = (int)(s.at(p++) - '0');
  if (s.digit(p)) hour = hour * 10 + (int)(s.at(p++) - '0');
  if (hour > 23) return false;
  if (!lit(':') || !num2(&minute) || minute > 59) return false;
  if (!lit(':') || !num2(&second) || second > 59) return false;
  unsigned nsec = 0;
  if (p + 1 < n && (s.at(p) == '.' || s.at(p) == ',') && s.digit(p + 1)) {
    ++p;
    int k = 0;
    while (s.digit(p)) { if (k < 9) { nsec = nsec * 10u + (s.at(p) - '0'); ++k; } ++p; }
    for (; k < 9; ++k) nsec *= 10u;
  }
  int offset = 0;
  if (p < n && s.at(p) == 'Z') {
    ++p;
  } else {
    if (n - p < 6 || s.at(p + 3) != ':') return false;
    const unsigned sign = s.at(p);
    ++p;
    int hh, mm;
    if (!num2(&hh)) return false;
    ++p;
    if (!num2(&mm)) return false;
    if (hh > 24 || mm > 60) return false;
    offset = (hh * 60 + mm) * 60;
    if (sign == '-') offset = -offset;
    else if (sign != '+') return false;
  }
  if (p != n) return false;
  if (day < 1 || day > days_in(month, year)) return false;
  *sec_out = days_from_civil(year, month, day) * 86400ll + hour * 3600 + minute * 60 + second - offset;
  *nsec_out = nsec;
  *off_out = offset;
  return true;
}

__global__ void __launch_bounds__(128)
parse_records_kernel(const uint32_t* __restrict__ t_off, const unsigned char* __restrict__ text,
                     const uint32_t* __restrict__ s_off, const unsigned char* __restrict__ state,
                     unsigned long long R, unsigned t_base, unsigned s_base,
                     long long* __restrict__ sec, uint32_t* __restrict__ nsf, int32_t* __restrict__ zone) {
  __shared__ __align__(16) unsigned char stage[4][kStageBytes];
  const unsigned lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
  const unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned long long base = r - lane;
  if (base >= R) return;
  const bool live = r < R;
  const unsigned long long last = min(base + 31ull, R - 1ull);
  // ---- timestamp
  const unsigned lo = live ? __ldg(t_off + r) - t_base : 0u, hi = live ? __ldg(t_off + r + 1) - t_base : 0u;
  const unsigned run_lo = __ldg(t_off + base) - t_base, run_hi = __ldg(t_off + last + 1) - t_base;
  const Bytes ts = stage_strings(text, live ? lo : run_lo, live ? hi : run_lo, run_lo, run_hi, stage[w], lane);
  long long sc = 0; unsigned ns = 0; int off = 0;
  unsigned word = 0;
  if (live && parse_rfc3339(ts, &sc, &ns, &off)) word = ns | NBG_REC_PARSE_OK; else { sc = 0; off = 0; }
  // ---- execution_state == "idle" (optional column)
  if (s_off) {
    __syncwarp();
    const unsigned slo = live ? __ldg(s_off + r) - s_base : 0u, shi = live ? __ldg(s_off + r + 1) - s_base : 0u;
    const unsigned srun_lo = __ldg(s_off + base) - s_base, srun_hi = __ldg(s_off + last + 1) - s_base;
    const Bytes st = stage_strings(state, live ? slo : srun_lo, live ? shi : srun_lo, srun_lo, srun_hi, stage[w], lane);
    if (live && st.len == 4 && st.at(0) == 'i' && st.at(1) == 'd' && st.at(2) == 'l' && st.at(3) == 'e') word |= NBG_REC_IDLE;
  }
  if (live) {
    sec[r] = sc;
    nsf[r] = word;
    if (zone) zone[r] = off;
  }
}

// nbg_hash_bytes (csrc/hash.h) on the device: same constants, same word order.
__device__ __forceinline__ unsigned long long hash_bytes_dev(const Bytes& s) {
  if (s.len == 0) return 0ull;
  unsigned long long h = 0x6E62677261667431ull ^ ((unsigned long long)s.len * 0x9E3779B97F4A7C15ull);
  unsigned i = 0;
  for (; i + 8 <= s.len; i += 8) {
    unsigned long long wv = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b) wv |= (unsigned long long)s.p[i + b] << (8 * b);
    h = nbg_mix64(h ^ wv) + 0x2545F4914F6CDD1Dull;
  }
  if (i < s.len) {
    unsigned long long wv = 0;
    for (unsigned b = 0; i + b < s.len; ++b) wv |= (unsigned long long)s.p[i + b] << (8 * b);
    h = nbg_mix64(h ^ wv ^ ((unsigned long long)(s.len - i) << 56));
  }
  h = nbg_mix64(h);
  return h ? h : 1ull;
}

__global__ void __launch_bounds__(128)
hash_strings_kernel(const uint32_t* __restrict__ off, const unsigned char* __restrict__ text, unsigned long long R,
                    unsigned base_off, unsigned long long* __restrict__ out) {
  __shared__ __align__(16) unsigned char stage[4][kStageBytes];
  const unsigned lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
  const unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned long long base = r - lane;
  if (base >= R) return;
  const bool live = r < R;
  const unsigned long long last = min(base + 31ull, R - 1ull);
  const unsigned lo = live ? __ldg(off + r) - base_off : 0u, hi = live ? __ldg(off + r + 1) - base_off : 0u;
  const unsigned run_lo = __ldg(off + base) - base_off, run_hi = __ldg(off + last + 1) - base_off;
  const Bytes s = stage_strings(text, live ? lo : run_lo, live ? hi : run_lo, run_lo, run_hi, stage[w], lane);
  if (live) out[r] = hash_bytes_dev(s);
}

}  // namespace nbg

extern "C" cudaError_t nbg_launch_parse_records(const uint32_t* t_off, const unsigned char* text, const uint32_t* s_off,
                                                const unsigned char* state, unsigned long long R, unsigned t_base,
                                                unsigned s_base, int64_t* sec, uint32_t* nsf, int32_t* zone,
                                                cudaStream_t stream) {
  if (R == 0) return cudaSuccess;
  const unsigned block = 128;
  nbg::parse_records_kernel<<<(unsigned)((R + block - 1) / block), block, 0, stream>>>(
      t_off, text, s_off, state, R, t_base, s_base, reinterpret_cast<long long*>(sec), nsf, zone);
  return cudaGetLastError();
}

extern "C" cudaError_t nbg_launch_hash_strings(const uint32_t* off, const unsigned char* text, unsigned long long R,
                                               unsigned base_off, uint64_t* out, cudaStream_t stream) {
  if (R == 0) return cudaSuccess;
  const unsigned block = 128;
  nbg::hash_strings_kernel<<<(unsigned)((R + block - 1) / block), block, 0, stream>>>(
      off, text, R, base_off, reinterpret_cast<unsigned long long*>(out));
  return cudaGetLastError();
}
