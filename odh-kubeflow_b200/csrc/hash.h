/*
 * hash.h — the one 64-bit string hash every packer of this repo uses, and the
 * mixers built on it.  Shared by libnbgraft.so (nbg_hash64, nbg_shard_of) and
 * libnbgraft_host.so (the object packer).  Plain C99/C++; no dependencies.
 *
 * Contract (include/nbgraft.h): hash("") == 0 and no non-empty string hashes
 * to 0, so that Go's `from.Labels[k] != v` with a missing key (which reads as
 * "", COMMON:110) is a plain integer compare on the device.
 */
#ifndef NBG_HASH_H
#define NBG_HASH_H
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifdef __CUDACC__
#define NBG_HD __host__ __device__
#else
#define NBG_HD
#endif

NBG_HD static inline uint64_t nbg_mix64(uint64_t x) { /* splitmix64 finaliser */
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return x;
}

static inline uint64_t nbg_rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }

/* Order-dependent combine of two 64-bit words (digests of sequences/trees). */
static inline uint64_t nbg_combine(uint64_t h, uint64_t v) {
  return nbg_mix64(nbg_rotl64(h, 23) ^ (v * 0x9E3779B97F4A7C15ull) ^ 0xD6E8FEB86659FD93ull);
}

static inline uint64_t nbg_hash_bytes(const void* data, size_t len) {
  if (len == 0) return 0;
  const unsigned char* p = (const unsigned char*)data;
  uint64_t h = 0x6E62677261667431ull ^ ((uint64_t)len * 0x9E3779B97F4A7C15ull);
  size_t i = 0;
  for (; i + 8 <= len; i += 8) {
    uint64_t w; memcpy(&w, p + i, 8);
    h = nbg_mix64(h ^ w) + 0x2545F4914F6CDD1Dull;
  }
  if (i < len) {
    uint64_t w = 0; memcpy(&w, p + i, len - i);
    h = nbg_mix64(h ^ w ^ ((uint64_t)(len - i) << 56));
  }
  h = nbg_mix64(h);
  return h ? h : 1;
}

static inline uint32_t nbg_shard_of_inline(uint64_t ns_hash, uint32_t n_shards) {
  if (n_shards == 0) return 0;
  return (uint32_t)(nbg_mix64(ns_hash ^ 0x73686172646F6621ull) % n_shards);
}

#ifdef __cplusplus
}
#endif
#endif
