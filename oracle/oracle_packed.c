/*
 * oracle_packed.c — L2 "packed" oracle: the same predicates as the CUDA
 * kernels, as plain C loops over the same structure-of-arrays batches
 * (include/nbgraft.h).  Bit-exact target for the kernels at every size, and a
 * timed CPU line in bench.py.
 *
 * TEST INFRASTRUCTURE ONLY: nothing under odh-kubeflow_b200/ may include, link
 * or call this file.
 *
 * It follows the reference lines the kernels follow:
 *   components/common/reconcilehelper/util.go:107-134 (StatefulSet), :166-195
 *   (Service), :199-219 (VirtualService);
 *   components/notebook-controller/controllers/notebook_controller.go:188-194;
 *   components/notebook-controller/controllers/culling_controller.go:86-203,
 *   :207-241, :315-428.
 * Parity of THIS file with the reference rests on oracle_semantic.cpp (L1): the
 * tests check L2 == L1 on every generated and hand-built case; see the pinning
 * note at the top of oracle_semantic.cpp ("PARITY UNPINNED" outside the
 * reference's own known-answer cases).
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../include/nbgraft.h"

#define ORC_API __attribute__((visibility("default")))

/* end of object i's run: off[i+1] (CSR) or end[i] (spans), see nbgraft.h */
#define END_OF(m, i) ((m)->end ? (m)->end[i] : (m)->off[(i) + 1])

static int live_differs(const nbg_kvmap* from, const nbg_kvmap* to, uint64_t i) {
  /* for k, v := range to { if from[k] != v { requireUpdate = true } } */
  int differs = 0;
  for (uint32_t t = to->off[i]; t < END_OF(to, i); ++t) {
    uint64_t want = 0;
    for (uint32_t f = from->off[i]; f < END_OF(from, i); ++f)
      if (from->ent[f].key == to->ent[t].key) { want = from->ent[f].val; break; }
    if (want != to->ent[t].val) differs = 1;
  }
  return differs;
}

static int map_not_equal(const nbg_kvmap* a, int a_nil, const nbg_kvmap* b, int b_nil, uint64_t i) {
  if (a_nil != b_nil) return 1;
  const uint32_t na = END_OF(a, i) - a->off[i], nb = END_OF(b, i) - b->off[i];
  if (na != nb) return 1;
  for (uint32_t k = 0; k < na; ++k) {
    const nbg_kv* x = &a->ent[a->off[i] + k];
    int found = 0;
    for (uint32_t q = 0; q < nb; ++q) {
      const nbg_kv* y = &b->ent[b->off[i] + q];
      if (x->key == y->key) { found = (x->val == y->val); break; }
    }
    if (!found) return 1;
  }
  return 0;
}

static void set_bit(uint64_t* bitmap, uint64_t i, int v) {
  if (v) __atomic_fetch_or(&bitmap[i >> 6], 1ull << (i & 63), __ATOMIC_RELAXED);
}

/* ------------------------------------------------------------- ranges */
static void sts_range(const nbg_sts_batch* in, uint64_t lo, uint64_t hi, uint8_t* reason, uint64_t* bitmap) {
  for (uint64_t i = lo; i < hi; ++i) {
    unsigned r = 0;
    const int32_t rf = in->replicas_from[i], rt = in->replicas_to[i];
    if (rf == NBG_REPLICAS_NIL || rt == NBG_REPLICAS_NIL) r |= NBG_STS_NILREPLICAS;
    else if (rf != rt) {
      r |= NBG_STS_REPLICAS;
      if (in->tmpl_from[i] != in->tmpl_to[i]) r |= NBG_STS_TMPL_LABELS;
    }
    if (live_differs(&in->labels_from, &in->labels_to, i)) r |= NBG_STS_LABELS;
    if (live_differs(&in->ann_from, &in->ann_to, i)) r |= NBG_STS_ANNOTATIONS;
    if (memcmp(in->pod_from + i * NBG_POD_SLOTS, in->pod_to + i * NBG_POD_SLOTS, NBG_POD_SLOTS * 8) != 0)
      r |= NBG_STS_PODSPEC;
    reason[i] = (uint8_t)r;
    set_bit(bitmap, i, (r & NBG_STS_UPDATE_MASK) != 0);
  }
}

static void sts_sparse_range(const nbg_sts_sparse_batch* in, uint64_t lo, uint64_t hi, uint8_t* reason, uint64_t* bitmap) {
  for (uint64_t i = lo; i < hi; ++i) {
    unsigned r = 0;
    const int32_t rf = in->replicas_from[i], rt = in->replicas_to[i];
    if (rf == NBG_REPLICAS_NIL || rt == NBG_REPLICAS_NIL) r |= NBG_STS_NILREPLICAS;
    else if (rf != rt) {
      r |= NBG_STS_REPLICAS;
      if (in->tmpl_from[i] != in->tmpl_to[i]) r |= NBG_STS_TMPL_LABELS;
    }
    if (live_differs(&in->labels_from, &in->labels_to, i)) r |= NBG_STS_LABELS;
    if (live_differs(&in->ann_from, &in->ann_to, i)) r |= NBG_STS_ANNOTATIONS;
    /* rebuild both dense tiles and compare them: the sparse form must mean exactly the dense one */
    uint64_t a[NBG_POD_SLOTS], b[NBG_POD_SLOTS];
    for (int side = 0; side < 2; ++side) {
      const nbg_podsparse* p = side ? &in->pod_to : &in->pod_from;
      uint64_t* t = side ? b : a;
      uint32_t at = p->off[i];
      for (unsigned s = 0; s < NBG_POD_SLOTS; ++s)
        t[s] = (p->mask[i * 3 + (s >> 6)] >> (s & 63) & 1) ? p->val[at++] : 0;
      if (at & 1) { if (p->val[at] != 0) r |= 0x40; ++at; }   /* runs are zero-padded to an even count */
      if (at != END_OF(p, i)) r |= 0x40;                      /* malformed input: popcount(mask) != run length */
    }
    if (memcmp(a, b, sizeof a) != 0) r |= NBG_STS_PODSPEC;
    reason[i] = (uint8_t)r;
    set_bit(bitmap, i, (r & NBG_STS_UPDATE_MASK) != 0);
  }
}

static void svc_range(const nbg_svc_batch* in, uint64_t lo, uint64_t hi, uint8_t* reason, uint64_t* bitmap) {
  for (uint64_t i = lo; i < hi; ++i) {
    unsigned r = 0;
    const unsigned fl = in->flags[i];
    if (live_differs(&in->labels_from, &in->labels_to, i)) r |= NBG_SVC_LABELS;
    if (live_differs(&in->ann_from, &in->ann_to, i)) r |= NBG_SVC_ANNOTATIONS;
    if (map_not_equal(&in->sel_from, (fl & NBG_F_FROM_SEL_NIL) != 0, &in->sel_to, (fl & NBG_F_TO_SEL_NIL) != 0, i))
      r |= NBG_SVC_SELECTOR;
    if (memcmp(in->ports_from + i * NBG_SVC_SLOTS, in->ports_to + i * NBG_SVC_SLOTS, NBG_SVC_SLOTS * 8) != 0)
      r |= NBG_SVC_PORTS;
    reason[i] = (uint8_t)r;
    set_bit(bitmap, i, (r & NBG_SVC_UPDATE_MASK) != 0);
  }
}

static void vs_range(const nbg_vs_batch* in, uint64_t lo, uint64_t hi, uint8_t* reason, uint64_t* bitmap) {
  for (uint64_t i = lo; i < hi; ++i) {
    unsigned r = 0;
    int update = 0;
    const unsigned fl = in->flags[i];
    if (fl & NBG_F_FROM_SPEC_MISS) r = NBG_VS_FROM_MISSING;
    else if (fl & NBG_F_TO_SPEC_MISS) { r = NBG_VS_TO_MISSING; update = 1; }
    else {
      int differs = memcmp(in->spec_from + i * NBG_VS_SLOTS, in->spec_to + i * NBG_VS_SLOTS, NBG_VS_SLOTS * 8) != 0;
      if (map_not_equal(&in->hdr_from, (fl & NBG_F_FROM_HDR_NIL) != 0, &in->hdr_to, (fl & NBG_F_TO_HDR_NIL) != 0, i))
        differs = 1;
      if (differs) { r = NBG_VS_SPEC_DIFFERS; update = 1; }
    }
    reason[i] = (uint8_t)r;
    set_bit(bitmap, i, update);
  }
}

typedef struct { int64_t sec; uint32_t nsec; } inst;
static int after(inst a, inst b) { return a.sec > b.sec || (a.sec == b.sec && a.nsec > b.nsec); }

static void scan(const int64_t* sec, const uint32_t* nsf, uint32_t lo, uint32_t hi, int* all_ok, int* all_idle,
                 inst* best, uint32_t* idx) {
  *all_ok = 1; *all_idle = 1; best->sec = 0; best->nsec = 0; *idx = 0;
  for (uint32_t r = lo; r < hi; ++r) {
    inst t = {sec[r], nsf[r] & NBG_NSEC_MASK};
    if (!(nsf[r] & NBG_REC_PARSE_OK)) *all_ok = 0;
    if (!(nsf[r] & NBG_REC_IDLE)) *all_idle = 0;
    if (r == lo || after(t, *best)) { *best = t; *idx = r - lo; }
  }
}

static void cull_range(const nbg_cull_batch* in, const nbg_cull_params* p, uint64_t lo, uint64_t hi, uint8_t* action,
                       uint32_t* argmax, uint64_t* cull_bitmap, uint64_t* period_bitmap) {
  const inst now = {p->now_sec, (uint32_t)p->now_nsec}, now_trunc = {p->now_sec, 0};
  for (uint64_t i = lo; i < hi; ++i) {
    unsigned act = 0; uint32_t amax = NBG_ARGMAX_NONE;
    inst ann = {in->ann_sec[i], in->ann_nsf[i] & NBG_NSEC_MASK}, chk = {in->chk_sec[i], in->chk_nsf[i] & NBG_NSEC_MASK};
    int ann_ok = (in->ann_nsf[i] & NBG_ANN_PARSE_OK) != 0, stop = (in->ann_nsf[i] & NBG_ANN_STOP_SET) != 0;
    int chk_present = (in->chk_nsf[i] & NBG_CHK_PRESENT) != 0, chk_ok = (in->chk_nsf[i] & NBG_CHK_PARSE_OK) != 0;
    int proceed = 1;
    if (p->mode == NBG_CULL_RECONCILE) {
      if (stop) { act = NBG_ACT_STOPPED; proceed = 0; }
      else if (in->nb_flags[i] & NBG_NB_POD_MISSING) { act = NBG_ACT_NOPOD; proceed = 0; }
      else if (!((in->nb_flags[i] & NBG_NB_LA_PRESENT) && chk_present)) {
        act |= NBG_ACT_INIT; ann = now_trunc; ann_ok = 1; chk = now_trunc; chk_ok = 1; chk_present = 1;
      }
    }
    if (proceed) {
      int period = 0;
      if (chk_present) {
        inst next = {(chk_ok ? chk.sec : -62135596800ll) + (int64_t)p->check_period_min * 60, chk_ok ? chk.nsec : 0};
        period = after(now, next);
      }
      if (period) act |= NBG_ACT_PERIOD;
      if (p->mode == NBG_CULL_PREDICATES || period) {
        unsigned src = 0; int all_ok, all_idle; inst best; uint32_t idx;
        if ((in->k_end ? in->k_end[i] : in->k_off[i + 1]) > in->k_off[i]) {
          scan(in->k_sec, in->k_nsf, in->k_off[i], (in->k_end ? in->k_end[i] : in->k_off[i + 1]), &all_ok, &all_idle, &best, &idx);
          if (!all_idle) { ann = now_trunc; ann_ok = 1; src = 1; }
          else if (all_ok) { inst t = {best.sec, 0}; if (ann_ok && !after(ann, t)) { ann = t; src = 2; amax = idx; } }
        }
        if ((in->t_end ? in->t_end[i] : in->t_off[i + 1]) > in->t_off[i]) {
          scan(in->t_sec, in->t_nsf, in->t_off[i], (in->t_end ? in->t_end[i] : in->t_off[i + 1]), &all_ok, &all_idle, &best, &idx);
          if (all_ok) { inst t = {best.sec, 0}; if (ann_ok && !after(ann, t)) { ann = t; src = 3; amax = idx; } }
        }
        act |= src;
        if (!stop && ann_ok) {
          inst cap = {ann.sec + (int64_t)p->cull_idle_min * 60, ann.nsec};
          if (after(now, cap)) act |= NBG_ACT_CULL;
        }
      }
    }
    action[i] = (uint8_t)act; argmax[i] = amax;
    set_bit(cull_bitmap, i, (act & NBG_ACT_CULL) != 0);
    set_bit(period_bitmap, i, (act & NBG_ACT_PERIOD) != 0);
  }
}

/* ------------------------------------------------------------ threading */
typedef struct {
  int kind; const void* in; const nbg_cull_params* p; uint64_t lo, hi;
  uint8_t* out8; uint32_t* out32; uint64_t* bm1; uint64_t* bm2;
} job;
static void* run_job(void* arg) {
  job* j = (job*)arg;
  switch (j->kind) {
    case 0: sts_range((const nbg_sts_batch*)j->in, j->lo, j->hi, j->out8, j->bm1); break;
    case 1: svc_range((const nbg_svc_batch*)j->in, j->lo, j->hi, j->out8, j->bm1); break;
    case 4: sts_sparse_range((const nbg_sts_sparse_batch*)j->in, j->lo, j->hi, j->out8, j->bm1); break;
    case 2: vs_range((const nbg_vs_batch*)j->in, j->lo, j->hi, j->out8, j->bm1); break;
    default: cull_range((const nbg_cull_batch*)j->in, j->p, j->lo, j->hi, j->out8, j->out32, j->bm1, j->bm2); break;
  }
  return NULL;
}
static double run(int kind, const void* in, const nbg_cull_params* p, uint64_t n, int threads, uint8_t* out8,
                  uint32_t* out32, uint64_t* bm1, uint64_t* bm2) {
  if (threads < 1) threads = 1;
  if (threads > 512) threads = 512;
  memset(bm1, 0, (n + 63) / 64 * 8);
  if (bm2) memset(bm2, 0, (n + 63) / 64 * 8);
  job* jobs = (job*)calloc((size_t)threads, sizeof(job));
  pthread_t* tid = (pthread_t*)calloc((size_t)threads, sizeof(pthread_t));
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (int t = 0; t < threads; ++t) {
    /* 64-aligned ranges so threads never share a bitmap word */
    const uint64_t words = (n + 63) / 64;
    uint64_t lo = words * (uint64_t)t / (uint64_t)threads * 64, hi = words * (uint64_t)(t + 1) / (uint64_t)threads * 64;
    if (lo > n) lo = n;
    if (hi > n) hi = n;
    jobs[t] = (job){kind, in, p, lo, hi, out8, out32, bm1, bm2};
    if (t > 0) pthread_create(&tid[t], NULL, run_job, &jobs[t]);
  }
  run_job(&jobs[0]);
  for (int t = 1; t < threads; ++t) pthread_join(tid[t], NULL);
  clock_gettime(CLOCK_MONOTONIC, &t1);
  free(jobs); free(tid);
  return (double)(t1.tv_sec - t0.tv_sec) + (double)(t1.tv_nsec - t0.tv_nsec) * 1e-9;
}

/* Each returns the seconds spent in the compare loops (for bench.py). */
ORC_API double orc2_diff_sts(const nbg_sts_batch* in, uint64_t n, int threads, uint8_t* reason, uint64_t* bitmap) {
  return run(0, in, NULL, n, threads, reason, NULL, bitmap, NULL);
}
/* CopyDeploymentSetFields (COMMON:136-163): as sts_range, but Replicas are compared as POINTERS (:152) —
 * different unless both nil — and there is neither a panic nor a template-label pre-step. */
ORC_API void orc2_diff_deploy(const nbg_sts_batch* in, uint64_t n, uint8_t* reason, uint64_t* bitmap) {
  for (uint64_t i = 0; i < n; ++i) {
    unsigned r = 0;
    if (in->replicas_from[i] != NBG_REPLICAS_NIL || in->replicas_to[i] != NBG_REPLICAS_NIL) r |= NBG_STS_REPLICAS;
    if (live_differs(&in->labels_from, &in->labels_to, i)) r |= NBG_STS_LABELS;
    if (live_differs(&in->ann_from, &in->ann_to, i)) r |= NBG_STS_ANNOTATIONS;
    if (memcmp(in->pod_from + i * NBG_POD_SLOTS, in->pod_to + i * NBG_POD_SLOTS, NBG_POD_SLOTS * 8) != 0)
      r |= NBG_STS_PODSPEC;
    reason[i] = (uint8_t)r;
    set_bit(bitmap, i, (r & NBG_STS_UPDATE_MASK) != 0);
  }
}
/* generateStatefulSet as a patch of the Notebook's dense PodSpec tile (notebook_controller.go:433-523; slot layout of
 * csrc/host/pack.cpp: pod-level slots 0..15, containers[0] at 40: workingDir 42, count tags 46, ports 52.., env 56..).
 * Written for the tile as a plain array, one slot at a time — not the way the kernel splits it over lanes. */
ORC_API void orc2_generate_sts_tile(const uint64_t* nb_tile, uint64_t prefix_hash, unsigned flags,
                                    const nbg_sts_gen_consts* k, uint64_t* out) {
  memcpy(out, nb_tile, NBG_POD_SLOTS * 8);
  if (flags & NBG_GEN_PREPATCHED) return;                     /* the host generated it */
  if (out[42] == 0) out[42] = k->working_dir;                  /* :495-497 */
  uint64_t tag = out[46];
  unsigned ports = (unsigned)(tag & 0xFFFF), env = (unsigned)((tag >> 16) & 0xFFFF);
  if (ports == 0xFFFF) {                                       /* :498-506: nil, not merely empty */
    out[52] = k->port_name; out[53] = k->port_rest;
    ports = 1;
  }
  const unsigned env_n = env == 0xFFFF ? 0 : env;
  int found = 0;                                               /* :417-431: an existing NB_PREFIX is left alone */
  for (unsigned e = 0; e < env_n && e < 14; ++e) found |= out[56 + 2 * e] == k->env_name;
  if (!found && env_n < 14) {
    out[56 + 2 * env_n] = k->env_name;
    out[57 + 2 * env_n] = prefix_hash;
    env = env_n + 1;
  }
  out[46] = (tag & ~0xFFFFFFFFull) | ports | ((uint64_t)env << 16);
  if (k->add_fsgroup && !(out[14] & 2)) {                      /* :514-521: securityContext == nil */
    out[7] = 100;
    out[9] = 0;                                                /* the rest of &PodSecurityContext{FSGroup: &g}: empty (a nil pointer has a marker there) */
    out[14] |= 6;
  }
}
ORC_API void orc2_diff_sts_gen(const nbg_sts_batch* in, const nbg_sts_gen_inputs* gen, const nbg_sts_gen_consts* k, uint64_t n,
                               uint8_t* reason, uint64_t* bitmap) {
  uint64_t want[NBG_POD_SLOTS];
  for (uint64_t i = 0; i < n; ++i) {
    unsigned r = 0;
    const int rf = (gen->flags[i] & NBG_GEN_STOPPED) ? 0 : 1, rt = in->replicas_to[i];
    if (rt == NBG_REPLICAS_NIL) r |= NBG_STS_NILREPLICAS;
    else if (rf != rt) { r |= NBG_STS_REPLICAS; if (in->tmpl_from[i] != in->tmpl_to[i]) r |= NBG_STS_TMPL_LABELS; }
    if (live_differs(&in->labels_from, &in->labels_to, i)) r |= NBG_STS_LABELS;
    for (uint32_t t = in->ann_to.off[i]; t < in->ann_to.off[i + 1]; ++t)
      if (in->ann_to.ent[t].val != 0) r |= NBG_STS_ANNOTATIONS;           /* desired has none */
    orc2_generate_sts_tile(in->pod_from + i * NBG_POD_SLOTS, gen->prefix_hash[i], gen->flags[i], k, want);
    if (memcmp(want, in->pod_to + i * NBG_POD_SLOTS, NBG_POD_SLOTS * 8) != 0) r |= NBG_STS_PODSPEC;
    reason[i] = (uint8_t)r;
    set_bit(bitmap, i, (r & NBG_STS_UPDATE_MASK) != 0);
  }
}
ORC_API double orc2_diff_sts_sparse(const nbg_sts_sparse_batch* in, uint64_t n, int threads, uint8_t* reason, uint64_t* bitmap) {
  return run(4, in, NULL, n, threads, reason, NULL, bitmap, NULL);
}
ORC_API double orc2_diff_svc(const nbg_svc_batch* in, uint64_t n, int threads, uint8_t* reason, uint64_t* bitmap) {
  return run(1, in, NULL, n, threads, reason, NULL, bitmap, NULL);
}
ORC_API double orc2_diff_vs(const nbg_vs_batch* in, uint64_t n, int threads, uint8_t* reason, uint64_t* bitmap) {
  return run(2, in, NULL, n, threads, reason, NULL, bitmap, NULL);
}
ORC_API double orc2_cull(const nbg_cull_batch* in, uint64_t n, const nbg_cull_params* p, int threads, uint8_t* action,
                         uint32_t* argmax, uint64_t* cull_bitmap, uint64_t* period_bitmap) {
  return run(3, in, p, n, threads, action, argmax, cull_bitmap, period_bitmap);
}

/* Owner join on keys (notebook_controller.go:157-170): first list entry with the
 * notebook's key.  Sort-free and obviously right: a chained table of list
 * positions, filled back to front so each chain head is the lowest index. */
ORC_API double orc2_owner_join(const uint64_t* nb_key, uint64_t n, const uint64_t* sts_key, uint64_t m, uint32_t* match) {
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  uint64_t buckets = 1024;
  while (buckets < 2 * m) buckets <<= 1;
  int64_t* head = (int64_t*)malloc(buckets * sizeof(int64_t));
  int64_t* next = (int64_t*)malloc((m ? m : 1) * sizeof(int64_t));
  for (uint64_t b = 0; b < buckets; ++b) head[b] = -1;
  for (int64_t j = (int64_t)m - 1; j >= 0; --j) {
    if (sts_key[j] == 0) continue;
    const uint64_t b = (sts_key[j] * 0x9E3779B97F4A7C15ull >> 20) & (buckets - 1);
    next[j] = head[b];
    head[b] = j;
  }
  for (uint64_t i = 0; i < n; ++i) {
    uint32_t found = NBG_ARGMAX_NONE;
    if (nb_key[i] != 0) {
      const uint64_t b = (nb_key[i] * 0x9E3779B97F4A7C15ull >> 20) & (buckets - 1);
      for (int64_t j = head[b]; j >= 0; j = next[j])
        if (sts_key[j] == nb_key[i]) { found = (uint32_t)j; break; }
    }
    match[i] = found;
  }
  free(head); free(next);
  clock_gettime(CLOCK_MONOTONIC, &t1);
  return (double)(t1.tv_sec - t0.tv_sec) + (double)(t1.tv_nsec - t0.tv_nsec) * 1e-9;
}

/* ODH comparator family on packed vectors (include/nbgraft.h: nbg_generic_batch). */
ORC_API void orc2_diff_generic(const nbg_generic_batch* in, uint64_t n, int mode, uint8_t* reason, uint64_t* bitmap) {
  memset(bitmap, 0, (n + 63) / 64 * 8);
  for (uint64_t i = 0; i < n; ++i) {
    unsigned r = 0;
    if (mode == NBG_GEN_DEEPEQUAL) {
      if (map_not_equal(&in->labels_a, 0, &in->labels_b, 0, i)) r |= NBG_GEN_LABELS;
      if (map_not_equal(&in->spec_a, 0, &in->spec_b, 0, i)) r |= NBG_GEN_SPEC;
    } else {
      if (END_OF(&in->labels_a, i) - in->labels_a.off[i] != END_OF(&in->labels_b, i) - in->labels_b.off[i] ||
          live_differs(&in->labels_b, &in->labels_a, i)) r |= NBG_GEN_LABELS;
      if (END_OF(&in->spec_a, i) - in->spec_a.off[i] != END_OF(&in->spec_b, i) - in->spec_b.off[i] ||
          live_differs(&in->spec_b, &in->spec_a, i)) r |= NBG_GEN_SPEC;
    }
    reason[i] = (uint8_t)r;
    set_bit(bitmap, i, r != 0);
  }
}

/* Webhook restart-block predicate on packed sparse tiles (nbg_restart_batch). */
static int sparse_differs(const nbg_podsparse* a, const nbg_podsparse* b, uint64_t i) {
  if (memcmp(a->mask + i * 3, b->mask + i * 3, 24) != 0) return 1;
  const uint32_t na = END_OF(a, i) - a->off[i];
  if (na != END_OF(b, i) - b->off[i]) return 1;
  return memcmp(a->val + a->off[i], b->val + b->off[i], (size_t)na * 8) != 0;
}
ORC_API void orc2_restart_block(const nbg_restart_batch* in, uint64_t n, uint8_t* decision, uint64_t* bitmap) {
  memset(bitmap, 0, (n + 63) / 64 * 8);
  for (uint64_t i = 0; i < n; ++i) {
    unsigned d = 0;
    if (!(in->flags[i] & (NBG_RB_CREATE | NBG_RB_STOPPED | NBG_RB_RESTARTING))) {
      if (sparse_differs(&in->old_spec, &in->updated_spec, i)) d = NBG_RB_EXTERNAL_CHANGE;
      else if (!sparse_differs(&in->old_spec, &in->mutated_spec, i)) d = NBG_RB_NO_CHANGE;
      else d = NBG_RB_BLOCKED;
    }
    decision[i] = (uint8_t)d;
    set_bit(bitmap, i, (d & NBG_RB_BLOCKED) != 0);
  }
}
