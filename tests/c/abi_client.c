/*
 * abi_client.c — libnbgraft.so driven from plain C99, the way a cgo shim (or any other FFI)
 * would drive it: include/nbgraft.h only, hand-built flat arrays, no Python, no C++.
 *
 *   no CUDA device : nbg_init must fail with NBG_E_NODEVICE (there is no CPU path) -> "nodevice"
 *   with a device  : known answers through nbg_diff_sts and nbg_cull -> "ok <n> checks"
 *
 * Known answers: the reference's TestNotebookIsIdle table (culling_controller_test.go:143-263) and
 * its BDD label assertion (notebook_controller_bdd_test.go:90), plus the label-loop asymmetries of
 * CopyStatefulSetFields (components/common/reconcilehelper/util.go:109-113).
 * Built and run by tests/test_c_client.py:  gcc -std=c99 -pedantic -Wall -Wextra -Werror.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "nbgraft.h"
#include "nbgraft_host.h"

static int checks = 0, failures = 0;
#define CHECK(cond, what)                                                        \
  do {                                                                           \
    ++checks;                                                                    \
    if (!(cond)) { ++failures; fprintf(stderr, "FAIL %s (line %d)\n", what, __LINE__); } \
  } while (0)

static uint64_t H(const char* s) { return nbg_hash64(s, strlen(s)); }

static void sort_kv(nbg_kv* e, int n) { /* entries of one object are sorted by key */
  int i, j;
  for (i = 1; i < n; ++i) {
    nbg_kv x = e[i];
    for (j = i; j > 0 && e[j - 1].key > x.key; --j) e[j] = e[j - 1];
    e[j] = x;
  }
}

#define N_STS 5
static void sts_known_answers(nbg_ctx* ctx) {
  static uint64_t pod_from[N_STS * NBG_POD_SLOTS], pod_to[N_STS * NBG_POD_SLOTS];
  int32_t rep_from[N_STS] = {1, 1, 1, 0, 1}, rep_to[N_STS] = {1, 1, 1, 1, 1};
  uint64_t tmpl_from[N_STS], tmpl_to[N_STS];
  /* labels: object 0 equal both sides (BDD case); 1 live-only label; 2 desired-only label; 3, 4 none */
  nbg_kv lf[8], lt[8];
  uint32_t lf_off[N_STS + 1], lt_off[N_STS + 1], none_off[N_STS + 1] = {0, 0, 0, 0, 0, 0};
  uint8_t reason[N_STS];
  uint64_t bitmap[1];
  nbg_sts_batch b;
  unsigned i, s;
  int nf = 0, nt = 0, rc;

  for (i = 0; i < N_STS; ++i) {
    for (s = 0; s < NBG_POD_SLOTS; ++s) pod_from[i * NBG_POD_SLOTS + s] = pod_to[i * NBG_POD_SLOTS + s] = (s % 7 == 0) ? H("busybox") + s : 0;
    tmpl_from[i] = tmpl_to[i] = H("statefulset=test-notebook");
  }
  lf_off[0] = lt_off[0] = 0;
  lf[nf].key = H("testLabel"); lf[nf++].val = H("testLabelValue");          /* 0: from */
  lt[nt].key = H("testLabel"); lt[nt++].val = H("testLabelValue");          /* 0: to   */
  lf_off[1] = (uint32_t)nf; lt_off[1] = (uint32_t)nt;
  lt[nt].key = H("added-by-someone"); lt[nt++].val = H("x");                /* 1: live only */
  lf_off[2] = (uint32_t)nf; lt_off[2] = (uint32_t)nt;
  lf[nf].key = H("only-desired"); lf[nf++].val = H("y");                    /* 2: desired only */
  lf_off[3] = (uint32_t)nf; lt_off[3] = (uint32_t)nt;
  lf_off[4] = lf_off[5] = (uint32_t)nf; lt_off[4] = lt_off[5] = (uint32_t)nt;
  sort_kv(lf, 1); sort_kv(lt, 1);
  tmpl_to[3] = H("statefulset=other");                                      /* 3: replicas 0 vs 1 + template labels differ */
  pod_to[4 * NBG_POD_SLOTS + 191] ^= 1;                                     /* 4: one PodSpec slot differs */

  memset(&b, 0, sizeof b);
  b.replicas_from = rep_from; b.replicas_to = rep_to;
  b.tmpl_from = tmpl_from; b.tmpl_to = tmpl_to;
  b.pod_from = pod_from; b.pod_to = pod_to;
  b.labels_from.off = lf_off; b.labels_from.ent = lf;
  b.labels_to.off = lt_off; b.labels_to.ent = lt;
  b.ann_from.off = none_off; b.ann_from.ent = lf;
  b.ann_to.off = none_off; b.ann_to.ent = lt;
  rc = nbg_diff_sts(ctx, &b, N_STS, reason, bitmap);
  CHECK(rc == NBG_OK, "nbg_diff_sts status");
  CHECK(reason[0] == 0, "BDD: StatefulSet labels equal the Notebook's -> no update");
  CHECK(reason[1] == NBG_STS_LABELS, "live-only label -> update");
  CHECK(reason[2] == 0, "desired-only label is not seen by the loop over live keys");
  CHECK(reason[3] == (NBG_STS_REPLICAS | NBG_STS_TMPL_LABELS), "replicas differ, template labels copied");
  CHECK(reason[4] == NBG_STS_PODSPEC, "one PodSpec slot differs");
  CHECK(bitmap[0] == 0x1Au, "bitmap = objects 1, 3, 4");

  rep_to[2] = NBG_REPLICAS_NIL;                                             /* Go would panic at COMMON:123 */
  rc = nbg_diff_sts(ctx, &b, N_STS, reason, bitmap);
  CHECK(rc == NBG_E_NILREPLICAS && (reason[2] & NBG_STS_NILREPLICAS), "nil replicas is reported, not computed");
  CHECK(nbg_diff_sts(ctx, NULL, N_STS, reason, bitmap) == NBG_E_INVAL, "NULL batch");
  CHECK(nbg_diff_sts(ctx, &b, 0, NULL, NULL) == NBG_OK, "empty batch");
}

#define N_NB 10
static void cull_known_answers(nbg_ctx* ctx) {
  /* TestNotebookIsIdle, in table order; "now" is whatever instant the batch is evaluated at */
  const int64_t now = 1767225600; /* 2026-01-01T00:00:00Z */
  const char* last[N_NB] = {NULL, NULL, NULL, NULL, "should-fail", "2021-08-30T15:37:36.990063Z", "1900-08-30T15:37:36.990063Z",
                            "now+0", "now-360", "now-180"};
  const int has_map[N_NB] = {0, 1, 1, 1, 1, 1, 1, 1, 1, 1}, stop[N_NB] = {0, 0, 1, 0, 0, 0, 0, 0, 0, 0};
  const int idle_min[N_NB] = {1440, 1440, 1440, 1440, 1440, 1440, 1440, 5, 5, 5};
  const int want[N_NB] = {0, 0, 0, 0, 0, 1, 1, 0, 1, 0};
  int64_t ann_sec[N_NB], chk_sec[N_NB], no_sec[1] = {0};
  uint32_t ann_nsf[N_NB], chk_nsf[N_NB], off[N_NB + 1], no_nsf[1] = {0};
  uint8_t nb_flags[N_NB], action[N_NB];
  uint32_t argmax[N_NB];
  uint64_t cull_bits[1], period_bits[1];
  nbg_cull_batch b;
  nbg_cull_params p;
  int i, pass;

  for (i = 0; i < N_NB; ++i) {
    int64_t sec = 0; int32_t nsec = 0, zone = 0; int ok = 0;
    if (last[i] && strncmp(last[i], "now", 3) == 0) { sec = now + atoi(last[i] + 3); ok = 1; }
    else if (last[i]) ok = nbh_parse_rfc3339(last[i], &sec, &nsec, &zone);
    ann_sec[i] = ok ? sec : 0;
    ann_nsf[i] = (ok ? ((uint32_t)nsec | NBG_ANN_PARSE_OK) : 0u) | (stop[i] ? NBG_ANN_STOP_SET : 0u);
    nb_flags[i] = (uint8_t)((has_map[i] && last[i]) ? NBG_NB_LA_PRESENT : 0);
    chk_sec[i] = 0; chk_nsf[i] = 0;                      /* no check-timestamp: irrelevant to notebookIsIdle */
    off[i] = 0;
  }
  off[N_NB] = 0;
  CHECK(ann_sec[5] == 1630337856 && (ann_nsf[5] & NBG_NSEC_MASK) == 990063000u, "host RFC3339 parse of the 2021 literal");
  CHECK(ann_sec[6] == -2188110144, "host RFC3339 parse of the 1900 literal");
  memset(&b, 0, sizeof b);
  b.k_off = off; b.t_off = off; b.k_sec = no_sec; b.k_nsf = no_nsf; b.t_sec = no_sec; b.t_nsf = no_nsf;
  b.ann_sec = ann_sec; b.ann_nsf = ann_nsf; b.chk_sec = chk_sec; b.chk_nsf = chk_nsf; b.nb_flags = nb_flags;
  for (pass = 0; pass < 2; ++pass) {
    const int idle = pass == 0 ? 1440 : 5;
    p.now_sec = now; p.now_nsec = 0; p.cull_idle_min = idle; p.check_period_min = 1; p.mode = NBG_CULL_PREDICATES;
    CHECK(nbg_cull(ctx, &b, N_NB, &p, action, argmax, cull_bits, period_bits) == NBG_OK, "nbg_cull status");
    for (i = 0; i < N_NB; ++i)
      if (idle_min[i] == idle) CHECK(((action[i] & NBG_ACT_CULL) != 0) == want[i], "TestNotebookIsIdle row");
  }
}

/* The resident fleet from C: the StatefulSet rows of sts_known_answers() would need the sparse form, so this
 * drives the culler rows only (a NULL batch leaves that predicate out): put three notebooks into slots 5, 0, 2
 * of an 8-slot store, replace one of them, step. */
static void store_known_answers(nbg_ctx* ctx) {
  const int64_t now = 1767225600;
  nbg_store* st = NULL;
  uint64_t idx[3] = {5, 0, 2}, one[1] = {0};
  /* last-activity: now-6 min, now-3 min, now (CULL_IDLE_TIME = 5) -> idle, not idle, not idle */
  int64_t ann_sec[3] = {now - 360, now - 180, now}, chk_sec[3] = {0, 0, 0}, no_sec[1] = {0};
  uint32_t ann_nsf[3] = {NBG_ANN_PARSE_OK, NBG_ANN_PARSE_OK, NBG_ANN_PARSE_OK}, chk_nsf[3] = {0, 0, 0}, off[4] = {0, 0, 0, 0}, no_nsf[1] = {0};
  uint8_t nb_flags[3] = {NBG_NB_LA_PRESENT, NBG_NB_LA_PRESENT, NBG_NB_LA_PRESENT};
  uint8_t action[8];
  uint32_t argmax[8];
  uint64_t cull_bits[1], period_bits[1];
  nbg_cull_batch b;
  nbg_cull_params p;
  nbg_step_outputs out;
  nbg_store_stats stats;

  CHECK(nbg_store_create(ctx, 0, 8, &st) == NBG_OK && st != NULL, "nbg_store_create");
  if (!st) return;
  memset(&b, 0, sizeof b);
  b.k_off = off; b.t_off = off; b.k_sec = no_sec; b.k_nsf = no_nsf; b.t_sec = no_sec; b.t_nsf = no_nsf;
  b.ann_sec = ann_sec; b.ann_nsf = ann_nsf; b.chk_sec = chk_sec; b.chk_nsf = chk_nsf; b.nb_flags = nb_flags;
  CHECK(nbg_store_put(st, idx, 3, NULL, NULL, NULL, &b) == NBG_OK, "nbg_store_put");
  memset(&out, 0, sizeof out);
  out.action = action; out.argmax = argmax; out.cull_bitmap = cull_bits; out.period_bitmap = period_bits;
  p.now_sec = now; p.now_nsec = 0; p.cull_idle_min = 5; p.check_period_min = 1; p.mode = NBG_CULL_PREDICATES;
  CHECK(nbg_store_step(st, 8, NULL, &p, &out) == NBG_OK, "nbg_store_step");
  CHECK((action[5] & NBG_ACT_CULL) && !(action[0] & NBG_ACT_CULL) && !(action[2] & NBG_ACT_CULL), "idle slot is 5");
  CHECK(cull_bits[0] == (1u << 5), "cull bitmap");
  ann_sec[0] = now - 3600;                                    /* slot 0 now holds a notebook idle for an hour */
  CHECK(nbg_store_put(st, one, 1, NULL, NULL, NULL, &b) == NBG_OK, "replace one row");
  CHECK(nbg_store_step(st, 8, NULL, &p, &out) == NBG_OK, "step after the replacement");
  CHECK(cull_bits[0] == ((1u << 5) | 1u), "slots 5 and 0 are idle now");
  one[0] = 8;
  CHECK(nbg_store_put(st, one, 1, NULL, NULL, NULL, &b) == NBG_E_RANGE, "slot past the capacity");
  CHECK(nbg_store_get_stats(st, &stats) == NBG_OK && stats.puts == 2 && stats.rows_put == 4 && stats.capacity == 8, "stats");
  nbg_store_destroy(st);
}

int main(void) {
  nbg_ctx* ctx = NULL;
  int rc;
  CHECK(nbg_abi_version() == NBG_ABI_VERSION, "ABI version");
  CHECK(nbg_hash64("", 0) == 0 && nbg_hash64("a", 1) != 0, "hash of the empty string is 0, of nothing else");
  CHECK(nbg_shard_of(H("kubeflow-user"), 1) == 0 && nbg_shard_of(H("kubeflow-user"), 8) < 8, "shard map range");
  CHECK(strlen(nbg_strerror(NBG_E_NODEVICE)) > 0, "strerror");
  rc = nbg_init(NULL, 0, &ctx);
  if (rc == NBG_E_NODEVICE) {
    CHECK(ctx == NULL, "no context without a device");
    if (failures) return 1;
    printf("nodevice %d checks\n", checks);
    return 0;
  }
  CHECK(rc == NBG_OK && ctx != NULL, "nbg_init");
  if (rc != NBG_OK) return 1;
  sts_known_answers(ctx);
  cull_known_answers(ctx);
  store_known_answers(ctx);
  CHECK(nbg_launch_count(ctx) >= 6, "kernels were launched");
  nbg_destroy(ctx);
  if (failures) return 1;
  printf("ok %d checks\n", checks);
  return 0;
}
