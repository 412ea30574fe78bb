"""The resident store's tick: replace 1 % of the rows (host buffers) + step over all slots with host outputs.
usage: [NBGRAFT_LIB=...] python tools/time_store.py 1250000"""
import importlib, os, statistics, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
pkg = importlib.import_module("odh-kubeflow_b200")
host = pkg.host
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_250_000
ctx = pkg.Context([0])
f = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, 4 * n, max(200, n // 50))
cf = host.CullFleet(host.SEED_BASE + 5, 4 * n)
now = (cf.now_sec, cf.now_nsec)
a, b, c = f.pack(0, n, lean=True)
cb = cf.pack(0, n)
store = pkg.engine.Store(ctx, n)
t0 = time.perf_counter(); store.put(np.arange(n, dtype=np.uint64), a.sparse(), b.gen(), c.gen(), cb); load = time.perf_counter() - t0
nbytes = sum(x.algorithmic_bytes() for x in (a.sparse(), b.gen(), c.gen(), cb))
consts = host.vs_gen_consts()
out = pkg.engine.StepBuffers(n, pinned=True)
rng = np.random.default_rng(7)
m = max(n // 100, 1)
puts, ticks = [], []
for t in range(8):
    slots = rng.choice(n, size=m, replace=False).astype(np.uint64)
    alt = (n + t * m + np.arange(m)).astype(np.uint64)
    s1, s2, s3 = f.pack(idx=alt, lean=True)
    s4 = cf.pack(idx=alt)
    rows = (s1.sparse(), s2.gen(), s3.gen(), s4)
    t0 = time.perf_counter(); store.put(slots, *rows); p = time.perf_counter() - t0
    r = store.step(n, consts, now[0], now[1], out=out); ticks.append(time.perf_counter() - t0); puts.append(p)
ok = int(r["sts"][0].astype(np.uint64).sum()) + int(r["cull"][0].astype(np.uint64).sum())   # to compare builds; parity is the tests' job
print("%s n=%d  load %.3f s (%.1f GB/s)  put of %d rows %.2f ms  tick %.2f ms  result-sum=%d  %s" % (
    os.path.basename(os.environ.get("NBGRAFT_LIB", "default")), n, load, nbytes / load / 1e9, m,
    statistics.median(puts[1:]) * 1e3, statistics.median(ticks[1:]) * 1e3, ok, store.stats()))
store.close()
