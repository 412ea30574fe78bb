"""The resident step of a given libnbgraft build: whole-step time, the StatefulSet kernel's time INSIDE the running
step (events recorded by the library around that launch) and alone.  Prints a sum of the result bytes so that
builds can be compared with each other; parity is the tests' job.
usage: [NBGRAFT_LIB=build/variants/libnbgraft_X.so] python tools/time_step.py 2500000"""
import importlib, os, statistics, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
pkg = importlib.import_module("odh-kubeflow_b200")
dev = importlib.import_module("odh-kubeflow_b200.device")
host, N = pkg.host, pkg._native
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_500_000
ctx = pkg.Context([0])
f = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, n, max(200, n // 50))
sts, svc, vs = f.pack(lean=True)
cf = host.CullFleet(host.SEED_BASE + 5, n)
cb = cf.pack()
now = (cf.now_sec, cf.now_nsec)
d = [dev.DeviceBatch.from_host(b) for b in (sts.sparse(), svc.gen(), vs.gen(), cb)]
total = sum(b.algorithmic_bytes for b in d)
step = dev.ResidentStep(ctx, *d, now=now, time_sts=True)
for _ in range(5): step.run()
torch.cuda.synchronize(); step.sts_ms(); step.reserve(20)
runs, inside = [], []
for rep in range(5):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(20): step.run()
    e.record(); torch.cuda.synchronize()
    runs.append(s.elapsed_time(e) / 20); inside.append(statistics.mean(step.sts_ms())); step.reserve(20)
alone = []
for _ in range(12):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); d[0].run(ctx); e.record(); torch.cuda.synchronize(); alone.append(s.elapsed_time(e))
ok = int(d[0].outs[0].cpu().numpy().astype(np.uint64).sum())
peak = 6544.0
print("%s n=%d  step %.1f us (%.3f of peak)  sts inside %.1f us (%.3f)  sts alone %.1f us (%.3f)  result-sum=%d" % (
    os.path.basename(os.environ.get("NBGRAFT_LIB", "default")), n, statistics.median(runs) * 1e3,
    total / (statistics.median(runs) * 1e-3) / 1e9 / peak, statistics.median(inside) * 1e3,
    d[0].algorithmic_bytes / (statistics.median(inside) * 1e-3) / 1e9 / peak, statistics.median(alone[2:]) * 1e3,
    d[0].algorithmic_bytes / (statistics.median(alone[2:]) * 1e-3) / 1e9 / peak, ok))
