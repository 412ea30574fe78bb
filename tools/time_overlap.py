"""Step time with the three small kernels on side streams (fork/join), resident inputs.
usage: [NBGRAFT_LIB=...] [ORDER=sts_first] python tools/time_overlap.py 1250000"""
import importlib, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
pkg = importlib.import_module("odh-kubeflow_b200")
dev = importlib.import_module("odh-kubeflow_b200.device")
host = pkg.host
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_250_000
ctx = pkg.Context([0])
f = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, n, max(200, n // 50))
sts, svc, vs = f.pack()
cf = host.CullFleet(host.SEED_BASE + 3, n)
now = (cf.now_sec, cf.now_nsec)
d = [dev.DeviceBatch.from_host(b) for b in (sts.sparse(), svc.gen(), vs.gen(), cf.pack())]
total = sum(b.algorithmic_bytes for b in d)
main = torch.cuda.current_stream()
side = [torch.cuda.Stream() for _ in range(3)]
fork, joins = torch.cuda.Event(), [torch.cuda.Event() for _ in range(3)]
order = os.environ.get("ORDER", "small_first")


def step():
    fork.record(main)
    if order == "sts_first":
        d[0].run(ctx, stream=main.cuda_stream)
    for k in range(3):
        side[k].wait_event(fork)
        if k == 2:
            d[3].run(ctx, now=now, stream=side[k].cuda_stream)
        else:
            d[1 + k].run(ctx, stream=side[k].cuda_stream)
        joins[k].record(side[k])
    if order != "sts_first":
        d[0].run(ctx, stream=main.cuda_stream)
    for k in range(3):
        main.wait_event(joins[k])


for _ in range(5): step()
torch.cuda.synchronize()
ms = []
for rep in range(5):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(20): step()
    e.record(); torch.cuda.synchronize()
    ms.append(s.elapsed_time(e) / 20)
med = statistics.median(ms)
print("%s order=%s  step %.1f us (min %.1f)  %.0f GB/s (%.3f of 6551)  %.2f G objects/s" % (
    os.path.basename(os.environ.get("NBGRAFT_LIB", "default")), order,
    med * 1e3, min(ms) * 1e3, total / (med * 1e-3) / 1e9, total / (med * 1e-3) / 1e9 / 6551, n / (med * 1e-3) / 1e9))
