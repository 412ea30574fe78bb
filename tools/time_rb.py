"""Time the restart-block kernel of a given libnbgraft build on ~1 M resident pairs.
usage: NBGRAFT_LIB=build/variants/libnbgraft_X.so python tools/time_rb.py"""
import importlib, os, statistics, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
pkg = importlib.import_module("odh-kubeflow_b200"); host, N = pkg.host, pkg._native
dev = importlib.import_module("odh-kubeflow_b200.device")
ctx = pkg.Context([0])
fo = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 11, 16384, 400)
rb = host.RestartBatch()
for k, o in enumerate(fo.objects(0, 16384)):
    m = o["notebook"].clone(); host.odh_mutate(m, 2 | 4 | 8 | 16)
    rb.add(o["notebook"], o["notebook"], m)
d = dev.DeviceBatch.from_host(rb, tile=64)
ms = []
for it in range(23):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); d.run(ctx); e.record(); torch.cuda.synchronize()
    if it >= 3: ms.append(s.elapsed_time(e))
digest = int(d.outs[0].cpu().numpy().astype(np.uint64).sum())   # to compare builds with each other; parity is tests' job
med = statistics.median(ms)
print("%s  pairs=%d  median %.1f us  %.0f GB/s (%.3f of 6551)  result-sum=%d" % (
    os.path.basename(os.environ.get("NBGRAFT_LIB", "default")), d.n, med * 1e3, d.algorithmic_bytes / (med * 1e-3) / 1e9,
    d.algorithmic_bytes / (med * 1e-3) / 1e9 / 6551, digest))
