"""Time the dense StatefulSet kernel against the one that generates the desired tile from the Notebook's
(nbg_diff_sts_gen_device), same notebooks, resident inputs, CUDA events, L2 flushed between launches.
usage: python tools/time_sts_gen.py 300000"""
import importlib, os, statistics, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
pkg = importlib.import_module("odh-kubeflow_b200")
dev = importlib.import_module("odh-kubeflow_b200.device")
host = pkg.host
n = int(sys.argv[1]) if len(sys.argv) > 1 else 300000
fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, n, max(200, n // 50))
g, h = host.StsBatch(), host.StsBatch()
t_gen = t_host = 0.0
for i in range(n):
    nb, live = fleet.get(i, "notebook"), fleet.get(i, "sts_to")
    t0 = time.perf_counter(); g.add_gen(nb, live); t1 = time.perf_counter()
    h.add(host.generate_statefulset(nb), live); t2 = time.perf_counter()
    t_gen += t1 - t0; t_host += t2 - t1
print("host side, one thread, per notebook: add_gen %.1f us   generate_statefulset + add %.1f us" % (t_gen / n * 1e6, t_host / n * 1e6))
ctx = pkg.Context([0])
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
for name, b in (("sts (dense)", h), ("sts_gen", host.StsGenView(g))):
    d = dev.DeviceBatch.from_host(b)
    ms = []
    for it in range(13):
        flush.fill_(it)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); outs = d.run(ctx); e.record(); torch.cuda.synchronize()
        if it >= 3: ms.append(s.elapsed_time(e))
    med = statistics.median(ms)
    print("%-12s n=%d  median %.1f us  min %.1f us  %.0f GB/s of %d algorithmic bytes  result-sum=%d" % (
        name, n, med * 1e3, min(ms) * 1e3, d.algorithmic_bytes / (med * 1e-3) / 1e9, d.algorithmic_bytes,
        int(outs[0].cpu().numpy().astype(np.uint64).sum())))
