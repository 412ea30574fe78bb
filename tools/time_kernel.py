"""Time one kernel of a given libnbgraft build on resident inputs (CUDA events, L2 > inputs).
usage: NBGRAFT_LIB=build/variants/libnbgraft_X.so python tools/time_kernel.py cull|sts|sts_sparse|svc|vs|svc_gen|vs_gen 1250000"""
import importlib, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
pkg = importlib.import_module("odh-kubeflow_b200")
dev = importlib.import_module("odh-kubeflow_b200.device")
host = pkg.host
kind, n = sys.argv[1], int(sys.argv[2])
ctx = pkg.Context([0])
if kind == "cull":
    cf = host.CullFleet(host.SEED_BASE + 3, n); b = cf.pack(); kw = dict(now=(int(os.environ.get("NOW_SEC", cf.now_sec)), cf.now_nsec))
else:
    f = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, n, max(200, n // 50))
    base = kind.replace("_gen", "").replace("_sparse", "")
    b = f.pack(sts=base == "sts", svc=base == "svc", vs=base == "vs")[{"sts": 0, "svc": 1, "vs": 2}[base]]; kw = {}
    if kind == "sts_sparse": b = b.sparse()
    if kind.endswith("_gen"): b = b.gen()
d = dev.DeviceBatch.from_host(b)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
for mode in ((1, 0) if kind == "cull" else (None,)):
    if mode is not None: kw["mode"] = mode
    ms = []
    for it in range(13):
        flush.fill_(it)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); outs = d.run(ctx, **kw); e.record(); torch.cuda.synchronize()
        if it >= 3: ms.append(s.elapsed_time(e))
    digest = int(outs[0].cpu().numpy().astype(np.uint64).sum())      # to compare builds with each other; parity is tests' job
    med = statistics.median(ms)
    print("%s %s mode=%s n=%d  median %.1f us  min %.1f us  %.0f GB/s  result-sum=%d" % (
        os.path.basename(os.environ.get("NBGRAFT_LIB", "default")), kind, mode, n, med * 1e3, min(ms) * 1e3,
        d.algorithmic_bytes / (med * 1e-3) / 1e9, digest))
