"""nbg_step_device issued directly vs replayed from a captured CUDA graph, small and large fleets.
usage: python tools/time_graph.py"""
import importlib, os, statistics, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
pkg = importlib.import_module("odh-kubeflow_b200")
dev = importlib.import_module("odh-kubeflow_b200.device")
host = pkg.host
ctx = pkg.Context([0])
for n in (10_000, 100_000, 1_250_000):
    f = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 2, n, max(200, n // 50))
    sts, svc, vs = f.pack()
    cf = host.CullFleet(host.SEED_BASE + 3, n)
    d = [dev.DeviceBatch.from_host(b) for b in (sts.sparse(), svc.gen(), vs.gen(), cf.pack())]
    step = dev.ResidentStep(ctx, *d, now=(cf.now_sec, cf.now_nsec))
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3): step.run()
    torch.cuda.synchronize()
    want = [o.cpu().numpy().copy() for b in d for o in b.outs]
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        step.run()
    for b in d:
        for o in b.outs: o.zero_()
    g.replay(); torch.cuda.synchronize()
    same = all(np.array_equal(o.cpu().numpy(), w) for o, w in zip((o for b in d for o in b.outs), want))
    res = {}
    for name, fn in (("direct", step.run), ("graph", g.replay)):
        ms = []
        with torch.cuda.stream(s):
            for rep in range(7):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(50): fn()
                b.record(); torch.cuda.synchronize()
                ms.append(a.elapsed_time(b) / 50)
        res[name] = statistics.median(ms)
    print("n=%d  direct %.1f us/step  graph %.1f us/step  graph outputs identical: %s" % (n, res["direct"] * 1e3, res["graph"] * 1e3, same))
