"""One small launch of every SURVEY 8(f) kernel (for ncu captures)."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
pkg = importlib.import_module("odh-kubeflow_b200"); host, N = pkg.host, pkg._native
dev = importlib.import_module("odh-kubeflow_b200.device")
ctx = pkg.Context([0])
n = 200_000
f = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 6, n, n // 50)
nb, st, _ = f.join_lists(n)
for _ in range(3): ctx.owner_join(nb, st)
cf = host.CullFleet(host.SEED_BASE + 3, 60_000)
nbs = [json.loads(cf.json(i)) for i in range(60_000)]
tb = pkg.engine.CullTextBatch(nbs)
for _ in range(3): ctx.cull_text(tb, cf.now_sec, cf.now_nsec, 1440, 1, 1)
fo = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 11, 8192, 200)
objs = fo.objects(0, 8192)
gb, rb = host.GenBatch(N.GEN_DEEPEQUAL, "metadata.labels", "spec.template.spec"), host.RestartBatch()
for o in objs:
    gb.add(o["sts_from"], o["sts_to"]); m = o["notebook"].clone(); host.odh_mutate(m, 30); rb.add(o["notebook"], o["notebook"], m)
for hb in (gb, rb):
    d = dev.DeviceBatch.from_host(hb, tile=32)
    for _ in range(3): d.run(ctx)
torch.cuda.synchronize()
_, svc, vs = f.pack(sts=False)
for hb in (svc.gen(), vs.gen()):
    d = dev.DeviceBatch.from_host(hb)
    for _ in range(3): d.run(ctx)
torch.cuda.synchronize()
print("ok")
