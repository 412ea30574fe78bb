"""Time the host-buffer C-ABI step (pinned buffers) — for tuning the copy pipeline."""
import ctypes as C, importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench
pkg = importlib.import_module("odh-kubeflow_b200"); host, N = pkg.host, pkg._native
per = int(sys.argv[1]) if len(sys.argv) > 1 else 1_250_000
ctx = pkg.Context([0]); lib = N.nbg()
fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, per, max(200, per // 50))
sts, svc, vs = fleet.pack(); cf = host.CullFleet(host.SEED_BASE + 5, per); cull = cf.pack()
keep, views = [], {}
vs_consts = host.vs_gen_consts()
for kind, b in (("sts", sts.sparse()), ("svc", svc.gen()), ("vs", vs.gen()), ("cull", cull)):
    arr = b.arrays(); views[kind] = (bench.struct_over(N, kind, bench.pinned_copy(lib, arr, keep)), sum(a.nbytes for a in arr.values()))
words = (per + 63) // 64
out = bench.pinned_copy(lib, {"r": np.zeros(per, np.uint8), "b": np.zeros(words, np.uint64), "a": np.zeros(per, np.uint32), "b2": np.zeros(words, np.uint64)}, keep)
rp, bp, ap_, b2p = (out[k].ctypes.data_as(C.c_void_p) for k in ("r", "b", "a", "b2"))
params = N.nbg_cull_params(cf.now_sec, cf.now_nsec, 1440, 1, 1)
calls = {"sts": lambda: lib.nbg_diff_sts_sparse(ctx.handle, C.byref(views["sts"][0]), per, rp, bp),
         "svc": lambda: lib.nbg_diff_svc_gen(ctx.handle, C.byref(views["svc"][0]), per, rp, bp),
         "vs": lambda: lib.nbg_diff_vs_gen(ctx.handle, C.byref(views["vs"][0]), C.byref(vs_consts), per, rp, bp),
         "cull": lambda: lib.nbg_cull(ctx.handle, C.byref(views["cull"][0]), per, C.byref(params), rp, ap_, bp, b2p)}
tot = 0
for name, fn in calls.items():
    for _ in range(2): assert fn() == 0
    t0 = time.perf_counter()
    for _ in range(5): fn()
    dt = (time.perf_counter() - t0) / 5; tot += dt
    print("scale=%s %-5s %.2f ms  %.1f GB/s" % (os.environ.get("NBG_CHUNK_SCALE", "1"), name, dt * 1e3, views[name][1] / dt / 1e9))
print("scale=%s step %.2f ms  %.1f M objects/s" % (os.environ.get("NBG_CHUNK_SCALE", "1"), tot * 1e3, per / tot / 1e6))
import torch
h = torch.empty(1 << 30, dtype=torch.uint8, pin_memory=True); d = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
for _ in range(2): d.copy_(h, non_blocking=True)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5): d.copy_(h, non_blocking=True)
torch.cuda.synchronize(); print("raw pinned H2D 1 GiB: %.1f GB/s" % (5 * (1 << 30) / (time.perf_counter() - t0) / 1e9))
