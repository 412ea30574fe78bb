"""RFC3339 parse kernel at a size where launch overhead no longer matters: the 1 M-record tables of the bench,
tiled on the device.  usage: python tools/time_parse.py [tiles=16]"""
import ctypes as C, importlib, json, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
pkg = importlib.import_module("odh-kubeflow_b200")
host, N = pkg.host, pkg._native
tiles = int(sys.argv[1]) if len(sys.argv) > 1 else 16
ctx = pkg.Context([0])
cf = host.CullFleet(host.SEED_BASE + 3, 250_000)
la, st = [], []
for i in range(250_000):
    for k in json.loads(cf.json(i))["kernels"]:
        la.append(k["last_activity"]); st.append(k["execution_state"])


def tiled(strings):
    t = pkg.engine.StringTable(strings)
    total = int(t.off[-1])
    body = np.frombuffer(b"".join(s.encode() for s in strings), dtype=np.uint8)
    off = np.concatenate([t.off[:-1].astype(np.int64) + k * total for k in range(tiles)] + [np.array([tiles * total])])
    assert off[-1] < 2 ** 32
    raw = np.concatenate([body] * tiles + [np.zeros(32, np.uint8)])
    return torch.from_numpy(off.astype(np.uint32).view(np.int32)).cuda(), torch.from_numpy(raw).cuda(), int(off[-1])


oa, ba, na = tiled(la)
os_, bs, ns = tiled(st)
R = len(la) * tiles
va = N.nbg_strings(C.cast(oa.data_ptr(), N.u32p), C.cast(ba.data_ptr(), N.u8p))
vs = N.nbg_strings(C.cast(os_.data_ptr(), N.u32p), C.cast(bs.data_ptr(), N.u8p))
sec = torch.zeros(R, dtype=torch.int64, device="cuda"); nsf = torch.zeros(R, dtype=torch.int32, device="cuda")
zone = torch.zeros(R, dtype=torch.int32, device="cuda")
stream = torch.cuda.current_stream().cuda_stream
lib = N.nbg()
ms = []
for it in range(13):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    assert lib.nbg_parse_records_device(ctx.handle, 0, C.byref(va), C.byref(vs), R, sec.data_ptr(), nsf.data_ptr(), zone.data_ptr(), stream) == 0
    e.record(); torch.cuda.synchronize()
    if it >= 3: ms.append(s.elapsed_time(e))
med = statistics.median(ms)
nbytes = na + ns + 8 * (R + 1) + 16 * R
print("records=%d  median %.1f us  %.1f G records/s  %.0f GB/s (%.3f of 6551)  %.1f B/record" % (
    R, med * 1e3, R / (med * 1e-3) / 1e9, nbytes / (med * 1e-3) / 1e9, nbytes / (med * 1e-3) / 1e9 / 6551, nbytes / R))
