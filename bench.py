#!/usr/bin/env python
"""bench.py — the graft's one headline measurement (contract: README of the task).

A "step" is one pass of the hot path over one GPU's share of the fleet:
diff_sts + diff_svc + diff_vs (Copy*Fields, COMMON:107-219) + cull (CULL:86-241,
315-428) over the same notebooks.  Unit: objects/s where one object = one Notebook
(its StatefulSet pair, Service pair, VirtualService pair and culler record set).

  value      inputs resident in HBM, CUDA-event timed, max over ranks
  e2e        the same step through the host-buffer C ABI (nbg_diff_*, nbg_cull):
             pinned host buffers -> H2D -> kernel -> D2H inside the timed region
  roofline   dominant kernel (diff_sts): algorithmic bytes / its own CUDA-event time
  cpu_baseline  the semantic oracle (CPU restatement of the Go predicates on whole
             objects) on all host threads, bounded sample; N=1, rank 0 only

--impl reference times that CPU restatement alone (the reference is Go and cannot
be built here: no toolchain, un-vendored modules — see DESIGN.md).
"""
import argparse
import ctypes as C
import importlib
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "objects/sec (batched reconcile-diff + cull predicate per Notebook)"
UNIT = "objects/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--per-gpu", type=int, default=1_250_000, help="notebooks per GPU (10M / 8)")
    ap.add_argument("--cpu-sample", type=int, default=40_000, help="notebooks in the CPU baseline sample")
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary configs (#2, #3, #4)")
    ap.add_argument("--no-numa", action="store_true", help="multi-rank: do not bind each rank to the CPUs next to its GPU")
    ap.add_argument("--no-overlap", action="store_true", help="launch the four kernels of a step on one stream")
    return ap.parse_args()


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            pass

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        out = self.p.communicate()[0]
        sm, mx, reasons = [], [], set()
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def pinned_copy(lib, arrays, keep):
    """Copy numpy arrays into cudaHostAlloc'ed memory; returns name -> numpy view."""
    out = {}
    for name, a in arrays.items():
        nbytes = max(a.nbytes, 1)
        p = lib.nbg_alloc_pinned(nbytes)
        if not p:
            raise MemoryError("nbg_alloc_pinned(%d)" % nbytes)
        keep.append(p)
        buf = (C.c_uint8 * nbytes).from_address(p)
        v = np.frombuffer(buf, dtype=np.uint8, count=a.nbytes).view(a.dtype)
        v[...] = a
        out[name] = v
    return out


def struct_over(N, kind, arr):
    """The nbg_* batch struct over host arrays (here: pinned copies of a batch's arrays)."""
    kind = {"sts": "StsSparseBatch", "svc": "svc_gen", "vs": "vs_gen", "cull": "CullBatch"}[kind]
    return N.batch_struct(kind, lambda name, ct: arr[name].ctypes.data_as(C.POINTER(ct)), lambda name: name in arr)


class CpuReference:
    """The CPU restatement of the Go path (semantic oracle: whole objects, reflect.DeepEqual
    walk, RFC3339 strings) on `threads` threads over a bounded sample of notebooks."""

    def __init__(self, pkg, O, sample, seed, threads):
        host = pkg.host
        self.sample, self.threads, self.L = sample, threads, O.lib()
        fleet = host.Fleet(host.FLEET_KUBEFLOW, seed, sample, max(sample // 50, 1))
        self.objs = fleet.objects(0, sample)
        vp = C.c_void_p
        self.arrs = []
        for a, b in (("sts_from", "sts_to"), ("svc_from", "svc_to"), ("vs_from", "vs_to")):
            self.arrs.append(((vp * sample)(*[o[a]._h for o in self.objs]), (vp * sample)(*[o[b]._h for o in self.objs])))
        self.cf = host.CullFleet(seed + 1, sample)
        self.texts = [self.cf.json(i).encode() for i in range(sample)]
        self.jarr = (C.c_char_p * sample)(*self.texts)

    def step(self):
        """One pass of all four predicates over the sample; seconds inside the oracle loops."""
        t = 0.0
        for kind in range(3):
            t += self.L.orc_bench_pairs(kind, self.arrs[kind][0], self.arrs[kind][1], self.sample, self.threads, None)
        t += self.L.orc_bench_cull(self.jarr, self.sample, self.cf.now_sec, self.cf.now_nsec, 1440, 1, 1, self.threads, None, None)
        return t

    def rate(self, min_cpu_seconds=10.0, max_reps=64):
        total, reps = 0.0, 0
        while reps < max_reps and (total * self.threads < min_cpu_seconds or reps < 2):
            total += self.step()
            reps += 1
        return self.sample * reps / total, {"sample_notebooks": self.sample, "repetitions": reps,
                                            "cpu_seconds": round(total * self.threads, 2)}


def run_reference(args, rank, world):
    """--impl reference: the reference's algorithm on the host cores (the Go code itself
    cannot be built in this image; the oracle port stands in, kind = "port")."""
    if rank != 0:
        return
    pkg = importlib.import_module("odh-kubeflow_b200")
    import oracle as O
    threads = os.cpu_count() or 1
    ref = CpuReference(pkg, O, args.cpu_sample, pkg.host.SEED_BASE + 5, threads)
    secs = []
    t0 = time.time()
    for s in range(args.warmup + args.steps):
        t = ref.step()
        if s >= args.warmup:
            secs.append(t)
        if time.time() - t0 > 200 and len(secs) >= 1:
            break
    value = args.cpu_sample / statistics.median(secs)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": len(secs),
            "warmup": args.warmup, "ms_per_step": 1000.0 * statistics.median(secs), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "config #5 share: %d-notebook sample of the 10M-notebook fleet per step, all 4 predicates per notebook"
                                   % args.cpu_sample, "note": "CPU port of the Go predicates (semantic oracle); Go itself unavailable on this box"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "%d notebooks per step, object-level DeepEqual + RFC3339 strings" % args.cpu_sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def bind_near_gpu(torch, local):
    """Multi-rank runs: keep this rank's threads (and so the pinned buffers it first-touches) on the CPUs
    NVML reports as local to its GPU, so that eight ranks do not pull their H2D traffic across sockets.
    Returns the CPU list, or None when the topology cannot be read (then nothing is changed)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        p = torch.cuda.get_device_properties(local)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(("%08x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)).encode())
        words = pynvml.nvmlDeviceGetCpuAffinity(h, ((os.cpu_count() or 1) + 63) // 64)
        cpus = [64 * i + bit for i, w in enumerate(words) for bit in range(64) if (w >> bit) & 1]
        cpus = sorted(set(cpus) & os.sched_getaffinity(0))
        if not cpus or len(cpus) == len(os.sched_getaffinity(0)):
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    pkg = importlib.import_module("odh-kubeflow_b200")
    host, N = pkg.host, pkg._native
    dev = importlib.import_module("odh-kubeflow_b200.device")
    import oracle as O  # checker + cpu_baseline leg only

    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    near = bind_near_gpu(torch, local) if (world > 1 and not args.no_numa) else None
    # host threads are shared by the ranks of this node
    if "NBH_THREADS" not in os.environ:
        os.environ["NBH_THREADS"] = str(max(1, (os.cpu_count() or 1) // world))
    ctx = pkg.Context([local])
    lib = N.nbg()

    # ---- this rank's shard of the fleet: namespaces are partitioned by nbg_shard_of, every
    # shard takes the same number of notebooks (weak scaling)
    per = args.per_gpu
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, per * world * 4, max(200, per * world // 50))
    t0 = time.time()
    shard_imbalance = None
    if world > 1:
        shard = fleet.shard_indices(world, rank)
        sizes = torch.zeros(world, dtype=torch.float64, device="cuda")
        sizes[rank] = float(len(shard))
        dist.all_reduce(sizes)                                   # set-up only, not in any timed region
        shard_imbalance = float(sizes.max() / sizes.mean())      # what hash sharding of Zipf namespaces gives
        idx = shard[:per]
    else:
        idx = np.arange(per, dtype=np.uint64)
    assert len(idx) == per, "shard too small"
    sts, svc, vs = fleet.pack(idx=idx)
    make_cpu_s, pack_cpu_s = fleet.pack_timing()
    cfleet = host.CullFleet(host.SEED_BASE + 5 + 1000 * rank, per)
    cull = cfleet.pack()
    pack_s = time.time() - t0
    now = (cfleet.now_sec, cfleet.now_nsec)

    # the step compares StatefulSets on SPARSE PodSpec tiles (same answer as the dense tiles, ~4x fewer bytes)
    # and Services / VirtualServices against a desired side generated on the device from the few
    # Notebook words generateService / generateVirtualService depend on
    sts_sp, svc_g, vs_g = sts.sparse(), svc.gen(), vs.gen()
    vs_consts = host.vs_gen_consts()
    d_sts, d_svc, d_vs, d_cull = (dev.DeviceBatch.from_host(b, device=local) for b in (sts_sp, svc_g, vs_g, cull))
    batches = [("diff_sts", d_sts), ("diff_svc", d_svc), ("diff_vs", d_vs), ("cull", d_cull)]
    alg_bytes = {name: b.algorithmic_bytes for name, b in batches}

    def launch(name, b):
        if name == "cull":
            return b.run(ctx, now=now, cull_idle_min=1440, check_period_min=1, mode=N.CULL_RECONCILE)
        return b.run(ctx)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: inputs resident in HBM.  One step = one nbg_step_device call: the StatefulSet kernel on the
    # main stream, the three small latency-bound kernels beside it on the context's side streams (the
    # library forks and joins them with events).  --no-overlap: four launches on one stream instead.
    main = torch.cuda.current_stream()
    resident = dev.ResidentStep(ctx, d_sts, d_svc, d_vs, d_cull, now=now, cull_idle_min=1440, check_period_min=1,
                                mode=N.CULL_RECONCILE, time_sts=True) if not args.no_overlap else None
    serial_ev = []

    def step(timed=False):
        if resident is not None:
            resident.run()
            return
        for k, (name, b) in enumerate(batches):
            if k == 0 and timed:
                serial_ev.append((torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)))
                serial_ev[-1][0].record(main)
            launch(name, b)
            if k == 0 and timed:
                serial_ev[-1][1].record(main)

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local)
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = ctx.launch_count()
    if resident is not None:
        resident.sts_ms()                                  # drop the warm-up pairs
        resident.reserve(args.steps)
        torch.cuda.synchronize()
    start.record()
    for s in range(args.steps):
        step(timed=True)
    stop.record()
    barrier()
    # dominant kernel, inside the timed region, on its stream (events recorded by the library around that launch)
    dom_ms = statistics.mean(resident.sts_ms() if resident is not None else [a.elapsed_time(b) for a, b in serial_ev])
    gpu_launches = ctx.launch_count() - launches0
    ms_total = start.elapsed_time(stop)
    # per-kernel durations: a separate, serialised pass (same launches, one stream, CUDA events around each)
    ev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in batches] for _ in range(args.steps)]
    for s in range(args.steps):
        for k, (name, b) in enumerate(batches):
            ev[s][k][0].record()
            launch(name, b)
            ev[s][k][1].record()
    torch.cuda.synchronize()
    kernel_ms = {name: statistics.mean(ev[s][k][0].elapsed_time(ev[s][k][1]) for s in range(args.steps))
                 for k, (name, _) in enumerate(batches)}
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    units = torch.tensor([float(per)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(units, op=dist.ReduceOp.SUM)
    ms_step = t.item() / args.steps
    value = units.item() / (ms_step / 1000.0)

    # ---- parity of what was just computed (outside the timed region)
    want = {"diff_sts": O.packed_diff_sts_sparse(sts_sp, threads=8), "diff_svc": O.packed_diff_svc(svc, threads=8),
            "diff_vs": O.packed_diff_vs(vs, threads=8), "cull": O.packed_cull(cull, now, 1440, 1, 1, threads=8)}
    parity = True
    for name, b in batches:
        outs = [o.cpu().numpy() for o in b.outs]
        parity &= bool(np.array_equal(outs[0], want[name][0]))
        parity &= bool(np.array_equal(outs[-1].view(np.uint64), want[name][3 if name == "cull" else 1]))
    dense_want = O.packed_diff_sts(sts, threads=8)
    parity &= bool(np.array_equal(dense_want[0], want["diff_sts"][0]))          # sparse tiles == dense tiles
    if not parity:
        raise SystemExit("bench: device results differ from the packed oracle")
    # the dense-tile kernel on the same pairs, timed alone for comparison
    d_dense = dev.DeviceBatch.from_host(sts, device=local)
    for _ in range(3):
        d_dense.run(ctx)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        d_dense.run(ctx)
    e1.record()
    torch.cuda.synchronize()
    dense_ms = e0.elapsed_time(e1) / 5
    parity_dense = bool(np.array_equal(d_dense.outs[0].cpu().numpy(), dense_want[0]))
    dense_bytes = d_dense.algorithmic_bytes
    del d_dense
    torch.cuda.empty_cache()

    # ---- e2e: the host-buffer C ABI with pinned host memory, copies inside the timed region
    keep = []
    views, outs_h = {}, {}
    words = (per + 63) // 64
    for kind, b in (("sts", sts_sp), ("svc", svc_g), ("vs", vs_g), ("cull", cull)):
        views[kind] = (struct_over(N, kind, pinned_copy(lib, b.arrays(), keep)), sum(a.nbytes for a in b.arrays().values()))
    out_arr = pinned_copy(lib, {"r": np.zeros(per, np.uint8), "b": np.zeros(words, np.uint64), "a": np.zeros(per, np.uint32),
                                "b2": np.zeros(words, np.uint64)}, keep)
    rp, bp, ap_, b2p = (out_arr[k].ctypes.data_as(C.c_void_p) for k in ("r", "b", "a", "b2"))
    params = N.nbg_cull_params(now[0], now[1], 1440, 1, N.CULL_RECONCILE)

    o_sts = pinned_copy(lib, {"r": np.zeros(per, np.uint8), "b": np.zeros(words, np.uint64)}, keep)
    o_svc = pinned_copy(lib, {"r": np.zeros(per, np.uint8), "b": np.zeros(words, np.uint64)}, keep)
    o_vs = pinned_copy(lib, {"r": np.zeros(per, np.uint8), "b": np.zeros(words, np.uint64)}, keep)
    step_out = N.nbg_step_outputs(o_sts["r"].ctypes.data, o_sts["b"].ctypes.data, o_svc["r"].ctypes.data, o_svc["b"].ctypes.data,
                                  o_vs["r"].ctypes.data, o_vs["b"].ctypes.data, out_arr["r"].ctypes.data, out_arr["a"].ctypes.data,
                                  out_arr["b"].ctypes.data, out_arr["b2"].ctypes.data)

    def e2e_step():
        # the public one-call step: all four batches as one copy/compute pipeline
        rc = lib.nbg_step(ctx.handle, C.byref(views["sts"][0]), C.byref(views["svc"][0]), C.byref(views["vs"][0]), C.byref(vs_consts),
                          C.byref(views["cull"][0]), C.byref(params), per, C.byref(step_out))
        if rc:
            raise SystemExit("bench: e2e call failed (%d)" % rc)
    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    clocks = sampler.stop()  # sampled from the start of the device-timed steps to the end of the e2e steps
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = units.item() / te.item()
    h2d = sum(v[1] for v in views.values())
    d2h = 3 * (per + words * 8) + (per + 4 * per + 2 * words * 8)  # 3x (reason + bitmap); action + argmax + 2 bitmaps
    assert np.array_equal(out_arr["r"], want["cull"][0]), "e2e cull action differs from the oracle"
    assert np.array_equal(o_sts["r"], want["diff_sts"][0]) and np.array_equal(o_svc["r"], want["diff_svc"][0]) and \
        np.array_equal(o_vs["r"], want["diff_vs"][0]), "e2e diff results differ from the oracle"
    # ---- gather (outside every timed region): the shards' StatefulSet bitmaps come together on the host through the
    # per-shard index maps, as a sharded controller would assemble its fleet-wide "needs update" list
    gathered = None
    if world > 1:
        idx_t = torch.from_numpy(idx.astype(np.int64)).cuda()
        upd_t = torch.from_numpy(((o_sts["r"] & 0x0F) != 0).astype(np.uint8)).cuda()
        all_idx = [torch.empty_like(idx_t) for _ in range(world)]
        all_upd = [torch.empty_like(upd_t) for _ in range(world)]
        dist.all_gather(all_idx, idx_t)            # plumbing only: 11 bytes per notebook, once, after the measurements
        dist.all_gather(all_upd, upd_t)
        if rank == 0:
            fleet_bits = np.zeros(fleet.n, dtype=np.uint8)
            seen = np.zeros(fleet.n, dtype=np.uint8)
            for gi, gu in zip(all_idx, all_upd):
                gi = gi.cpu().numpy()
                seen[gi] += 1
                fleet_bits[gi] = gu.cpu().numpy()
            gathered = {"notebooks": int(seen.sum()), "each_notebook_in_exactly_one_shard": bool((seen <= 1).all() and seen.sum() == per * world),
                        "statefulsets_needing_update": int(fleet_bits.sum())}
    for p in keep:
        lib.nbg_free_pinned(p)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    traffic = None
    try:  # per-launch DRAM bytes of the dominant kernel from the committed ncu capture (same launch size only)
        tr = json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json")))["diff_sts_sparse_kernel"]
        if tr["objects_per_launch"] == per:
            traffic = tr["dram_bytes_read"] + tr["dram_bytes_write"]
    except Exception:
        pass
    sts_gbs = alg_bytes["diff_sts"] / (dom_ms * 1e-3) / 1e9
    iso_gbs = alg_bytes["diff_sts"] / (kernel_ms["diff_sts"] * 1e-3) / 1e9
    step_gbs = sum(alg_bytes.values()) / (ms_step * 1e-3) / 1e9          # this rank's bytes / max-over-ranks step time
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic",
        "config": {"workload": "config #5 share: %d Notebooks per GPU (10M / 8), StatefulSet+Service+VirtualService diff + culler "
                               "decision per Notebook" % per,
                   "object": "one Notebook = 3 desired/live pairs + its kernel/terminal records",
                   "tiles": "PodSpec tiles stored sparsely (192-bit mask + non-zero slots); dense tiles reported under dense_tiles",
                   "desired_side": "StatefulSet: packed from generateStatefulSet's output; Service and VirtualService: generated on the device "
                                   "from name/port/prefix/rewrite/host words (same answer as shipping generateService/VirtualService outputs)",
                   "sharding": "namespace hash (nbg_shard_of), no collective, equal notebooks per shard",
                   "shard_imbalance_max_over_mean": shard_imbalance, "gathered_on_rank0": gathered,
                   "cpu_binding": ("rank 0 bound to the %d CPUs NVML lists next to its GPU" % len(near)) if near else "none",
                   "streams": "one stream" if args.no_overlap else "one nbg_step_device call per step: StatefulSet kernel on the main stream, the three small kernels "
                              "on the context's side streams (event fork/join inside the library); per-kernel times are from a separate serialised pass",
                   "serial_sum_ms": sum(kernel_ms.values()),
                   "l2": "inputs (%.1f GB per GPU) exceed L2; no flush needed" % (sum(alg_bytes.values()) / 1e9),
                   "bytes_per_object": {k: round(v / per, 1) for k, v in alg_bytes.items()}},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d) * world, "d2h_bytes_per_step": int(d2h) * world,
                "ms_per_step": te.item() * 1e3, "effective_h2d_GBps_per_gpu": h2d / te.item() / 1e9,
                "bytes_note": "whole job: rank 0's bytes x ranks (every shard holds the same number of notebooks)",
                "timer": "host wall clock around the synchronous C-ABI call nbg_step, max over ranks"},
        "gpu_launches": int(gpu_launches),
        "roofline": {"bound": "hbm", "kernel": "diff_sts_sparse_kernel", "achieved": sts_gbs, "peak": peak, "unit": "GB/s",
                     "frac": sts_gbs / peak, "frac_of_nominal_8TBps": sts_gbs / 8000.0, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": int(alg_bytes["diff_sts"]), "avg_launch_ms": dom_ms,
                     "note": "timed inside the timed region on its own stream" + ("" if args.no_overlap else
                             ", while the three small kernels run concurrently on side streams and share HBM with it"),
                     "isolated": {"achieved": iso_gbs, "frac": iso_gbs / peak, "avg_launch_ms": kernel_ms["diff_sts"],
                                  "note": "same launch alone on the device (serialised pass after the timed region)"},
                     "whole_step": {"achieved": step_gbs, "frac": step_gbs / peak,
                                    "note": "algorithmic bytes of all four kernels / step time, per GPU"}},
        "kernels": {name: {"avg_ms": kernel_ms[name], "GBps": alg_bytes[name] / (kernel_ms[name] * 1e-3) / 1e9,
                           "frac_of_peak": alg_bytes[name] / (kernel_ms[name] * 1e-3) / 1e9 / peak} for name, _ in batches},
        "parity": "bit-exact vs packed oracle on all %d objects x 4 kernels (device path and e2e path); "
                  "sparse-tile results == dense-tile results" % per,
        "dense_tiles": {"kernel": "diff_sts_kernel", "avg_ms": dense_ms, "GBps": dense_bytes / (dense_ms * 1e-3) / 1e9,
                        "frac_of_peak": dense_bytes / (dense_ms * 1e-3) / 1e9 / peak, "bytes_per_object": round(dense_bytes / per, 1),
                        "parity": parity_dense, "note": "same pairs, dense 192-slot tiles (timed alone, back to back)"},
        "host_pack": {"objects_per_s": per * int(os.environ["NBH_THREADS"]) / pack_cpu_s, "threads": int(os.environ["NBH_THREADS"]),
                      "us_per_object_per_thread": pack_cpu_s / per * 1e6,
                      "note": "packing three desired/live pairs per Notebook into the batch arrays (objects already in memory), outside "
                              "the timed region; rate = objects x threads / summed thread-seconds",
                      "synthesis_us_per_object_per_thread": make_cpu_s / per * 1e6, "wall_seconds_with_synthesis": pack_s},
    }
    if world == 1:
        threads = os.cpu_count() or 1
        cpu_ref = CpuReference(pkg, O, args.cpu_sample, host.SEED_BASE + 5, threads)
        rate, info = cpu_ref.rate()
        cpu_ref.threads = 1                                   # the same port on ONE thread: the per-object cost (SURVEY 8d ii)
        t1 = cpu_ref.step()
        rate_1 = cpu_ref.sample / t1
        t_packed_1 = sum(x[-1] for x in (O.packed_diff_sts_sparse(sts_sp, threads=1), O.packed_diff_svc(svc, threads=1),
                                         O.packed_diff_vs(vs, threads=1), O.packed_cull(cull, now, 1440, 1, 1, threads=1)))
        # also the packed loops (the CPU version of OUR algorithm) for context
        t_packed = sum(x[-1] for x in (O.packed_diff_sts_sparse(sts_sp, threads=threads), O.packed_diff_svc(svc, threads=threads),
                                       O.packed_diff_vs(vs, threads=threads), O.packed_cull(cull, now, 1440, 1, 1, threads=threads)))
        line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": "%(sample_notebooks)d notebooks x %(repetitions)d repetitions (%(cpu_seconds)s CPU-s): "
                                          "object-level CopyStatefulSetFields/ServiceFields/VirtualService + culler on RFC3339 strings" % info,
                                "single_thread_value": rate_1, "single_thread_us_per_object": 1e6 / rate_1,
                                "packed_loops_value": per / t_packed, "packed_loops_single_thread_value": per / t_packed_1,
                                "packed_loops_note": "plain C loops over the same SoA batches, all %d objects, %d threads" % (per, threads)}
        if not args.no_extra:
            line["other_configs"] = other_configs(pkg, dev, ctx, torch, peak)
            del d_sts, d_svc, d_vs, d_cull, batches, resident
            torch.cuda.empty_cache()
            line["other_configs"]["config5_10M_on_one_gpu"] = ten_million_on_one_gpu(
                pkg, dev, ctx, torch, (sts_sp, svc_g, vs_g, cull), now, want, per, peak)
            line["other_configs"]["resident_store_1.25M"] = resident_store(
                pkg, O, ctx, torch, fleet, per, (sts_sp, svc_g, vs_g, cull), now, want, sum(alg_bytes.values()), peak)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def ten_million_on_one_gpu(pkg, dev, ctx, torch, host_batches, now, want, per, peak):
    """BASELINE config #5 at its full size on ONE GPU: the 1.25 M-notebook batches tiled 8x on the device (10 M
    notebooks, ~11 GB resident), one nbg_step_device per step.  Parity through the tiling property: every tile must
    repeat the 1.25 M results the oracle was checked against."""
    N = pkg._native
    tiles = 8
    d = [dev.DeviceBatch.from_host(b, tile=tiles) for b in host_batches]
    n = d[0].n
    step = dev.ResidentStep(ctx, *d, now=now, cull_idle_min=1440, check_period_min=1, mode=N.CULL_RECONCILE)
    for _ in range(3):
        step.run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        step.run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    ok = True
    for b, w in zip(d, (want["diff_sts"][0], want["diff_svc"][0], want["diff_vs"][0], want["cull"][0])):
        ok &= bool((b.outs[0].cpu().numpy().reshape(tiles, per) == w[None, :]).all())
    if not ok:
        raise SystemExit("bench: 10 M-notebook results do not repeat the checked 1.25 M results tile by tile")
    nbytes = sum(b.algorithmic_bytes for b in d)
    return {"notebooks": n, "ms_per_step": ms, "objects_per_s": n / (ms * 1e-3), "resident_GB": nbytes / 1e9,
            "GBps": nbytes / (ms * 1e-3) / 1e9, "frac_of_peak": nbytes / (ms * 1e-3) / 1e9 / peak,
            "frac_of_nominal_8TBps": nbytes / (ms * 1e-3) / 1e9 / 8000.0,
            "parity": "every one of the 8 tiles repeats the 1.25 M results that were checked against the packed oracle"}


def resident_store(pkg, O, ctx, torch, fleet, per, host_batches, now, want, alg_bytes, peak):
    """The same fleet held in an nbg_store: load, step from the store (spans instead of CSR), then ticks in which
    1 % of the slots are replaced by other notebooks' rows before the step.  Parity after the last tick."""
    host, N = pkg.host, pkg._native
    sts_sp, svc_g, vs_g, cull = host_batches
    store = pkg.engine.Store(ctx, per)
    t0 = time.perf_counter()
    store.put(np.arange(per, dtype=np.uint64), sts_sp, svc_g, vs_g, cull)
    load_s = time.perf_counter() - t0
    words = (per + 63) // 64
    outs = {k: torch.empty(sz, dtype=torch.uint8, device="cuda") for k, sz in
            (("sts_r", per), ("sts_b", words * 8), ("svc_r", per), ("svc_b", words * 8), ("vs_r", per), ("vs_b", words * 8),
             ("act", per), ("amax", per * 4), ("cb", words * 8), ("pb", words * 8))}
    o = N.nbg_step_outputs(*[outs[k].data_ptr() for k in ("sts_r", "sts_b", "svc_r", "svc_b", "vs_r", "vs_b", "act", "amax", "cb", "pb")])
    consts = host.vs_gen_consts()
    p = N.nbg_cull_params(now[0], now[1], 1440, 1, N.CULL_RECONCILE)
    stream = torch.cuda.current_stream().cuda_stream

    def device_step_ms(reps=20):
        for _ in range(3):
            store.step_device(per, consts, p, o, stream)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            store.step_device(per, consts, p, o, stream)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    fresh_ms = device_step_ms()
    exp = {"sts": want["diff_sts"][0].copy(), "svc": want["diff_svc"][0].copy(), "vs": want["diff_vs"][0].copy(),
           "act": want["cull"][0].copy(), "amax": want["cull"][1].copy()}
    # ticks: 1 % of the slots take the rows of notebooks that were not in the fleet so far
    rng = np.random.default_rng(7)
    m = max(per // 100, 1)
    ticks = []
    for t in range(6):
        slots = rng.choice(per, size=m, replace=False).astype(np.uint64)
        alt = (per + t * m + np.arange(m)).astype(np.uint64) % np.uint64(fleet.n)
        a_sts, a_svc, a_vs = fleet.pack(idx=alt)
        a_cf = host.CullFleet(host.SEED_BASE + 900 + t, m)
        a_cull = a_cf.pack()
        rows = (a_sts.sparse(), a_svc.gen(), a_vs.gen(), a_cull)
        t0 = time.perf_counter()
        store.put(slots, *rows)
        put_s = time.perf_counter() - t0
        r = store.step(per, consts, now[0], now[1])                      # host outputs: the tick a controller would run
        ticks.append((put_s, time.perf_counter() - t0))
        exp["sts"][slots] = O.packed_diff_sts_sparse(rows[0])[0]
        exp["svc"][slots] = O.packed_diff_svc(a_svc)[0]
        exp["vs"][slots] = O.packed_diff_vs(a_vs)[0]
        c = O.packed_cull(a_cull, now, 1440, 1, 1)
        exp["act"][slots], exp["amax"][slots] = c[0], c[1]
    ok = (np.array_equal(r["sts"][0], exp["sts"]) and np.array_equal(r["svc"][0], exp["svc"]) and np.array_equal(r["vs"][0], exp["vs"])
          and np.array_equal(r["cull"][0], exp["act"]) and np.array_equal(r["cull"][1], exp["amax"]))
    if not ok:
        raise SystemExit("bench: the resident store differs from the packed oracle after replacements")
    churned_ms = device_step_ms()
    st = store.stats()
    store.close()
    put_ms = statistics.median(x[0] for x in ticks[1:]) * 1e3
    tick_ms = statistics.median(x[1] for x in ticks[1:]) * 1e3
    return {"load_seconds": load_s, "load_GBps": (alg_bytes / load_s) / 1e9,
            "step_ms_fresh": fresh_ms, "step_objects_per_s_fresh": per / (fresh_ms * 1e-3), "step_frac_of_peak_fresh": alg_bytes / (fresh_ms * 1e-3) / 1e9 / peak,
            "step_ms_after_6_ticks": churned_ms, "rows_replaced_per_tick": m,
            "tick_ms": tick_ms, "tick_put_ms": put_ms, "tick_objects_per_s": per / (tick_ms * 1e-3),
            "store": st, "parity": "bit-exact vs packed oracle after the last tick (replaced slots included)",
            "note": "tick = nbg_store_put of 1 % of the rows (host buffers) + nbg_store_step over all 1.25 M slots with host outputs; "
                    "packing the replaced rows is outside the tick (it happens when the informer delivers the object)"}


def other_configs(pkg, dev, ctx, torch, peak):
    """Device-timed numbers for BASELINE configs #2, #3, #4 (parity for these is in tests/)."""
    host, N = pkg.host, pkg._native
    import oracle as O   # the checker of these side measurements

    def O_packed_cull(batch, now):
        return O.packed_cull(batch, now, 1440, 1, 1, threads=8)[0]
    out = {}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def timed(b, reps=20, **kw):
        ms = []
        for _ in range(3):
            b.run(ctx, **kw)
        for _ in range(reps):
            flush.fill_(1)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); b.run(ctx, **kw); e.record()
            torch.cuda.synchronize()
            ms.append(s.elapsed_time(e))
        return statistics.median(ms)

    f2 = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 2, 10_000, 200)
    sts, svc, _ = f2.pack(vs=False)
    ms = timed(dev.DeviceBatch.from_host(sts)) + timed(dev.DeviceBatch.from_host(svc))
    out["config2_10k_sts_svc"] = {"objects_per_s": 10_000 / (ms * 1e-3), "ms": ms, "roofline": "n/a (36 MB, launch-bound; L2 flushed between launches)"}
    # the same 10 k notebooks as a whole step (sparse tiles, device-generated Service), issued directly and
    # replayed from a captured CUDA graph: this size is launch-bound, so the graph is what helps
    cf2 = host.CullFleet(host.SEED_BASE + 2, 10_000)
    small = [dev.DeviceBatch.from_host(b) for b in (sts.sparse(), svc.gen(), cf2.pack())]
    rstep = dev.ResidentStep(ctx, sts=small[0], svc=small[1], cull=small[2], now=(cf2.now_sec, cf2.now_nsec))
    cap_stream = torch.cuda.Stream()
    with torch.cuda.stream(cap_stream):
        for _ in range(3):
            rstep.run()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=cap_stream):
        rstep.run()
    per_step = {}
    for name, fn in (("direct", rstep.run), ("graph_replay", graph.replay)):
        runs = []
        with torch.cuda.stream(cap_stream):
            for _ in range(7):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(50):
                    fn()
                b.record()
                torch.cuda.synchronize()
                runs.append(a.elapsed_time(b) / 50)
        per_step[name] = statistics.median(runs)
    out["config2_10k_step_back_to_back"] = {
        "direct_us_per_step": per_step["direct"] * 1e3, "graph_replay_us_per_step": per_step["graph_replay"] * 1e3,
        "objects_per_s_graph": 10_000 / (per_step["graph_replay"] * 1e-3),
        "note": "StatefulSet + Service + culler over 10 k notebooks, 50 steps back to back, inputs L2-resident (no flush): "
                "launch-bound; nbg_step_device captured once into a CUDA graph and replayed"}
    del small, rstep, graph
    cf = host.CullFleet(host.SEED_BASE + 3, 250_000)
    cb = cf.pack()
    d = dev.DeviceBatch.from_host(cb)
    ms = timed(d, now=(cf.now_sec, cf.now_nsec))
    out["config3_1M_kernel_records"] = {"records_per_s": sum(cb.records()) / (ms * 1e-3), "notebooks_per_s": 250_000 / (ms * 1e-3), "ms": ms,
                                        "GBps": d.algorithmic_bytes / (ms * 1e-3) / 1e9, "roofline": "n/a (22 MB, launch-bound; L2 flushed)"}
    # the same culler batch scaled to ~1 G kernel/terminal records (SURVEY 8d, config #3 note): tiled 1000x on the device
    del d
    big = dev.DeviceBatch.from_host(cb, tile=1000)
    ms_big = timed(big, reps=5, now=(cf.now_sec, cf.now_nsec))
    got = big.outs[0].cpu().numpy().reshape(1000, -1)
    want_c = O_packed_cull(cb, (cf.now_sec, cf.now_nsec))
    assert (got == want_c[None, :]).all(), "1 G-record culler run does not repeat the checked 250 k results"
    out["config3_1G_kernel_records"] = {"records": int(sum(cb.records())) * 1000, "notebooks": big.n, "ms": ms_big,
                                        "records_per_s": sum(cb.records()) * 1000 / (ms_big * 1e-3),
                                        "resident_GB": big.algorithmic_bytes / 1e9, "GBps": big.algorithmic_bytes / (ms_big * 1e-3) / 1e9,
                                        "frac_of_peak": big.algorithmic_bytes / (ms_big * 1e-3) / 1e9 / peak,
                                        "parity": "every one of the 1000 tiles repeats the 250 k results checked against the packed oracle"}
    del big
    torch.cuda.empty_cache()
    f4 = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 4, 100_000, 2000)
    sts, _, vs = f4.pack(svc=False)
    d1, d2 = dev.DeviceBatch.from_host(sts), dev.DeviceBatch.from_host(vs)
    m1, m2 = timed(d1), timed(d2)
    out["config4_100k_odh_sts_vs"] = {"objects_per_s": 100_000 / ((m1 + m2) * 1e-3), "ms": m1 + m2,
                                      "sts_GBps": d1.algorithmic_bytes / (m1 * 1e-3) / 1e9, "sts_frac_of_peak": d1.algorithmic_bytes / (m1 * 1e-3) / 1e9 / peak}
    # the owner-reference join (NBCTL:157-170) as a hash join, 1.25 M notebooks
    lib = N.nbg()
    n = 1_250_000
    fj = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 6, n, n // 50)
    nb_keys, sts_keys, _ = fj.join_lists(n)
    d_nb, d_st = torch.from_numpy(nb_keys.view(np.int64)).cuda(), torch.from_numpy(sts_keys.view(np.int64)).cuda()
    d_match = torch.empty(n, dtype=torch.int32, device="cuda")
    need = lib.nbg_owner_join_scratch_bytes(len(sts_keys))
    scratch = torch.empty(need, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    def join():
        rc = lib.nbg_owner_join_device(ctx.handle, 0, d_nb.data_ptr(), n, d_st.data_ptr(), len(sts_keys), d_match.data_ptr(),
                                       scratch.data_ptr(), need, stream)
        assert rc == 0
    ms = []
    for it in range(13):
        flush.fill_(it)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); join(); e.record()
        torch.cuda.synchronize()
        if it >= 3:
            ms.append(s.elapsed_time(e))
    want, cpu_s = O.packed_owner_join(nb_keys, sts_keys)
    assert np.array_equal(d_match.cpu().numpy().view(np.uint32), want), "owner join differs from the oracle"
    med = statistics.median(ms)
    stream_bytes = 8 * n + 8 * len(sts_keys) + 4 * n
    out["owner_join_1.25M"] = {"notebooks_per_s": n / (med * 1e-3), "ms": med, "statefulsets": int(len(sts_keys)),
                               "algorithmic_GBps": stream_bytes / (med * 1e-3) / 1e9,
                               "table_bytes": int(need), "note": "hash table (scratch) is L2-resident; random access, not a streaming roofline",
                               "cpu_packed_1thread_notebooks_per_s": n / cpu_s, "parity": "bit-exact vs packed oracle"}
    # ODH comparator family and the webhook restart-block predicate, ~1 M pairs (tiled 64x on device)
    fo = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 11, 16384, 400)
    objs = fo.objects(0, 16384)
    gb = host.GenBatch(N.GEN_DEEPEQUAL, "metadata.labels", "spec.template.spec")
    rb = host.RestartBatch()
    for o in objs:
        gb.add(o["sts_from"], o["sts_to"])
        mutated = o["notebook"].clone()
        host.odh_mutate(mutated, 2 | 4 | 8 | 16)
        rb.add(o["notebook"], o["notebook"], mutated)
    for name, hb, want in (("odh_comparators_1M", gb, O.packed_diff_generic(gb)[0]), ("restart_block_1M", rb, O.packed_restart_block(rb)[0])):
        db = dev.DeviceBatch.from_host(hb, tile=64)
        ms = timed(db)
        got = db.outs[0].cpu().numpy().reshape(64, -1)
        assert (got == want[None, :]).all(), name + " differs from the oracle"
        out[name] = {"pairs_per_s": db.n / (ms * 1e-3), "ms": ms, "pairs": db.n, "GBps": db.algorithmic_bytes / (ms * 1e-3) / 1e9,
                     "frac_of_peak": db.algorithmic_bytes / (ms * 1e-3) / 1e9 / peak, "bytes_per_pair": round(db.algorithmic_bytes / db.n, 1),
                     "parity": "bit-exact vs packed oracle (16384 distinct pairs tiled 64x)"}
        del db
    # RFC3339 parse + "idle" compare of kernel records on the device
    import json as _json
    cf2 = host.CullFleet(host.SEED_BASE + 3, 250_000)
    la, st = [], []
    for i in range(0, 250_000):
        nb = _json.loads(cf2.json(i))
        for k in nb["kernels"]:
            la.append(k["last_activity"]); st.append(k["execution_state"])
    ta, ts = pkg.engine.StringTable(la), pkg.engine.StringTable(st)
    R = ta.n
    dv = {}
    for name, t in (("a", ta), ("s", ts)):
        dv[name] = (torch.from_numpy(t.off.view(np.int32)).cuda(), torch.from_numpy(t.bytes.copy()).cuda())
    va = N.nbg_strings(C.cast(dv["a"][0].data_ptr(), N.u32p), C.cast(dv["a"][1].data_ptr(), N.u8p))
    vs_ = N.nbg_strings(C.cast(dv["s"][0].data_ptr(), N.u32p), C.cast(dv["s"][1].data_ptr(), N.u8p))
    d_sec = torch.zeros(R, dtype=torch.int64, device="cuda")
    d_nsf = torch.zeros(R, dtype=torch.int32, device="cuda")
    d_zone = torch.zeros(R, dtype=torch.int32, device="cuda")

    def parse():
        assert lib.nbg_parse_records_device(ctx.handle, 0, C.byref(va), C.byref(vs_), R, d_sec.data_ptr(), d_nsf.data_ptr(),
                                            d_zone.data_ptr(), stream) == 0
    ms = []
    for it in range(13):
        flush.fill_(it)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); parse(); e.record()
        torch.cuda.synchronize()
        if it >= 3:
            ms.append(s.elapsed_time(e))
    med = statistics.median(ms)
    # CPU baseline: the host parser (C++, all threads) on the same tables
    h_sec, h_nsf, h_zone = np.zeros(R, np.int64), np.zeros(R, np.uint32), np.zeros(R, np.int32)
    nbh = N.nbh()
    best = 1e9
    for _ in range(5):
        t0 = time.perf_counter()
        nbh.nbh_parse_records(ta.off.ctypes.data_as(N.u32p), ta.bytes.ctypes.data_as(N.u8p), ts.off.ctypes.data_as(N.u32p),
                              ts.bytes.ctypes.data_as(N.u8p), R, h_sec.ctypes.data_as(C.c_void_p), h_nsf.ctypes.data_as(C.c_void_p),
                              h_zone.ctypes.data_as(C.c_void_p))
        best = min(best, time.perf_counter() - t0)
    assert np.array_equal(d_sec.cpu().numpy(), h_sec) and np.array_equal(d_nsf.cpu().numpy().view(np.uint32), h_nsf) and \
        np.array_equal(d_zone.cpu().numpy(), h_zone), "device RFC3339 parse differs from the host parser"
    text_bytes = int(ta.off[-1]) + int(ts.off[-1]) + 8 * (R + 1) + 16 * R
    out["rfc3339_parse_1M_records"] = {"records_per_s": R / (med * 1e-3), "ms": med, "records": R,
                                       "algorithmic_GBps": text_bytes / (med * 1e-3) / 1e9, "frac_of_peak": text_bytes / (med * 1e-3) / 1e9 / peak,
                                       "bytes_per_record": round(text_bytes / R, 1),
                                       "cpu_host_parser_records_per_s": R / best, "cpu_threads": nbh.nbh_threads(),
                                       "parity": "bit-exact vs the host parser on all records (tests: vs both CPU parsers + fuzz)"}
    # the culler end to end from the strings the reference handles: device parse + cull in one call,
    # against host parse (all threads) + nbg_cull on the packed words
    nbs = [_json.loads(cf2.json(i)) for i in range(250_000)]
    tb = pkg.engine.CullTextBatch(nbs)
    for _ in range(2):
        r_text = ctx.cull_text(tb, cf2.now_sec, cf2.now_nsec, 1440, 1, 1)
    t0 = time.perf_counter()
    for _ in range(5):
        r_text = ctx.cull_text(tb, cf2.now_sec, cf2.now_nsec, 1440, 1, 1)
    text_s = (time.perf_counter() - t0) / 5
    packed = cf2.pack()
    r_packed = ctx.cull(packed, cf2.now_sec, cf2.now_nsec, 1440, 1, 1)
    assert np.array_equal(r_text[0], r_packed[0]) and np.array_equal(r_text[1], r_packed[1])
    # host parse cost of the same strings (kernels + terminals + 2 annotations), then the packed call
    tabs = tb.tables
    t0 = time.perf_counter()
    for t_ in (tabs[0], tabs[2], tabs[3], tabs[4]):
        hs, hn = np.zeros(t_.n, np.int64), np.zeros(t_.n, np.uint32)
        nbh.nbh_parse_records(t_.off.ctypes.data_as(N.u32p), t_.bytes.ctypes.data_as(N.u8p), None, None, t_.n,
                              hs.ctypes.data_as(C.c_void_p), hn.ctypes.data_as(C.c_void_p), None)
    host_parse_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    for _ in range(5):
        ctx.cull(packed, cf2.now_sec, cf2.now_nsec, 1440, 1, 1)
    packed_s = (time.perf_counter() - t0) / 5
    out["cull_from_strings_250k"] = {"notebooks_per_s": 250_000 / text_s, "ms": text_s * 1e3, "h2d_bytes": int(tb.input_bytes()),
                                     "host_parse_plus_packed_call_notebooks_per_s": 250_000 / (host_parse_s + packed_s),
                                     "host_parse_ms": host_parse_s * 1e3, "packed_call_ms": packed_s * 1e3,
                                     "parity": "bit-exact vs nbg_cull on host-parsed words"}
    return out


if __name__ == "__main__":
    main()
