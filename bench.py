#!/usr/bin/env python
"""bench.py — the graft's one headline measurement (contract: README of the task).

Workload: BASELINE config #5 as stated — 10 M DISTINCT synthetic Notebooks in 200 k Zipf(1.1) namespaces.  A "step"
is one pass of the hot path over the whole fleet: diff_sts + diff_svc + diff_vs (Copy*Fields, COMMON:107-219) +
cull (CULL:86-241, 315-428).  Unit: objects/s where one object = one Notebook (its StatefulSet pair, Service pair,
VirtualService pair and culler record set).  One GPU holds the whole fleet; N GPUs split the SAME fleet by
namespace (NBCTL:159), at the sizes the partition gives (strong scaling): headline with the size-aware table of
nbg_balance_shards, and beside it the stateless nbg_shard_of hash.

  value      fleet / slowest rank's step time; inputs resident in HBM, CUDA events on the launch stream
  e2e        the same step through the host-buffer C ABI (nbg_step): pinned host buffers -> H2D -> kernels -> D2H
             inside the timed region, max over ranks
  roofline   dominant kernel (diff_sts_sparse): algorithmic bytes / its own CUDA-event time inside the step;
             frac_by_traffic uses the DRAM bytes ncu saw for that launch (profiles/r2_traffic.json)
  parity     L2: every object of every shard against the packed oracle; L1: a sample of the same fleet
             (100 k notebooks at N=1) against the object-level port of the Go predicates — the bench fails on a difference
  cpu_baseline  that port (CPU restatement of the Go predicates on whole objects) on all host threads, on the
             same sample; like_for_like: the two ratios that start both sides from the same place

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
    ap.add_argument("--notebooks", type=int, default=10_000_000, help="notebooks in the fleet (BASELINE config #5: 10 M)")
    ap.add_argument("--cpu-sample", type=int, default=100_000, help="notebooks in the CPU baseline / L1 parity sample")
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary configs (#2, #3, #4)")
    ap.add_argument("--no-numa", action="store_true", help="multi-rank: do not bind each rank to the CPUs next to its GPU")
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
    walk, RFC3339 strings) on `threads` threads over a bounded sample of the SAME fleet the GPU arm
    runs: notebooks `idx` of `fleet` / `cfleet`.  Its verdicts are kept (self.out) so that the caller
    can compare them with the GPU's bytes for the same notebooks."""

    def __init__(self, pkg, O, fleet, cfleet, idx, threads):
        self.idx = np.ascontiguousarray(idx, dtype=np.uint64)
        self.sample, self.threads, self.L = len(self.idx), threads, O.lib()
        self.cf = cfleet
        sample = self.sample
        self.objs = fleet.objects(idx=self.idx)
        vp = C.c_void_p
        self.arrs = []
        for a, b in (("sts_from", "sts_to"), ("svc_from", "svc_to"), ("vs_from", "vs_to")):
            self.arrs.append(((vp * sample)(*[o[a]._h.value for o in self.objs]), (vp * sample)(*[o[b]._h.value for o in self.objs])))
        self.texts = [cfleet.json(int(i)).encode() for i in self.idx]
        self.jarr = (C.c_char_p * sample)(*self.texts)
        self.out = {k: np.zeros(sample, np.uint8) for k in ("sts", "svc", "vs", "act")}
        self.out["amax"] = np.zeros(sample, np.uint32)

    def step(self, diff_only=False):
        """One pass of all four predicates over the sample; seconds inside the oracle loops (the
        culler's JSON texts are decoded before its clock starts)."""
        t = 0.0
        for kind, key in enumerate(("sts", "svc", "vs")):
            t += self.L.orc_bench_pairs(kind, self.arrs[kind][0], self.arrs[kind][1], self.sample, self.threads,
                                        self.out[key].ctypes.data_as(C.c_void_p))
        if diff_only:
            return t
        tc = self.L.orc_bench_cull(self.jarr, self.sample, self.cf.now_sec, self.cf.now_nsec, 1440, 1, 1, self.threads,
                                   self.out["act"].ctypes.data_as(C.c_void_p), self.out["amax"].ctypes.data_as(C.c_void_p))
        if tc < 0:
            raise SystemExit("bench: the oracle could not decode a culler input")
        return t + tc

    def rate(self, min_cpu_seconds=10.0, max_reps=64):
        total, reps = 0.0, 0
        while reps < max_reps and (total * self.threads < min_cpu_seconds or reps < 2):
            total += self.step()
            reps += 1
        return self.sample * reps / total, {"sample_notebooks": self.sample, "repetitions": reps,
                                            "cpu_seconds": round(total * self.threads, 2)}


def sample_indices(total, want):
    """`want` notebooks spread evenly over the fleet."""
    want = max(1, min(want, total))
    return (np.arange(want, dtype=np.uint64) * np.uint64(total // want)).astype(np.uint64)


def run_reference(args, rank, world):
    """--impl reference: the reference's algorithm on the host cores (the Go code itself
    cannot be built in this image; the oracle port stands in, kind = "port").  Same fleet as the GPU arm,
    bounded sample of it per step; a second, four times smaller sample shows the rate does not depend on the size."""
    if rank != 0:
        return
    pkg = importlib.import_module("odh-kubeflow_b200")
    import oracle as O
    host = pkg.host
    threads = os.cpu_count() or 1
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, args.notebooks, max(200, args.notebooks // 50))
    cfleet = host.CullFleet(host.SEED_BASE + 5, args.notebooks)
    ref = CpuReference(pkg, O, fleet, cfleet, sample_indices(args.notebooks, args.cpu_sample), threads)
    secs = []
    t0 = time.time()
    for s in range(args.warmup + args.steps):
        t = ref.step()
        if s >= args.warmup:
            secs.append(t)
        if time.time() - t0 > 200 and len(secs) >= 1:
            break
    value = ref.sample / statistics.median(secs)
    small = CpuReference(pkg, O, fleet, cfleet, sample_indices(args.notebooks, max(args.cpu_sample // 4, 1)), threads)
    small_secs = [small.step() for _ in range(5)][1:]
    small_value = small.sample / statistics.median(small_secs)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": len(secs),
            "warmup": args.warmup, "ms_per_step": 1000.0 * statistics.median(secs), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "config #5: %d-notebook sample per step of the %d-notebook fleet the GPU arm runs, all 4 predicates per notebook"
                                   % (ref.sample, args.notebooks),
                       "note": "CPU port of the Go predicates (semantic oracle) on whole objects; Go itself unavailable on this box; "
                               "the culler's JSON is decoded outside the timed loops (as CULL:276-313 is outside C1-C9)"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "%d notebooks per step, object-level DeepEqual + RFC3339 strings" % ref.sample,
                             "other_sample": {"notebooks": small.sample, "value": small_value,
                                              "note": "same port, a four times smaller sample: the rate does not depend on the sample size"}},
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


class Shard:
    """One rank's part of the fleet: packed on the host (lean: sparse PodSpec tiles only), resident on the device."""

    def __init__(self, host, dev, fleet, cfleet, idx, local):
        t0 = time.time()
        self.idx = idx
        self.sts, self.svc, self.vs = fleet.pack(idx=idx, lean=True) if idx is not None else fleet.pack(lean=True)
        self.make_cpu_s, self.pack_cpu_s = fleet.pack_timing()
        self.cull = cfleet.pack(idx=idx) if idx is not None else cfleet.pack()
        self.pack_wall_s = time.time() - t0
        self.n = len(self.cull)
        self.host_batches = (self.sts.sparse(), self.svc.gen(), self.vs.gen(), self.cull)
        self.dev_batches = [dev.DeviceBatch.from_host(b, device=local) for b in self.host_batches]
        self.named = list(zip(("diff_sts", "diff_svc", "diff_vs", "cull"), self.dev_batches))
        self.alg_bytes = {name: b.algorithmic_bytes for name, b in self.named}

    def want(self, O, now, threads):
        """The packed oracle's answers for every object of the shard."""
        sp, _, _, cull = self.host_batches
        return {"diff_sts": O.packed_diff_sts_sparse(sp, threads=threads), "diff_svc": O.packed_diff_svc(self.svc, threads=threads),
                "diff_vs": O.packed_diff_vs(self.vs, threads=threads), "cull": O.packed_cull(cull, now, 1440, 1, 1, threads=threads)}

    def check(self, want):
        ok = True
        for name, b in self.named:
            outs = [o.cpu().numpy() for o in b.outs]
            ok &= bool(np.array_equal(outs[0], want[name][0]))
            ok &= bool(np.array_equal(outs[-1].view(np.uint64), want[name][3 if name == "cull" else 1]))
            if name == "cull":
                ok &= bool(np.array_equal(outs[1].view(np.uint32), want[name][1]))
        return ok


def timed_resident_steps(torch, dist, world, resident, steps, warmup):
    """W warm-up steps, then exactly K nbg_step_device calls between two events on the launch stream, bracketed by
    barrier + synchronize on both sides.  Returns (ms for the K steps on this rank, mean ms of the StatefulSet launch)."""
    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    for _ in range(max(warmup, 3)):
        resident.run()
    barrier()
    resident.sts_ms()                                      # drop the warm-up pairs
    resident.reserve(steps)
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(steps):
        resident.run()
    stop.record()
    barrier()
    dom = resident.sts_ms()
    return start.elapsed_time(stop), (statistics.mean(dom) if dom else None)


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
    host_threads = int(os.environ["NBH_THREADS"])
    ctx = pkg.Context([local])
    lib = N.nbg()

    # ---- the fleet: BASELINE config #5, `total` DISTINCT notebooks in Zipf(1.1) namespaces.  One GPU takes all of
    # it.  N GPUs split it by namespace (NBCTL:159: the namespace is the unit): the headline run uses the size-aware
    # table of nbg_balance_shards, a second run the stateless nbg_shard_of hash; both at the sizes the partition
    # gives (strong scaling: the step ends when the largest shard does).
    total = args.notebooks
    fleet = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 5, total, max(200, total // 50))
    cfleet = host.CullFleet(host.SEED_BASE + 5, total)
    now = (cfleet.now_sec, cfleet.now_nsec)
    vs_consts = host.vs_gen_consts()
    tables = {}
    if world > 1:
        for kind in ("balanced", "hash"):
            table, load = fleet.shard_table(world, balanced=(kind == "balanced"))
            tables[kind] = (table, load)
        idx = fleet.shard_indices_by_table(tables["balanced"][0], rank)
        assert len(idx) == int(tables["balanced"][1][rank])
    else:
        idx = None

    def all_max(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    def all_gather_floats(x):
        t = torch.zeros(world, dtype=torch.float64, device="cuda")
        t[rank] = x
        if world > 1:
            dist.all_reduce(t)
        return [float(v) for v in t.cpu()]

    shard = Shard(host, dev, fleet, cfleet, idx, local)
    per = shard.n
    d_sts, d_svc, d_vs, d_cull = shard.dev_batches
    batches, alg_bytes = shard.named, shard.alg_bytes
    sts_sp, svc_g, vs_g, cull = shard.host_batches

    def launch(name, b):
        if name == "cull":
            return b.run(ctx, now=now, cull_idle_min=1440, check_period_min=1, mode=N.CULL_RECONCILE)
        return b.run(ctx)

    # ---- value: inputs resident in HBM.  One step = one nbg_step_device call: the StatefulSet kernel on the
    # main stream, the three small latency-bound kernels beside it on the context's side streams (the
    # library forks and joins them with events).
    resident = dev.ResidentStep(ctx, d_sts, d_svc, d_vs, d_cull, now=now, cull_idle_min=1440, check_period_min=1,
                                mode=N.CULL_RECONCILE, time_sts=True)
    sampler = ClockSampler(local)
    launches0 = ctx.launch_count()
    ms_total, dom_ms = timed_resident_steps(torch, dist, world, resident, args.steps, args.warmup)
    gpu_launches = ctx.launch_count() - launches0 - max(args.warmup, 3) * resident.launches
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
    ms_step = all_max(ms_total) / args.steps
    rank_ms = all_gather_floats(ms_total / args.steps)
    rank_n = all_gather_floats(float(per))
    value = total / (ms_step / 1000.0)

    # ---- parity of what was just computed (outside the timed region): every object of the shard against the
    # packed oracle (L2)
    cpu_threads = max(1, (os.cpu_count() or 1) // world)
    want = shard.want(O, now, cpu_threads)
    if not shard.check(want):
        raise SystemExit("bench: device results differ from the packed oracle")

    # ---- e2e: the host-buffer C ABI with pinned host memory, copies inside the timed region
    keep = []
    views = {}
    words = (per + 63) // 64
    for kind, b in (("sts", sts_sp), ("svc", svc_g), ("vs", vs_g), ("cull", cull)):
        views[kind] = (struct_over(N, kind, pinned_copy(lib, b.arrays(), keep)), sum(a.nbytes for a in b.arrays().values()))
    out_arr = pinned_copy(lib, {"r": np.zeros(per, np.uint8), "b": np.zeros(words, np.uint64), "a": np.zeros(per, np.uint32),
                                "b2": np.zeros(words, np.uint64)}, keep)
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

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    e2e_steps = max(3, min(args.steps, 5))
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    clocks = sampler.stop()  # sampled from the start of the device-timed steps to the end of the e2e steps
    e2e_max = all_max(e2e_s)
    e2e_value = total / e2e_max
    h2d = sum(v[1] for v in views.values())
    d2h = 3 * (per + words * 8) + (per + 4 * per + 2 * words * 8)  # 3x (reason + bitmap); action + argmax + 2 bitmaps
    job_h2d, job_d2h = sum(all_gather_floats(float(h2d))), sum(all_gather_floats(float(d2h)))
    assert np.array_equal(out_arr["r"], want["cull"][0]) and np.array_equal(out_arr["a"], want["cull"][1]), \
        "e2e cull results differ from the oracle"
    assert np.array_equal(o_sts["r"], want["diff_sts"][0]) and np.array_equal(o_svc["r"], want["diff_svc"][0]) and \
        np.array_equal(o_vs["r"], want["diff_vs"][0]), "e2e diff results differ from the oracle"
    sts_reason_host = o_sts["r"].copy()
    for p in keep:
        lib.nbg_free_pinned(p)
    keep = []

    # ---- gather (outside every timed region): the shards' StatefulSet verdicts come together on rank 0 through the
    # per-shard index maps, as a sharded controller would assemble its fleet-wide "needs update" list
    gathered = None
    if world > 1:
        cap = int(max(rank_n))
        idx_t = torch.full((cap,), -1, dtype=torch.int64, device="cuda")
        idx_t[:per] = torch.from_numpy(idx.astype(np.int64)).cuda()
        upd_t = torch.zeros(cap, dtype=torch.uint8, device="cuda")
        upd_t[:per] = torch.from_numpy(((sts_reason_host & 0x0F) != 0).astype(np.uint8)).cuda()
        all_idx = [torch.empty_like(idx_t) for _ in range(world)]
        all_upd = [torch.empty_like(upd_t) for _ in range(world)]
        dist.all_gather(all_idx, idx_t)            # plumbing only: 9 bytes per notebook, once, after the measurements
        dist.all_gather(all_upd, upd_t)
        if rank == 0:
            fleet_bits = np.zeros(total, dtype=np.uint8)
            seen = np.zeros(total, dtype=np.uint8)
            for gi, gu in zip(all_idx, all_upd):
                gi, gu = gi.cpu().numpy(), gu.cpu().numpy()
                keep_ = gi >= 0
                seen[gi[keep_]] += 1
                fleet_bits[gi[keep_]] = gu[keep_]
            gathered = {"notebooks": int(seen.sum()), "each_notebook_in_exactly_one_shard": bool((seen == 1).all()),
                        "statefulsets_needing_update": int(fleet_bits.sum())}

    # ---- the same fleet cut by the stateless hash (nbg_shard_of): device-resident step only
    hash_run = None
    if world > 1:
        host_pack_first = (shard.make_cpu_s, shard.pack_cpu_s, shard.pack_wall_s)
        del resident, d_sts, d_svc, d_vs, d_cull, batches, sts_sp, svc_g, vs_g, cull, views
        shard_first_alg = dict(alg_bytes)
        shard = None
        torch.cuda.empty_cache()
        idx_h = fleet.shard_indices_by_table(tables["hash"][0], rank)
        shard_h = Shard(host, dev, fleet, cfleet, idx_h, local)
        res_h = dev.ResidentStep(ctx, *shard_h.dev_batches, now=now, cull_idle_min=1440, check_period_min=1, mode=N.CULL_RECONCILE)
        ms_h, _ = timed_resident_steps(torch, dist, world, res_h, args.steps, args.warmup)
        if not shard_h.check(shard_h.want(O, now, cpu_threads)):
            raise SystemExit("bench: device results differ from the packed oracle (hash shards)")
        ms_h_step = all_max(ms_h) / args.steps
        hash_run = {"value": total / (ms_h_step / 1000.0), "ms_per_step": ms_h_step,
                    "shard_notebooks": [int(x) for x in all_gather_floats(float(shard_h.n))],
                    "shard_ms": all_gather_floats(ms_h / args.steps),
                    "shard_imbalance_max_over_mean": float(tables["hash"][1].max() / tables["hash"][1].mean()),
                    "parity": "bit-exact vs packed oracle on every object of every shard",
                    "note": "namespace -> shard by nbg_shard_of(hash(namespace)): stateless, but a Zipf fleet lands unevenly and "
                            "the step ends when the largest shard does"}
        del res_h, shard_h
        torch.cuda.empty_cache()
    else:
        host_pack_first = (shard.make_cpu_s, shard.pack_cpu_s, shard.pack_wall_s)
        shard_first_alg = dict(alg_bytes)

    # ---- the single-process multi-device context (nbg_init with every local device), checked once per run
    multi_device = None
    if world > 1 and rank == 0:
        multi_device = multi_device_context_check(pkg, O, world)
    if world > 1:
        dist.barrier()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    traffic = None
    try:  # per-launch DRAM bytes of the dominant kernel from the committed ncu capture (same launch size only)
        tr = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))["diff_sts_sparse_kernel"]
        if tr["objects_per_launch"] == per:
            traffic = tr["dram_bytes_read"] + tr["dram_bytes_write"]
    except Exception:
        pass
    alg_bytes = shard_first_alg
    make_cpu_s, pack_cpu_s, pack_wall_s = host_pack_first
    sts_gbs = alg_bytes["diff_sts"] / (dom_ms * 1e-3) / 1e9
    iso_gbs = alg_bytes["diff_sts"] / (kernel_ms["diff_sts"] * 1e-3) / 1e9
    step_gbs = sum(alg_bytes.values()) / (rank_ms[0] * 1e-3) / 1e9          # rank 0's bytes / rank 0's step time
    names = ("diff_sts", "diff_svc", "diff_vs", "cull")
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": {"workload": "config #5: %d distinct synthetic Notebooks in %d Zipf(1.1) namespaces, StatefulSet+Service+VirtualService "
                               "diff + culler decision per Notebook%s" % (total, fleet.n_namespaces,
                                                                         "" if world == 1 else ", split by namespace over %d GPUs" % world),
                   "scaling_note": "strong: the fleet is the same %d notebooks at every N" % total,
                   "object": "one Notebook = 3 desired/live pairs + its kernel/terminal records",
                   "tiles": "PodSpec tiles stored sparsely (192-bit mask + non-zero slots)",
                   "desired_side": "StatefulSet: packed from generateStatefulSet's output; Service and VirtualService: generated on the device "
                                   "from name/port/prefix/rewrite/host words (same answer as shipping generateService/VirtualService outputs)",
                   "sharding": "one GPU holds the whole fleet" if world == 1 else
                               "whole namespaces per shard, size-aware table (nbg_balance_shards), no collective; shards at the sizes "
                               "the partition gives; value = fleet / slowest shard's step time",
                   "shard_notebooks": [int(x) for x in rank_n], "shard_ms": rank_ms,
                   "shard_imbalance_max_over_mean": max(rank_n) / (sum(rank_n) / len(rank_n)),
                   "largest_namespace_over_mean_shard": (float(fleet.ns_table()[1].max()) * world / total),
                   "hash_shards": hash_run, "gathered_on_rank0": gathered, "multi_device_context": multi_device,
                   "cpu_binding": ("rank 0 bound to the %d CPUs NVML lists next to its GPU" % len(near)) if near else "none",
                   "streams": "one nbg_step_device call per step: StatefulSet kernel on the main stream, the three small kernels "
                              "on the context's side streams (event fork/join inside the library); per-kernel times are from a separate serialised pass",
                   "serial_sum_ms": sum(kernel_ms.values()),
                   "l2": "inputs (%.1f GB on rank 0) exceed L2; no flush needed" % (sum(alg_bytes.values()) / 1e9),
                   "bytes_per_object": {k: round(v / per, 1) for k, v in alg_bytes.items()}},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(job_h2d), "d2h_bytes_per_step": int(job_d2h),
                "ms_per_step": e2e_max * 1e3, "effective_h2d_GBps_rank0": h2d / e2e_s / 1e9, "steps": e2e_steps,
                "bytes_note": "whole job: summed over the ranks' shards",
                "timer": "host wall clock around the synchronous C-ABI call nbg_step, max over ranks"},
        "gpu_launches": int(gpu_launches),
        "roofline": {"bound": "hbm", "kernel": "diff_sts_sparse_kernel", "achieved": sts_gbs, "peak": peak, "unit": "GB/s",
                     "frac": sts_gbs / peak, "frac_of_nominal_8TBps": sts_gbs / 8000.0, "traffic": traffic,
                     "frac_by_traffic": (traffic / (dom_ms * 1e-3) / 1e9 / peak) if traffic else None, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": int(alg_bytes["diff_sts"]), "avg_launch_ms": dom_ms,
                     "note": "timed inside the timed region on its own stream, while the three small kernels run concurrently on side "
                             "streams and share HBM with it; algorithmic bytes count every input array once, including value runs of "
                             "objects whose masks differ (which the kernel skips): frac_by_traffic uses the bytes ncu saw instead",
                     "isolated": {"achieved": iso_gbs, "frac": iso_gbs / peak, "avg_launch_ms": kernel_ms["diff_sts"],
                                  "note": "same launch alone on the device (serialised pass after the timed region)"},
                     "whole_step": {"achieved": step_gbs, "frac": step_gbs / peak,
                                    "note": "algorithmic bytes of all four kernels / step time, rank 0"}},
        "kernels": {name: {"avg_ms": kernel_ms[name], "GBps": alg_bytes[name] / (kernel_ms[name] * 1e-3) / 1e9,
                           "frac_of_peak": alg_bytes[name] / (kernel_ms[name] * 1e-3) / 1e9 / peak} for name in names},
        "parity": {"l2": "bit-exact vs packed oracle on all %d objects x 4 kernels of every shard (device path and e2e path)" % total},
        "host_pack": {"objects_per_s": per * host_threads / pack_cpu_s, "threads": host_threads,
                      "us_per_object_per_thread": pack_cpu_s / per * 1e6,
                      "note": "packing three desired/live pairs per Notebook into the batch arrays right after the objects were built "
                              "(cache-hot), outside the timed region; rate = objects x threads / summed thread-seconds; "
                              "from_objects below packs objects that sit in memory",
                      "synthesis_us_per_object_per_thread": make_cpu_s / per * 1e6, "wall_seconds_with_synthesis": pack_wall_s},
    }
    # ---- CPU legs on rank 0: the port of the Go predicates on whole objects (L1) over a sample of the SAME fleet.  Its
    # verdicts are compared with the GPU's bytes for the same notebooks: this is the check that ties packer + kernel to
    # the Go semantics.  world > 1: rank 0's shard only supplies the GPU bytes, so the sample is drawn from it.
    threads = os.cpu_count() or 1
    pool = idx if idx is not None else None
    want_n = args.cpu_sample if world == 1 else min(args.cpu_sample, 20_000)
    sidx = sample_indices(total, want_n) if pool is None else pool[sample_indices(len(pool), want_n)]
    cpu_ref = CpuReference(pkg, O, fleet, cfleet, sidx, threads)
    rate, info = cpu_ref.rate(min_cpu_seconds=10.0 if world == 1 else 2.0)
    l1 = dict(cpu_ref.out)
    # the GPU path on exactly these notebooks, from the objects the port walked: pack -> nbg_step (host buffers)
    rows = host.object_rows(cpu_ref.objs)
    t0 = time.perf_counter()
    p_sts, p_svc, p_vs = host.pack_objects(rows, lean=True)
    t_pack = time.perf_counter() - t0
    p_cull = cfleet.pack(idx=sidx)
    r = ctx.step(p_sts.sparse(), p_svc.gen(), p_vs.gen(), vs_consts, p_cull, now[0], now[1])     # warm-up + results
    t0 = time.perf_counter()
    r3 = ctx.step3(p_sts.sparse(), p_svc.gen(), p_vs.gen(), vs_consts)
    t_step3 = time.perf_counter() - t0
    l1_ok = (np.array_equal(r["sts"][0], l1["sts"]) and np.array_equal(r["svc"][0], l1["svc"]) and np.array_equal(r["vs"][0], l1["vs"])
             and np.array_equal(r["cull"][0], l1["act"]) and np.array_equal(r["cull"][1], l1["amax"])
             and np.array_equal(r3["sts"][0], l1["sts"]))
    if not l1_ok:
        raise SystemExit("bench: GPU results differ from the object-level port of the Go predicates (L1)")
    line["parity"]["l1"] = "bit-exact vs the object-level port of the Go predicates on %d notebooks x 4 predicates (reason bytes, action bytes, argmax)" % cpu_ref.sample
    line["parity"]["l1_objects_checked"] = int(cpu_ref.sample)
    if world == 1:
        t_diff = statistics.median(cpu_ref.step(diff_only=True) for _ in range(3))
        cpu_ref.threads = 1                                   # the same port on ONE thread: the per-object cost (SURVEY 8d ii)
        t1 = cpu_ref.step()
        rate_1 = cpu_ref.sample / t1
        small = CpuReference(pkg, O, fleet, cfleet, sample_indices(total, max(args.cpu_sample // 4, 1)), threads)
        small_rate = small.sample / statistics.median(small.step() for _ in range(3))
        del small
        # the packed loops (the CPU version of OUR algorithm) on a 1.25 M-notebook slice of the packed fleet
        sl = min(total, 1_250_000)
        s_sts, s_svc, s_vs = fleet.pack(0, sl, lean=True)
        s_cull = cfleet.pack(0, sl)
        def packed(th):
            return sum(x[-1] for x in (O.packed_diff_sts_sparse(s_sts.sparse(), threads=th), O.packed_diff_svc(s_svc, threads=th),
                                       O.packed_diff_vs(s_vs, threads=th), O.packed_cull(s_cull, now, 1440, 1, 1, threads=th)))
        t_packed, t_packed_1 = packed(threads), packed(1)
        del s_sts, s_svc, s_vs, s_cull
        line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": "%(sample_notebooks)d notebooks of the same fleet x %(repetitions)d repetitions (%(cpu_seconds)s CPU-s): "
                                          "object-level CopyStatefulSetFields/ServiceFields/VirtualService + culler on RFC3339 strings "
                                          "(JSON decoded outside the timed loops)" % info,
                                "single_thread_value": rate_1, "single_thread_us_per_object": 1e6 / rate_1,
                                "other_sample": {"notebooks": int(max(args.cpu_sample // 4, 1)), "value": small_rate},
                                "packed_loops_value": sl / t_packed, "packed_loops_single_thread_value": sl / t_packed_1,
                                "packed_loops_note": "plain C loops over the same SoA batches, %d objects, %d threads" % (sl, threads)}
        # like-for-like ratios: same starting point on both sides
        line["like_for_like"] = {
            "e2e_from_packed_vs_cpu_packed": {"gpu_e2e_objects_per_s": e2e_value, "cpu_packed_loops_objects_per_s": sl / t_packed,
                                              "ratio": e2e_value / (sl / t_packed),
                                              "note": "both start from the packed SoA buffers in host memory"},
            "e2e_from_objects": {"notebooks": int(cpu_ref.sample), "pack_seconds": t_pack, "step_seconds": t_step3,
                                 "gpu_objects_per_s": cpu_ref.sample / (t_pack + t_step3),
                                 "cpu_port_objects_per_s": cpu_ref.sample / t_diff,
                                 "ratio": (cpu_ref.sample / (t_pack + t_step3)) / (cpu_ref.sample / t_diff),
                                 "note": "both start from the same object trees in host memory; the three Copy*Fields predicates only "
                                         "(GPU side: nbh_pack_objects on %d threads + nbg_step with host buffers; CPU side: the port's "
                                         "DeepEqual walk on %d threads).  Packing is a tree walk like the port's own, so this ratio is "
                                         "about 1 whatever the kernels do" % (host_threads, threads)}}
    del cpu_ref, p_sts, p_svc, p_vs, p_cull
    if world == 1 and not args.no_extra:
        torch.cuda.empty_cache()
        line["other_configs"] = other_configs(pkg, dev, ctx, torch, peak)
        # the deployment this graft argues for, at the size of the headline: the whole fleet resident in an nbg_store
        del resident, d_sts, d_svc, d_vs, d_cull, batches
        shard.dev_batches = shard.named = None
        torch.cuda.empty_cache()
        line["other_configs"]["resident_store"] = resident_store(pkg, O, ctx, torch, fleet, cfleet, total, now, peak,
                                                                 prepacked=(shard.host_batches, shard.svc, shard.vs, want))
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def multi_device_context_check(pkg, O, n_devices):
    """One nbg_step through a context created with ALL local devices (nbg_init(devices, N): one process, the batch
    split into contiguous ranges, one per device), checked against the packed oracle.  Outside every timed region."""
    host = pkg.host
    try:
        ctx = pkg.Context(list(range(n_devices)))
    except Exception as e:   # another rank may hold a device exclusively on some set-ups
        return "unavailable: %s" % e
    n = 300_001
    f = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 12, n, 500)
    sts, svc, vs = f.pack(lean=True)
    cf = host.CullFleet(host.SEED_BASE + 12, n)
    cb = cf.pack()
    now = (cf.now_sec, cf.now_nsec)
    r = ctx.step(sts.sparse(), svc.gen(), vs.gen(), host.vs_gen_consts(), cb, now[0], now[1])
    ok = (np.array_equal(r["sts"][0], O.packed_diff_sts_sparse(sts.sparse())[0]) and np.array_equal(r["svc"][0], O.packed_diff_svc(svc)[0])
          and np.array_equal(r["vs"][0], O.packed_diff_vs(vs)[0]) and np.array_equal(r["cull"][0], O.packed_cull(cb, now, 1440, 1, 1)[0]))
    ctx.close()
    if not ok:
        raise SystemExit("bench: the multi-device context differs from the packed oracle")
    return "bit-exact (nbg_init with %d devices, %d notebooks through nbg_step)" % (n_devices, n)


def resident_store(pkg, O, ctx, torch, fleet, cfleet, per, now, peak, prepacked=None):
    """The first `per` notebooks of the fleet held in an nbg_store: load, step from the store (spans instead of CSR),
    then ticks in which 1 % of the slots are replaced by other notebooks' rows before the step.  Parity after the last
    tick.  prepacked = (host batches, SvcBatch, VsBatch, oracle answers) of exactly those notebooks, when the caller has them."""
    host, N = pkg.host, pkg._native
    if prepacked is not None:
        (sts_sp, svc_g, vs_g, cull), b_svc, b_vs, want = prepacked
        assert len(cull) == per
    else:
        b_sts, b_svc, b_vs = fleet.pack(0, per, lean=True)
        cull = cfleet.pack(0, per)
        sts_sp, svc_g, vs_g = b_sts.sparse(), b_svc.gen(), b_vs.gen()
        th = os.cpu_count() or 1
        want = {"diff_sts": O.packed_diff_sts_sparse(sts_sp, threads=th), "diff_svc": O.packed_diff_svc(b_svc, threads=th),
                "diff_vs": O.packed_diff_vs(b_vs, threads=th), "cull": O.packed_cull(cull, now, 1440, 1, 1, threads=th)}
    alg_bytes = sum(b.algorithmic_bytes() for b in (sts_sp, svc_g, vs_g, cull))
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
    tick_out = pkg.engine.StepBuffers(per, pinned=True)              # a controller keeps its result buffers
    for t in range(6):
        slots = rng.choice(per, size=m, replace=False).astype(np.uint64)
        alt = (per + t * m + np.arange(m)).astype(np.uint64) % np.uint64(fleet.n)
        a_sts, a_svc, a_vs = fleet.pack(idx=alt, lean=True)
        a_cf = host.CullFleet(host.SEED_BASE + 900 + t, m)
        a_cull = a_cf.pack()
        rows = (a_sts.sparse(), a_svc.gen(), a_vs.gen(), a_cull)
        t0 = time.perf_counter()
        store.put(slots, *rows)
        put_s = time.perf_counter() - t0
        r = store.step(per, consts, now[0], now[1], out=tick_out)        # host outputs: the tick a controller would run
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
    r = None
    tick_out.close()
    store.close()
    put_ms = statistics.median(x[0] for x in ticks[1:]) * 1e3
    tick_ms = statistics.median(x[1] for x in ticks[1:]) * 1e3
    return {"notebooks": per, "load_seconds": load_s, "load_GBps": (alg_bytes / load_s) / 1e9,
            "step_ms_fresh": fresh_ms, "step_objects_per_s_fresh": per / (fresh_ms * 1e-3), "step_frac_of_peak_fresh": alg_bytes / (fresh_ms * 1e-3) / 1e9 / peak,
            "step_ms_after_6_ticks": churned_ms, "rows_replaced_per_tick": m,
            "tick_ms": tick_ms, "tick_put_ms": put_ms, "tick_objects_per_s": per / (tick_ms * 1e-3),
            "store": st, "parity": "bit-exact vs packed oracle after the last tick (replaced slots included)",
            "note": "tick = nbg_store_put of 1 % of the rows (host buffers) + nbg_store_step over all slots with host outputs (pinned, kept across ticks); "
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
    th = os.cpu_count() or 1
    ok4 = (np.array_equal(d1.outs[0].cpu().numpy(), O.packed_diff_sts(sts, threads=th)[0]) and
           np.array_equal(d2.outs[0].cpu().numpy(), O.packed_diff_vs(vs, threads=th)[0]))
    if not ok4:
        raise SystemExit("bench: config #4 results differ from the packed oracle")
    out["config4_100k_odh_sts_vs"] = {"objects_per_s": 100_000 / ((m1 + m2) * 1e-3), "ms": m1 + m2,
                                      "sts_GBps": d1.algorithmic_bytes / (m1 * 1e-3) / 1e9, "sts_frac_of_peak": d1.algorithmic_bytes / (m1 * 1e-3) / 1e9 / peak,
                                      "parity": "bit-exact vs packed oracle on all 100 000 StatefulSet and VirtualService pairs "
                                                "(tests/test_gpu_parity.py::test_config4_odh_notebooks adds the object-level oracle on all of them)"}
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
