"""The N>1 path on CPU: two processes (gloo), each owning the namespaces that
nbg_shard_of assigns to it, pack their shard and gather result bitmaps on rank 0
with no data-path collective.  No GPU here, so each rank's predicate results come
from the packed oracle — what is under test is the host-side sharding/gather logic
that bench.py --gpus N and a multi-process deployment use."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
    import numpy as np, torch, torch.distributed as dist
    import oracle as O
    host = O.host
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%(port)d", rank=int(sys.argv[1]), world_size=2)
    rank, world = dist.get_rank(), dist.get_world_size()
    fleet = host.Fleet(host.FLEET_KUBEFLOW, 77, 3000, 41)
    idx = fleet.shard_indices(world, rank)
    sts = fleet.pack(idx=idx, svc=False, vs=False)[0]
    reason, bitmap, _ = O.packed_diff_sts(sts)
    # host gather of (indices, reasons): variable length => pad to the max
    n = torch.tensor([len(idx)]); sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sizes, n)
    cap = int(max(s.item() for s in sizes))
    pad_i = torch.zeros(cap, dtype=torch.int64); pad_i[:len(idx)] = torch.from_numpy(idx.astype(np.int64))
    pad_r = torch.zeros(cap, dtype=torch.uint8); pad_r[:len(idx)] = torch.from_numpy(reason)
    gi = [torch.zeros(cap, dtype=torch.int64) for _ in range(world)] if rank == 0 else None
    gr = [torch.zeros(cap, dtype=torch.uint8) for _ in range(world)] if rank == 0 else None
    dist.gather(pad_i, gi, dst=0); dist.gather(pad_r, gr, dst=0)
    if rank == 0:
        out = np.full(fleet.n, 255, np.uint8)
        for k in range(world):
            m = int(sizes[k].item())
            out[gi[k][:m].numpy()] = gr[k][:m].numpy()
        whole = O.packed_diff_sts(fleet.pack(svc=False, vs=False)[0])[0]
        assert np.array_equal(out, whole), "gathered shard results differ from the unsharded result"
        assert sum(int(s.item()) for s in sizes) == fleet.n
        print("GLOO_OK")
    dist.barrier()
    dist.destroy_process_group()
""")


def test_two_rank_shard_and_gather(tmp_path):
    port = 29500 + os.getpid() % 2000
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT, "port": port})
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(2)]
    outs = [p.communicate(timeout=300)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "GLOO_OK" in outs[0]
