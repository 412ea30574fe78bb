"""nbg_owner_join_static: the owner join for callers that cannot allocate device memory
(a reconcile path with a latency bound, or a controller sharing its GPU with the notebooks
it manages, where an on-demand cudaMalloc can stall or fail).  Its hash table is a fixed
array inside libnbgraft.so instead of caller-provided scratch.

The check: through the C ABI, its output must be identical (bit-exact, like every other
test of this integer path) to nbg_owner_join_device's on the same inputs — twice in a row
with different inputs, because the built-in table keeps the previous call's contents and
has to be cleared by the call itself.  The file is named so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu
host, N = O.host, O.N


def _both(ctx, nb_keys, sts_keys):
    import torch
    lib = N.nbg()
    torch.cuda.set_device(0)                      # the static variant works on the CURRENT device
    n, m = len(nb_keys), len(sts_keys)
    d_nb = torch.from_numpy(nb_keys.view(np.int64)).cuda()
    d_st = torch.from_numpy(sts_keys.view(np.int64)).cuda()
    got = torch.full((n,), 7, dtype=torch.int32, device="cuda")
    ref = torch.full((n,), 9, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    need = lib.nbg_owner_join_scratch_bytes(m)
    scratch = torch.empty(need, dtype=torch.uint8, device="cuda")
    assert lib.nbg_owner_join_device(ctx.handle, 0, d_nb.data_ptr(), n, d_st.data_ptr(), m, ref.data_ptr(),
                                     scratch.data_ptr(), need, stream) == N.NBG_OK
    assert lib.nbg_owner_join_static(d_nb.data_ptr(), n, d_st.data_ptr(), m, got.data_ptr(), stream) == N.NBG_OK
    torch.cuda.synchronize()
    return got.cpu().numpy().view(np.uint32), ref.cpu().numpy().view(np.uint32)


def test_static_join_matches_the_scratch_variant_twice_in_a_row(ctx):
    # first call: a list close to the built-in limit
    f1 = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 7, 120_000, 3000)
    nb1, st1, _ = f1.join_lists(120_000)
    assert len(st1) <= N.NBG_JOIN_STATIC_MAX_STS
    got, ref = _both(ctx, nb1, st1)
    assert np.array_equal(got, ref)
    assert np.array_equal(got, O.packed_owner_join(nb1, st1)[0])
    # second call right after it: other keys, fewer of them — anything the first call left
    # in the shared table would show up as false matches
    f2 = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 8, 5_000, 100)
    nb2, st2, _ = f2.join_lists(5_000)
    probe = np.concatenate([nb2, nb1[:5_000]])    # the second half must now find nothing
    got, ref = _both(ctx, probe, st2)
    assert np.array_equal(got, ref)
    assert (got[5_000:] == 0xFFFFFFFF).all()
    assert np.array_equal(got, O.packed_owner_join(probe, st2)[0])


def test_static_join_claims_no_memory_after_its_first_call(ctx):
    """The first call on a device is the set-up (it may load the module and claim the table); from then on a call
    must leave the device's free memory exactly as it found it — no cudaMalloc behind the caller's back."""
    import torch
    lib = N.nbg()
    torch.cuda.set_device(0)
    f = host.Fleet(host.FLEET_KUBEFLOW, host.SEED_BASE + 9, 20_000, 400)
    nb, st, _ = f.join_lists(20_000)
    d_nb = torch.from_numpy(nb.view(np.int64)).cuda()
    d_st = torch.from_numpy(st.view(np.int64)).cuda()
    out = torch.empty(len(nb), dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    call = lambda: lib.nbg_owner_join_static(d_nb.data_ptr(), len(nb), d_st.data_ptr(), len(st), out.data_ptr(), stream)
    assert call() == N.NBG_OK                      # warm-up: set-up done by now at the latest
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)          # cudaMemGetInfo: the driver's view, not torch's cache
    for _ in range(3):
        assert call() == N.NBG_OK
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 == free0
    assert np.array_equal(out.cpu().numpy().view(np.uint32), O.packed_owner_join(nb, st)[0])


def test_static_join_rejects_lists_beyond_its_limit(ctx):
    import torch
    lib = N.nbg()
    torch.cuda.set_device(0)
    m = N.NBG_JOIN_STATIC_MAX_STS + 1
    keys = torch.ones(m, dtype=torch.int64, device="cuda")
    out = torch.zeros(4, dtype=torch.int32, device="cuda")
    rc = lib.nbg_owner_join_static(keys.data_ptr(), 4, keys.data_ptr(), m, out.data_ptr(), None)
    assert rc == N.NBG_E_TOOBIG
    assert b"maximum" in lib.nbg_strerror(rc)
    # at the limit it still runs
    assert lib.nbg_owner_join_static(keys.data_ptr(), 4, keys.data_ptr(), m - 1, out.data_ptr(), None) == N.NBG_OK
    torch.cuda.synchronize()
    assert (out.cpu().numpy() == 0).all()         # every probe key 1 matches list entry 0
