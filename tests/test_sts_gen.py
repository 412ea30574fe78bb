"""generateStatefulSet on the device (include/nbgraft.h: nbg_diff_sts_gen): the desired PodSpec tile is made from the
Notebook's own tile instead of being generated and packed on the host.

Three things are pinned here.  The packed oracle's patch of the Notebook tile equals the tile of the StatefulSet the
host builder generates (notebook_controller.go:433-523, restated in csrc/host/builders.cpp and, separately, in
oracle/oracle_semantic.cpp) — that is the claim the feature rests on, and it is checked on the CPU.  The verdicts of
the generated path equal the verdicts of the host-generated path.  And, on the GPU, the kernel equals both."""
import copy
import importlib

import numpy as np
import pytest

import cases
import oracle as O

pkg = O.pkg
host = pkg.host
N = pkg._native

ENVS = [None, {"ADD_FSGROUP": "false"}, {"ADD_FSGROUP": "true"}]


def _nb(fn=None):
    o = copy.deepcopy(cases.BASE_NOTEBOOK)
    if fn:
        fn(o)
    return o


def _c0(nb):
    return nb["spec"]["template"]["spec"]["containers"][0]


def _pod(nb):
    return nb["spec"]["template"]["spec"]


def _env_of(k):
    return [{"name": f"V{j}", "value": str(j)} for j in range(k)]


# (name, edit of the base Notebook): every branch of generateStatefulSet's PodSpec patch, and the tile's limits
NOTEBOOKS = [
    ("base", None),
    ("workingDir set", lambda o: _c0(o).update(workingDir="/work")),
    ("workingDir empty string", lambda o: _c0(o).update(workingDir="")),
    ("ports given", lambda o: _c0(o).update(ports=[{"containerPort": 9999, "name": "http", "protocol": "TCP"}])),
    ("two ports given", lambda o: _c0(o).update(ports=[{"containerPort": 9999}, {"containerPort": 7, "name": "x"}])),
    ("ports empty, not nil", lambda o: _c0(o).update(ports=[])),
    ("env empty, not nil", lambda o: _c0(o).update(env=[])),
    ("env with other variables", lambda o: _c0(o).update(env=_env_of(3))),
    ("NB_PREFIX already there", lambda o: _c0(o).update(env=[{"name": "A", "value": "b"}, {"name": "NB_PREFIX", "value": "/mine"}])),
    ("NB_PREFIX there, from a secret", lambda o: _c0(o).update(env=[{"name": "NB_PREFIX", "valueFrom": {"secretKeyRef": {"name": "s", "key": "k"}}}])),
    ("env of 12", lambda o: _c0(o).update(env=_env_of(12))),
    ("env of 13: the appended one is the tile's last", lambda o: _c0(o).update(env=_env_of(13))),
    ("env of 14: prepatched on the host", lambda o: _c0(o).update(env=_env_of(14))),
    ("env of 15", lambda o: _c0(o).update(env=_env_of(15))),
    ("env of 14 with NB_PREFIX last", lambda o: _c0(o).update(env=_env_of(13) + [{"name": "NB_PREFIX", "value": "x"}])),
    ("env of 16 with NB_PREFIX beyond the tile", lambda o: _c0(o).update(env=_env_of(15) + [{"name": "NB_PREFIX", "value": "x"}])),
    ("securityContext given", lambda o: _pod(o).update(securityContext={"runAsUser": 1000})),
    ("securityContext with fsGroup", lambda o: _pod(o).update(securityContext={"fsGroup": 7})),
    ("securityContext empty, not nil", lambda o: _pod(o).update(securityContext={})),
    ("stopped", lambda o: o["metadata"]["annotations"].update({"kubeflow-resource-stopped": "2026-01-01T00:00:00Z"})),
    ("no labels, no annotations", lambda o: o["metadata"].pop("labels") and o["metadata"].pop("annotations")),
    ("a sidecar", lambda o: _pod(o)["containers"].append({"name": "side", "image": "s:1"})),
    ("a sidecar with its own workingDir/ports/env", lambda o: _pod(o)["containers"].append(
        {"name": "side", "image": "s:1", "workingDir": "/s", "ports": [{"containerPort": 1}], "env": _env_of(2)})),
    ("everything at once", lambda o: (_c0(o).update(workingDir="/w", ports=[{"containerPort": 80}], env=_env_of(5)),
                                        _pod(o).update(securityContext={"fsGroup": 100}))),
    ("name of 52", lambda o: o["metadata"].update(name="n" * 52)),
]


def _tiles(nb, live, env):
    """(generated batch, host batch) for one Notebook against one live StatefulSet."""
    g = host.StsBatch().add_gen(nb, live, env)
    h = host.StsBatch().add(host.generate_statefulset(nb, env), live)
    return g, h


@pytest.mark.parametrize("env", ENVS, ids=["default", "no-fsgroup", "fsgroup"])
@pytest.mark.parametrize("name,edit", NOTEBOOKS, ids=[c[0] for c in NOTEBOOKS])
def test_oracle_patch_of_the_notebook_tile_is_the_generated_statefulsets_tile(name, edit, env):
    nb = host.Notebook(_nb(edit))
    live = host.generate_statefulset(nb, env)
    g, h = _tiles(nb, live, env)
    consts = host.sts_gen_consts(env)
    gen = g.gen_inputs
    made = O.generate_sts_tile(g.arrays()["pod_from"], gen.prefix_hash[0], gen.flags[0], consts)
    assert np.array_equal(made, h.arrays()["pod_from"])
    # ...and so a live object equal to the desired one needs no update, whichever way the desired side was made
    assert O.packed_diff_sts_gen(g, consts)[0].tolist() == O.packed_diff_sts(h)[0].tolist() == [0]
    # the other fixed-size inputs agree too
    for k in ("replicas_from", "tmpl_from", "labels_from.off", "labels_from.ent", "ann_from.off"):
        assert np.array_equal(g.arrays()[k], h.arrays()[k]), k


@pytest.mark.parametrize("env", ENVS, ids=["default", "no-fsgroup", "fsgroup"])
def test_host_patched_pairs_are_the_generated_pairs_array_for_array(env):
    """nbh_sts_add_notebook (the same edits made by the host packer, for the sparse step) against
    add(generate_statefulset(nb), live): every dense and every sparse array, bit for bit."""
    a, b = host.StsBatch(), host.StsBatch()
    other = host.generate_statefulset(host.Notebook(_nb(lambda o: _c0(o).update(image="img:2"))), env)
    for name, edit in NOTEBOOKS:
        nb = host.Notebook(_nb(edit))
        desired = host.generate_statefulset(nb, env)
        for live in (desired, other):
            a.add_notebook(nb, live, env)
            b.add(desired, live)
    for x, y in ((a, b), (a.sparse(), b.sparse())):
        xa, ya = x.arrays(), y.arrays()
        assert xa.keys() == ya.keys()
        for k in xa:
            assert np.array_equal(xa[k], ya[k]), k
    with pytest.raises(host.HostError):
        host.StsBatch().add_notebook(host.Notebook(_nb(lambda o: _pod(o).update(containers=[]))), other, env)
    with pytest.raises(host.HostError):                           # one batch, one kind of `from` side
        host.StsBatch().add_gen(host.Notebook(_nb()), other).add_notebook(host.Notebook(_nb()), other)


def test_host_patched_fleet_pairs_in_a_lean_batch():
    n = 1500
    fleet = host.Fleet(host.FLEET_ODH, host.SEED_BASE + 49, n, 40)
    a, b = host.pack_objects([], lean=True)[0], host.StsBatch()
    assert a.lean and len(a) == 0
    for i in range(n):
        nb, live = fleet.get(i, "notebook"), fleet.get(i, "sts_to")
        a.add_notebook(nb, live)
        b.add(host.generate_statefulset(nb), live)
    xa, ya = a.sparse().arrays(), b.sparse().arrays()
    for k in xa:
        assert np.array_equal(xa[k], ya[k]), k
    assert np.array_equal(O.packed_diff_sts_sparse(a.sparse())[0], O.packed_diff_sts(b)[0])


def test_flags_say_what_the_host_did():
    g = host.StsBatch()
    for name in ("base", "stopped", "env of 13: the appended one is the tile's last", "env of 14: prepatched on the host"):
        nb = host.Notebook(_nb(dict(NOTEBOOKS)[name]))
        g.add_gen(nb, host.generate_statefulset(nb))
    gen = g.gen_inputs
    assert [gen.flags[i] for i in range(4)] == [0, N.GEN_STOPPED, 0, N.GEN_PREPATCHED]
    assert len({gen.prefix_hash[i] for i in range(4)}) == 1        # same namespace/name


def test_add_gen_refuses_what_generate_statefulset_refuses():
    with pytest.raises(host.HostError):
        nb = host.Notebook(_nb(lambda o: _pod(o).update(containers=[])))
        host.StsBatch().add_gen(nb, host.generate_statefulset(host.Notebook(_nb())))
    assert host.StsBatch().gen_inputs is not None                  # an empty batch is a (trivial) gen batch
    b = host.StsBatch().add(host.generate_statefulset(host.Notebook(_nb())), host.generate_statefulset(host.Notebook(_nb())))
    with pytest.raises(host.HostError):
        b.gen_inputs                                                # filled with add(), not add_gen()


def _fleet_batches(kind, seed, n, env=None):
    fleet = host.Fleet(kind, seed, n, max(n // 40, 1))
    g, h = host.StsBatch(), host.StsBatch()
    for i in range(n):
        nb, live = fleet.get(i, "notebook"), fleet.get(i, "sts_to")
        g.add_gen(nb, live, env)
        h.add(host.generate_statefulset(nb, env), live)
    return fleet, g, h


@pytest.mark.parametrize("kind,n", [(host.FLEET_KUBEFLOW, 3000), (host.FLEET_ODH, 2000)])
def test_fleet_verdicts_do_not_depend_on_where_the_desired_side_is_made(kind, n):
    """Kubeflow fleet: the Notebook is what the controller generates from.  (ODH fleet: the webhook has already
    mutated the Notebook the controller sees, and the fleet's `notebook` is that mutated object.)"""
    fleet, g, h = _fleet_batches(kind, host.SEED_BASE + 41, n)
    consts = host.sts_gen_consts()
    want = O.packed_diff_sts(h)
    got = O.packed_diff_sts_gen(g, consts)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    assert len(set(want[0].tolist())) > 3                          # the fleet's drift classes are all there
    # the generated tiles themselves, one by one
    a, gen = g.arrays()["pod_from"].reshape(n, -1), g.gen_inputs
    hh = h.arrays()["pod_from"].reshape(n, -1)
    for i in range(0, n, 7):
        assert np.array_equal(O.generate_sts_tile(a[i], gen.prefix_hash[i], gen.flags[i], consts), hh[i])
    # the semantic oracle (objects in, no tiles anywhere) agrees with both
    l1 = O.l1_diff_all(fleet, np.arange(0, n, 11, dtype=np.uint64))["sts"]
    assert np.array_equal(got[0][::11], l1)


# ------------------------------------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def ctx():
    c = pkg.engine.Context()
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("env", ENVS, ids=["default", "no-fsgroup", "fsgroup"])
def test_kernel_generates_what_the_host_generates(ctx, env):
    g, h = host.StsBatch(), host.StsBatch()
    for name, edit in NOTEBOOKS:
        nb = host.Notebook(_nb(edit))
        desired = host.generate_statefulset(nb, env)
        other = host.generate_statefulset(host.Notebook(_nb(lambda o: _c0(o).update(image="img:2"))), env)
        for live in (desired, other):
            g.add_gen(nb, live, env)
            h.add(desired, live)
    consts = host.sts_gen_consts(env)
    got = ctx.diff_sts_gen(g, consts)
    want = ctx.diff_sts(h)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    assert np.array_equal(got[0], O.packed_diff_sts_gen(g, consts)[0])
    assert got[0][0::2].tolist() == [0] * len(NOTEBOOKS)            # live == desired: nothing to do
    assert all(r & N.STS_PODSPEC for r in got[0][1::2].tolist()[1:])


@pytest.mark.gpu
@pytest.mark.parametrize("kind,n", [(host.FLEET_KUBEFLOW, 20001), (host.FLEET_ODH, 5000), (host.FLEET_KUBEFLOW, 1), (host.FLEET_ODH, 33)])
def test_kernel_fleet_verdicts(ctx, kind, n):
    fleet, g, h = _fleet_batches(kind, host.SEED_BASE + 43, n)
    consts = host.sts_gen_consts()
    got = ctx.diff_sts_gen(g, consts)
    want = O.packed_diff_sts(h)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    again = ctx.diff_sts(h)
    assert np.array_equal(got[0], again[0]) and np.array_equal(got[1], again[1])


@pytest.mark.gpu
def test_kernel_on_resident_batches_tiled(ctx):
    import torch
    device = importlib.import_module("odh-kubeflow_b200.device")
    n, tile = 4001, 5
    _, g, h = _fleet_batches(host.FLEET_KUBEFLOW, host.SEED_BASE + 47, n)
    want = O.packed_diff_sts(h)[0]
    db = device.DeviceBatch.from_host(host.StsGenView(g), tile=tile)
    assert db.n == n * tile
    reason, bitmap = db.run(ctx)
    torch.cuda.synchronize()
    assert np.array_equal(reason.cpu().numpy(), np.tile(want, tile))
    bits = np.unpackbits(bitmap.cpu().numpy(), bitorder="little")[:db.n].astype(bool)
    assert np.array_equal(bits, np.tile((want & N.STS_UPDATE_MASK) != 0, tile))


@pytest.mark.gpu
def test_kernel_with_the_wrong_constants_sees_drift(ctx):
    """The constants are inputs, not baked in: with ADD_FSGROUP's default the kernel adds fsGroup 100, which a desired
    side generated under ADD_FSGROUP=false does not have."""
    nb = host.Notebook(_nb())
    live = host.generate_statefulset(nb, {"ADD_FSGROUP": "false"})
    g = host.StsBatch().add_gen(nb, live, {"ADD_FSGROUP": "false"})
    assert ctx.diff_sts_gen(g, host.sts_gen_consts({"ADD_FSGROUP": "false"}))[0].tolist() == [0]
    assert ctx.diff_sts_gen(g, host.sts_gen_consts())[0].tolist() == [N.STS_PODSPEC]


@pytest.mark.gpu
def test_kernel_rejects_bad_arguments(ctx):
    lib = N.nbg()
    nb = host.Notebook(_nb())
    g = host.StsBatch().add_gen(nb, host.generate_statefulset(nb))
    view, gen, consts = g.view, g.gen_inputs, host.sts_gen_consts()
    r, b = np.zeros(1, np.uint8), np.zeros(1, np.uint64)
    import ctypes as C
    rp, bp = r.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p)
    assert lib.nbg_diff_sts_gen(ctx.handle, None, C.byref(gen), C.byref(consts), 1, rp, bp) == N.NBG_E_INVAL
    assert lib.nbg_diff_sts_gen(ctx.handle, C.byref(view), None, C.byref(consts), 1, rp, bp) == N.NBG_E_INVAL
    assert lib.nbg_diff_sts_gen(ctx.handle, C.byref(view), C.byref(gen), None, 1, rp, bp) == N.NBG_E_INVAL
    empty = N.nbg_sts_gen_inputs()
    assert lib.nbg_diff_sts_gen(ctx.handle, C.byref(view), C.byref(empty), C.byref(consts), 1, rp, bp) == N.NBG_E_INVAL
    assert lib.nbg_diff_sts_gen(ctx.handle, C.byref(view), C.byref(gen), C.byref(consts), 0, rp, bp) == N.NBG_OK
