"""SURVEY 8(f) rank 3, second half: the webhook's restart-block predicate
(maybeRestartRunningNotebook, odh notebook_webhook.go:505-564) with
equality.Semantic.DeepEqual, as semantic sparse tiles on the GPU."""
import copy
import json

import numpy as np
import pytest

import cases
import oracle as O

host, N = O.host, O.N
BLOCKED, EXTERNAL, NOCHANGE = 1, 2, 4


def _nb(fn=None):
    d = copy.deepcopy(cases.BASE_NOTEBOOK)
    if fn:
        fn(d)
    return d


def _c0(d):
    return d["spec"]["template"]["spec"]["containers"][0]


def _mutated(d, flags=2 | 4 | 8 | 16):
    nb = host.Notebook(d)
    host.odh_mutate(nb, flags)
    return nb


# semantic equality of two PodSpecs: (description, edit applied to a copy of the base, equal?)
SEMANTIC = [
    ("identical", lambda d: None, True),
    ("env nil vs [] (an empty slice IS a nil slice)", lambda d: _c0(d).update({"env": []}), True),
    ("nodeSelector {} vs nil", lambda d: d["spec"]["template"]["spec"].update({"nodeSelector": {}}), True),
    ("quantity 500m vs 0.5 (Cmp == 0)", lambda d: _c0(d)["resources"]["requests"].update({"cpu": "0.5"}), True),
    ("quantity 1Gi vs 1024Mi", lambda d: _c0(d)["resources"]["requests"].update({"memory": "1024Mi"}), True),
    ("quantity 1Gi vs 1G", lambda d: _c0(d)["resources"]["requests"].update({"memory": "1G"}), False),
    ("quantity 1Gi vs 1073741824", lambda d: _c0(d)["resources"]["requests"].update({"memory": "1073741824"}), True),
    ("image changed", lambda d: _c0(d).update({"image": "img:2"}), False),
    ("securityContext {} vs nil: pointers stay distinct", lambda d: d["spec"]["template"]["spec"].update({"securityContext": {}}), False),
    ("env [] vs one entry", lambda d: _c0(d).update({"env": [{"name": "A", "value": "1"}]}), False),
    ("limits {} vs absent", lambda d: _c0(d)["resources"].update({"limits": {}}), True),
]


@pytest.mark.parametrize("desc,edit,equal", SEMANTIC, ids=[s[0][:50] for s in SEMANTIC])
def test_semantic_equality_oracle_and_tiles(desc, edit, equal):
    a, b = host.Notebook(_nb()), host.Notebook(_nb(edit))
    pa, pb = a.get("spec.template.spec"), b.get("spec.template.spec")
    assert O.semantic_equal(pa, pb) == equal
    # semantic tiles: old == updated is "not EXTERNAL"; use mutated = old so the only other outcome is NOCHANGE
    rb = host.RestartBatch().add(a, b, a)
    assert (int(O.packed_restart_block(rb)[0][0]) == NOCHANGE) == equal
    # and reflect.DeepEqual is stricter or equal
    if O.deep_equal(pa, pb):
        assert equal


# decisions: (description, old, updated, mutated-builder, create, expected decision)
def _decisions():
    base = _nb()
    injected = _mutated(base).to_python()           # what a previously mutated, stored notebook looks like ... as typed dump
    out = []
    out.append(("create is never blocked (:515)", base, base, lambda u: _mutated(u), True, 0))
    out.append(("webhook would inject a sidecar into a running notebook => BLOCK (:559)", base, base, lambda u: _mutated(u), False, BLOCKED))
    out.append(("user changes the image => their update restarts anyway (:548)", base, _nb(lambda d: _c0(d).update({"image": "img:2"})), lambda u: _mutated(u), False, EXTERNAL))
    out.append(("nothing injected, nothing changed (:554)", base, base, lambda u: host.Notebook(u), False, NOCHANGE))
    out.append(("stopped notebooks may be updated (:521)", base, _nb(lambda d: d["metadata"]["annotations"].update({"kubeflow-resource-stopped": "t"})),
                lambda u: _mutated(u), False, 0))
    out.append(("restarting notebooks may be updated (:527)", base, _nb(lambda d: d["metadata"]["annotations"].update({"notebooks.opendatahub.io/notebook-restart": "true"})),
                lambda u: _mutated(u), False, 0))
    out.append(("the lock annotation the webhook adds counts as stopped", base, base, lambda u: _mutated(u, 1 | 16), False, 0))
    out.append(("only nil-vs-empty and quantity spelling differ => semantically unchanged (:554)", base,
                _nb(lambda d: (_c0(d).update({"env": []}), _c0(d)["resources"]["requests"].update({"cpu": "0.5"}))), lambda u: host.Notebook(u), False, NOCHANGE))
    return out


DECISIONS = _decisions()


@pytest.mark.parametrize("desc,old,upd,mut,create,want", DECISIONS, ids=[d[0][:60] for d in DECISIONS])
def test_restart_block_oracles(desc, old, upd, mut, create, want):
    o, u = host.Notebook(old), host.Notebook(upd)
    m = mut(upd)
    assert O.restart_block(o, u, m, create) == want
    rb = host.RestartBatch().add(o, u, m, create)
    d, bm = O.packed_restart_block(rb)
    assert int(d[0]) == want and int(bm[0]) == int(want == BLOCKED)


def _fleet_triples(n, seed=host.SEED_BASE + 12):
    """old / updated / mutated notebooks for a synthetic fleet: a third were stored before the
    webhook existed (old lacks the injections), a tenth get an external image change, some are
    stopped or being created."""
    fleet = host.Fleet(host.FLEET_KUBEFLOW, seed, n, max(n // 30, 1))
    objs = fleet.objects(0, n)
    rng = np.random.RandomState(7)
    triples = []
    for i, o in enumerate(objs):
        nb = o["notebook"]
        stored_mutated = rng.rand() < 0.65
        old = nb.clone()
        if stored_mutated:
            host.odh_mutate(old, 2 | 4 | 8 | 16)
        updated = old.clone()
        u = rng.rand()
        if u < 0.10:
            updated.set_str("spec.template.spec.containers.0.image", "quay.io/new:image")
        mutated = updated.clone()
        host.odh_mutate(mutated, 2 | 4 | 8 | 16 | (1 if rng.rand() < 0.05 else 0))
        triples.append((old, updated, mutated, bool(rng.rand() < 0.03)))
    return triples


def test_restart_block_fleet_packed_equals_semantic():
    triples = _fleet_triples(1500)
    rb = host.RestartBatch()
    for t in triples:
        rb.add(*t)
    d, bm = O.packed_restart_block(rb)
    for i, t in enumerate(triples):
        assert O.restart_block(*t) == d[i], i
    assert set(d.tolist()) == {0, BLOCKED, EXTERNAL, NOCHANGE}
    assert np.array_equal(O.pkg.engine.bitmap_to_bools(bm, len(triples)), d == BLOCKED)


@pytest.mark.gpu
@pytest.mark.parametrize("desc,old,upd,mut,create,want", DECISIONS, ids=[d[0][:60] for d in DECISIONS])
def test_restart_block_gpu_cases(ctx, desc, old, upd, mut, create, want):
    rb = host.RestartBatch().add(host.Notebook(old), host.Notebook(upd), mut(upd), create)
    d, bm = ctx.restart_block(rb)
    assert int(d[0]) == want and int(bm[0]) == int(want == BLOCKED)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 65, 12000])
def test_restart_block_gpu_fleet(ctx, n):
    triples = _fleet_triples(n)
    rb = host.RestartBatch()
    for t in triples:
        rb.add(*t)
    got, want = ctx.restart_block(rb), O.packed_restart_block(rb)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    for i in range(0, n, 97):
        assert O.restart_block(*triples[i]) == got[0][i]
