"""The JSON -> Go-value layer, pinned by hand-written expectations (tests/golden/decode_cases.json).

The semantic oracle walks value trees that the PRODUCT's decoder built (oracle_semantic.cpp includes gval.hpp and
json.hpp), so a wrong nil/empty/zero decision in that decoder changes the oracle and the packer alike and every
parity test stays green.  These cases state what Go does (encoding/json + reflect.DeepEqual on k8s.io/api types) for
exactly those decisions and hold three things to it: the oracle's DeepEqual, the packer's tiles, and (with a GPU) the
kernel's PodSpec bit."""
import json
import os

import numpy as np
import pytest

import oracle as O

host, N = O.host, O.N
G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "decode_cases.json")))
CASES = G["cases"]


def _sts(pod_json):
    pod = json.loads(pod_json)
    return host.StatefulSet({"metadata": {"name": "nb", "namespace": "ns"},
                             "spec": {"replicas": 1, "template": {"spec": pod}}})


@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_oracle_and_packer_follow_go(case):
    a, b = host.Obj.from_json(G["type"], case["a"]), host.Obj.from_json(G["type"], case["b"])
    assert O.deep_equal(a, b) == case["go_deep_equal"], case["why"]
    assert O.deep_equal(b, a) == case["go_deep_equal"]
    assert (a.digest() == b.digest()) == case["go_deep_equal"]
    # through the objects the predicates see: CopyStatefulSetFields' fourth test (COMMON:128) on both layers
    # (json.loads merges duplicate keys the way Go does — last wins — so the pair survives the round trip)
    sa, sb = _sts(case["a"]), _sts(case["b"])
    batch = host.StsBatch().add(sa, sb)
    l2 = int(O.packed_diff_sts(batch)[0][0])
    l2s = int(O.packed_diff_sts_sparse(batch.sparse())[0][0])
    l1 = O.copy_statefulset_fields(sa, sb, 1)
    want = 0 if case["go_deep_equal"] else N.STS_PODSPEC
    assert l1 == l2 == l2s == want, case["why"]


def test_unknown_field_is_the_documented_divergence():
    d = G["known_divergence"]
    a, b = host.Obj.from_json(G["type"], d["a"]), host.Obj.from_json(G["type"], d["b"])
    assert d["go_deep_equal"] is True and O.deep_equal(a, b) is d["here"]      # an extra update, never a missed one


@pytest.mark.gpu
def test_kernel_follows_go_on_the_decode_cases():
    ctx = O.pkg.Context()
    try:
        batch = host.StsBatch()
        for c in CASES:
            batch.add(_sts(c["a"]), _sts(c["b"]))
        want = np.array([0 if c["go_deep_equal"] else N.STS_PODSPEC for c in CASES], np.uint8)
        assert np.array_equal(ctx.diff_sts(batch)[0], want)
        assert np.array_equal(ctx.diff_sts_sparse(batch.sparse())[0], want)
    finally:
        ctx.close()
