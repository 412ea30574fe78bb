"""Structural fuzz of the packing rule.  Random edits anywhere in a StatefulSet's JSON (leaf
changes, dropped fields, nil <-> empty, added / removed / reordered list elements, type
changes of IntOrString, quantity spellings) — after each edit the three encodings must agree
with the object-level oracle:
    dense + sparse PodSpec tiles   <=> reflect.DeepEqual            (CopyStatefulSetFields' P bit)
    flattened value trees          <=> reflect.DeepEqual            (ODH comparators)
    semantic tiles                 <=> equality.Semantic.DeepEqual  (webhook restart-block)
CPU: packed oracle (L2) vs semantic oracle (L1).  GPU: the kernels vs L2."""
import copy
import json
import random

import numpy as np
import pytest

import oracle as O

host, N = O.host, O.N

RICH_NOTEBOOK = {
    "metadata": {"name": "nb-fuzz", "namespace": "ns-fuzz", "labels": {"app": "nb", "team": "x"}, "annotations": {"a": "b"}},
    "spec": {"template": {"spec": {
        "serviceAccountName": "sa", "nodeSelector": {"gpu": "true"}, "terminationGracePeriodSeconds": 30,
        "tolerations": [{"key": "k", "operator": "Exists", "effect": "NoSchedule"}],
        "imagePullSecrets": [{"name": "pull"}],
        "securityContext": {"runAsUser": 1000, "fsGroup": 100, "supplementalGroups": [1, 2]},
        "volumes": [{"name": "ws", "persistentVolumeClaim": {"claimName": "ws-pvc"}},
                    {"name": "shm", "emptyDir": {"medium": "Memory", "sizeLimit": "1Gi"}},
                    {"name": "cfg", "configMap": {"name": "cm", "defaultMode": 420, "optional": True,
                                                  "items": [{"key": "k", "path": "p"}]}},
                    {"name": "sec", "secret": {"secretName": "s"}}],
        "containers": [
            {"name": "nb-fuzz", "image": "img:1", "command": ["start"], "args": ["--x", "1"], "imagePullPolicy": "Always",
             "env": [{"name": "A", "value": "1"}, {"name": "B", "valueFrom": {"secretKeyRef": {"name": "s", "key": "k"}}},
                     {"name": "C", "value": ""}],
             "envFrom": [{"configMapRef": {"name": "cm"}}],
             "ports": [{"name": "http", "containerPort": 8888, "protocol": "TCP"}],
             "resources": {"requests": {"cpu": "500m", "memory": "1Gi"}, "limits": {"cpu": "2", "memory": "2Gi", "nvidia.com/gpu": "1"}},
             "volumeMounts": [{"name": "ws", "mountPath": "/home/jovyan"}, {"name": "shm", "mountPath": "/dev/shm"},
                              {"name": "cfg", "mountPath": "/etc/cfg", "readOnly": True, "subPath": "x"}],
             "livenessProbe": {"httpGet": {"path": "/api", "port": "http", "scheme": "HTTP"}, "initialDelaySeconds": 10, "periodSeconds": 5},
             "readinessProbe": {"tcpSocket": {"port": 8888}, "timeoutSeconds": 1},
             "securityContext": {"allowPrivilegeEscalation": False, "capabilities": {"drop": ["ALL"]}}},
            {"name": "sidecar", "image": "proxy:2", "args": ["--listen=:8443"],
             "resources": {"requests": {"cpu": "100m", "memory": "64Mi"}, "limits": {"cpu": "100m", "memory": "64Mi"}},
             "ports": [{"name": "https", "containerPort": 8443}]}]}}},
}

RESPELL = {"500m": "0.5", "1Gi": "1024Mi", "2Gi": "2048Mi", "2": "2000m", "100m": "0.1", "64Mi": "65536Ki"}
QUANTITIES = ["500m", "0.5", "1", "1000m", "1Gi", "1024Mi", "2Gi", "64Mi", "100m", "0", "1e3", "1k", "2"]


def _paths(node, prefix=()):
    """all (container, key) positions in a JSON tree"""
    out = []
    if isinstance(node, dict):
        for k, v in node.items():
            out.append((node, k))
            out.extend(_paths(v))
    elif isinstance(node, list):
        for i, v in enumerate(node):
            out.append((node, i))
            out.extend(_paths(v))
    return out


def mutate(rng, obj):
    """one random structural edit of the PodSpec (or of metadata now and then), in place"""
    root = obj["spec"]["template"]["spec"] if rng.random() < 0.9 else obj["metadata"]
    pos = _paths(root)
    for _ in range(20):
        cont, key = rng.choice(pos)
        v = cont[key]
        op = rng.randrange(11)
        if op == 0 and isinstance(v, str):
            cont[key] = rng.choice(QUANTITIES) if v in QUANTITIES else (v + "x" if rng.random() < 0.7 else "")
            return "leaf string"
        if op == 1 and isinstance(v, bool):
            cont[key] = not v
            return "bool flip"
        if op == 1 and isinstance(v, int):
            cont[key] = v + rng.choice([-1, 1, 0]) if rng.random() < 0.8 else str(v)   # IntOrString may change type
            return "int"
        if op == 2 and isinstance(cont, dict):
            del cont[key]
            return "drop field"
        if op == 3 and isinstance(v, list):
            cont[key] = [] if v else None
            return "list -> empty / nil"
        if op == 4 and isinstance(v, list) and v:
            if rng.random() < 0.5:
                v.append(copy.deepcopy(rng.choice(v)))
            else:
                v.pop(rng.randrange(len(v)))
            return "list grow/shrink"
        if op == 5 and isinstance(v, list) and len(v) > 1:
            rng.shuffle(v)
            return "list reorder"
        if op == 6 and isinstance(v, dict):
            cont[key] = {} if v else None
            return "dict -> empty / nil"
        if op == 7 and isinstance(cont, dict) and isinstance(v, (dict, list)):
            cont[key + "Extra" if isinstance(key, str) else key] = copy.deepcopy(v)      # a field outside the schema
            return "unknown field"
        if op == 9 and isinstance(v, str) and v in RESPELL:
            cont[key] = RESPELL[v]                        # same number, other spelling: only Semantic.DeepEqual agrees
            return "quantity respelled"
        if op == 10 and isinstance(v, dict) and "name" in v and "image" in v:
            for f in ("command", "args", "env", "envFrom", "volumeMounts"):
                if f not in v:
                    v[f] = []                             # nil -> empty: only Semantic.DeepEqual agrees
                    return "nil -> empty list"
        if op == 8 and isinstance(cont, list):
            cont.insert(key, copy.deepcopy(v))
            return "duplicate element"
    return "nothing"


def _pairs(n, seed):
    rng = random.Random(seed)
    base = O.host.generate_statefulset(host.Notebook(RICH_NOTEBOOK))
    from test_oracle_cases import _typed_json
    base_json = _typed_json(base)
    pairs = []
    while len(pairs) < n:
        a, b = copy.deepcopy(base_json), copy.deepcopy(base_json)
        for _ in range(rng.choice([0, 1, 1, 2, 3])):
            mutate(rng, b)
        if rng.random() < 0.3:
            mutate(rng, a)
        try:
            pairs.append((host.StatefulSet(a), host.StatefulSet(b)))
        except host.HostError:
            continue                                      # the edit produced JSON the typed decoder rejects
    return pairs


def _check_cpu(pairs):
    sts, gen, rb = host.StsBatch(), host.GenBatch(N.GEN_DEEPEQUAL, "metadata.labels", "spec.template.spec"), host.RestartBatch()
    for a, b in pairs:
        sts.add(a, b); gen.add(a, b); rb.add(a, b, a)
    r_dense, r_sparse = O.packed_diff_sts(sts)[0], O.packed_diff_sts_sparse(sts.sparse())[0]
    r_gen, r_rb = O.packed_diff_generic(gen)[0], O.packed_restart_block(rb)[0]
    assert np.array_equal(r_dense, r_sparse)
    n_eq = n_sem = 0
    for i, (a, b) in enumerate(pairs):
        pa, pb = a.get("spec.template.spec") or host.Obj.from_json("PodSpec", {}), b.get("spec.template.spec") or host.Obj.from_json("PodSpec", {})
        deq, seq = O.deep_equal(pa, pb), O.semantic_equal(pa, pb)
        assert bool(r_dense[i] & 8) == (not deq), i
        assert bool(r_gen[i] & 2) == (not deq), i
        assert (int(r_rb[i]) == 4) == seq, i          # old == updated (semantically) -> NO_CHANGE since mutated == old
        assert not deq or seq                          # DeepEqual implies Semantic.DeepEqual
        n_eq += deq; n_sem += seq
        l1 = O.copy_statefulset_fields(a, b.clone(), 1)
        assert l1 == r_dense[i], i
    return sts, gen, rb, n_eq, n_sem


def test_fuzz_encodings_agree_with_the_object_oracle():
    pairs = _pairs(1500, 20251106)
    _, _, _, n_eq, n_sem = _check_cpu(pairs)
    assert 100 < n_eq < 1200 and n_sem > n_eq           # both outcomes, and some pairs equal only semantically


@pytest.mark.gpu
def test_fuzz_gpu(ctx):
    pairs = _pairs(4000, 99)
    sts, gen, rb, _, _ = _check_cpu(pairs)
    assert np.array_equal(ctx.diff_sts(sts)[0], O.packed_diff_sts(sts)[0])
    assert np.array_equal(ctx.diff_sts_sparse(sts.sparse())[0], O.packed_diff_sts(sts)[0])
    assert np.array_equal(ctx.diff_generic(gen)[0], O.packed_diff_generic(gen)[0])
    assert np.array_equal(ctx.restart_block(rb)[0], O.packed_restart_block(rb)[0])
