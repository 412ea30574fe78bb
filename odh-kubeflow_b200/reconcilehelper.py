"""Same names, arguments and side effects as the reference's
components/common/reconcilehelper/util.go, one object pair at a time — each call
packs a batch of one and runs the CUDA kernel through the C ABI.  For fleets use
host.Fleet / host.*Batch with engine.Context directly.

Objects are Kubernetes JSON as Python dicts.  As in Go, the functions return
True when `to` needed an update and overwrite the owned fields of `to`.
"""
import copy

from . import _native as N
from . import engine, host

_ctx = None


def default_context():
    global _ctx
    if _ctx is None:
        _ctx = engine.Context()
    return _ctx


def _meta(o):
    return o.setdefault("metadata", {})


def _copy_meta_fields(from_, to):
    for f in ("labels", "annotations"):
        v = from_.get("metadata", {}).get(f)
        if v is None:
            _meta(to).pop(f, None)
        else:
            _meta(to)[f] = copy.deepcopy(v)


def CopyStatefulSetFields(from_, to, ctx=None, caller_prestep=False):
    """COMMON:107-134.  With caller_prestep=True also performs NBCTL:188-194
    (copy template labels iff replicas differ) first, as Reconcile does."""
    ctx = ctx or default_context()
    batch = host.StsBatch().add(host.StatefulSet(from_), host.StatefulSet(to))
    reason, _ = ctx.diff_sts(batch, allow_nil=True)
    r = int(reason[0])
    if r & N.STS_NILREPLICAS:
        raise RuntimeError("panic: runtime error: invalid memory address or nil pointer dereference (COMMON:123)")
    if caller_prestep and (r & N.STS_TMPL_LABELS):
        tl = from_.get("spec", {}).get("template", {}).get("metadata", {}).get("labels")
        tm = to.setdefault("spec", {}).setdefault("template", {}).setdefault("metadata", {})
        if tl is None:
            tm.pop("labels", None)
        else:
            tm["labels"] = copy.deepcopy(tl)
    _copy_meta_fields(from_, to)
    if r & N.STS_REPLICAS:
        to.setdefault("spec", {})["replicas"] = from_["spec"]["replicas"]
    to.setdefault("spec", {}).setdefault("template", {})["spec"] = copy.deepcopy(
        from_.get("spec", {}).get("template", {}).get("spec", {}))
    return (r & 0x0F) != 0


def CopyServiceFields(from_, to, ctx=None):
    """COMMON:166-195."""
    ctx = ctx or default_context()
    batch = host.SvcBatch().add(host.Service(from_), host.Service(to))
    reason, _ = ctx.diff_svc(batch)
    r = int(reason[0])
    _copy_meta_fields(from_, to)
    for f in ("selector", "ports"):
        v = from_.get("spec", {}).get(f)
        if v is None:
            to.setdefault("spec", {}).pop(f, None)
        else:
            to.setdefault("spec", {})[f] = copy.deepcopy(v)
    return (r & 0x0F) != 0


def CopyVirtualService(from_, to, ctx=None):
    """COMMON:199-219 on unstructured objects."""
    ctx = ctx or default_context()
    batch = host.VsBatch().add(host.Unstructured(from_), host.Unstructured(to))
    reason, _ = ctx.diff_vs(batch)
    r = int(reason[0])
    if r & N.VS_FROM_MISSING:
        return False
    if r & (N.VS_TO_MISSING | N.VS_SPEC_DIFFERS):
        to["spec"] = copy.deepcopy(from_["spec"])
        return True
    return False
