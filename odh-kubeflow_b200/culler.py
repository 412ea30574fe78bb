"""Same names and behaviour as the predicates of the reference's
components/notebook-controller/controllers/culling_controller.go, one notebook
at a time; each call runs the cull kernel on a batch of one through the C ABI.

`meta` is a dict with an optional "annotations" dict (metav1.ObjectMeta), kernels
and terminals are lists of dicts as /api/kernels and /api/terminals return them.
`now` is (unix_sec, nsec); default = the wall clock, as time.Now() in Go.
"""
import os
import time

from . import _native as N
from . import host
from .reconcilehelper import default_context

STOP_ANNOTATION = "kubeflow-resource-stopped"
LAST_ACTIVITY_ANNOTATION = "notebooks.kubeflow.org/last-activity"
LAST_ACTIVITY_CHECK_TIMESTAMP_ANNOTATION = "notebooks.kubeflow.org/last_activity_check_timestamp"
KERNEL_EXECUTION_STATE_IDLE = "idle"
KERNEL_EXECUTION_STATE_BUSY = "busy"
KERNEL_EXECUTION_STATE_STARTING = "starting"

CULL_IDLE_TIME = 0           # minutes; CULL:37
IDLENESS_CHECK_PERIOD = 0    # minutes; CULL:39


def initGlobalVars():
    """CULL:525-558 (the two variables the predicates read)."""
    global CULL_IDLE_TIME, IDLENESS_CHECK_PERIOD
    CULL_IDLE_TIME = int(os.environ.get("CULL_IDLE_TIME") or "1440")
    IDLENESS_CHECK_PERIOD = int(os.environ.get("IDLENESS_CHECK_PERIOD") or "1")


def _now(now):
    if now is not None:
        return now
    t = time.time_ns()
    return t // 1_000_000_000, t % 1_000_000_000


def _run(meta, kernels, terminals, now, mode=N.CULL_PREDICATES):
    sec, nsec = _now(now)
    batch = host.CullBatch().add({"annotations": (meta or {}).get("annotations"),
                                  "kernels": kernels or [], "terminals": terminals or []})
    action, argmax, _, _ = default_context().cull(batch, sec, nsec, CULL_IDLE_TIME, IDLENESS_CHECK_PERIOD, mode)
    return batch, int(action[0]), int(argmax[0]), sec


def StopAnnotationIsSet(meta):
    """CULL:493-502 (pure key presence; no device work involved)."""
    ann = (meta or {}).get("annotations")
    return ann is not None and STOP_ANNOTATION in ann


def allKernelsAreIdle(kernels, now=None):
    """CULL:315-327."""
    if not kernels:
        return True
    probe = {"annotations": {LAST_ACTIVITY_ANNOTATION: "0001-01-01T00:00:00Z"}}
    _, action, _, _ = _run(probe, kernels, None, now)
    return (action & N.ACT_SRC_MASK) != 1


def notebookIsIdle(meta, now=None):
    """CULL:220-241."""
    _, action, _, _ = _run(meta, None, None, now)
    return bool(action & N.ACT_CULL)


def cullingCheckPeriodHasPassed(meta, now=None):
    """CULL:207-217."""
    _, action, _, _ = _run(meta, None, None, now)
    return bool(action & N.ACT_PERIOD)


def updateTimestampFromKernelsActivity(meta, kernels, now=None):
    """CULL:371-402: may rewrite meta["annotations"][LAST_ACTIVITY_ANNOTATION]."""
    if not kernels:
        return False
    batch, action, argmax, sec = _run(meta, kernels, None, now)
    src = action & N.ACT_SRC_MASK
    if src:
        meta["annotations"][LAST_ACTIVITY_ANNOTATION] = batch.new_last_activity(0, action, argmax, sec)
    return src == 2


def updateTimestampFromTerminalsActivity(meta, terminals, now=None):
    """CULL:404-428."""
    if not terminals:
        return False
    batch, action, argmax, sec = _run(meta, None, terminals, now)
    src = action & N.ACT_SRC_MASK
    if src:
        meta["annotations"][LAST_ACTIVITY_ANNOTATION] = batch.new_last_activity(0, action, argmax, sec)
    return src == 3


def setStopAnnotation(meta, now=None):
    """CULL:475-491 (host-side write; metrics are out of scope)."""
    if meta is None:
        return
    sec, _ = _now(now)
    if not meta.get("annotations"):
        meta["annotations"] = {}
    meta["annotations"][STOP_ANNOTATION] = host.format_rfc3339(sec)
