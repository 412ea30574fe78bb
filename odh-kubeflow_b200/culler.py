"""Same names and behaviour as the predicates of the reference's
components/notebook-controller/controllers/culling_controller.go, one notebook
at a time; each call runs the cull kernel on a batch of one through the C ABI.

`meta` is a dict with an optional "annotations" dict (metav1.ObjectMeta), kernels
and terminals are lists of dicts as /api/kernels and /api/terminals return them.
`now` is (unix_sec, nsec); default = the wall clock, as time.Now() in Go.
"""
import os
import re
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


_ATOI = re.compile(r"[+-]?[0-9]+\Z")


def _atoi(text):
    """strconv.Atoi: an optional sign and decimal digits, nothing else (no spaces, no underscores, no
    empty string) and the value must fit a 64-bit int; None where Go returns an error."""
    if not _ATOI.match(text):
        return None
    v = int(text)
    return v if -(1 << 63) <= v < (1 << 63) else None


def initGlobalVars():
    """CULL:525-558 (the two variables the predicates read).  As there: an empty variable means the default
    (GetEnvDefault, CULL:505-511); a CULL_IDLE_TIME that is not an integer is logged and replaced by the default
    1440 (:535-542); an IDLENESS_CHECK_PERIOD that is not an integer is an error, which stops the controller's
    set-up (:550-554, :566-569) — ValueError here."""
    global CULL_IDLE_TIME, IDLENESS_CHECK_PERIOD
    idle = _atoi(os.environ.get("CULL_IDLE_TIME") or "1440")
    CULL_IDLE_TIME = 1440 if idle is None else idle
    text = os.environ.get("IDLENESS_CHECK_PERIOD") or "1"
    period = _atoi(text)
    if period is None:
        raise ValueError('strconv.Atoi: parsing "%s": invalid syntax' % text)
    IDLENESS_CHECK_PERIOD = period


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
    """CULL:475-491 (host-side write; metrics are out of scope).  The reference formats time.Now() in the process's
    LOCAL zone (t.Format(time.RFC3339)); this writes UTC ("...Z"), which is the same string only where TZ is UTC, as
    it is in the controller's container image.  The instant is the same either way, and that is all C7 reads back."""
    if meta is None:
        return
    sec, _ = _now(now)
    if not meta.get("annotations"):
        meta["annotations"] = {}
    meta["annotations"][STOP_ANNOTATION] = host.format_rfc3339(sec)
