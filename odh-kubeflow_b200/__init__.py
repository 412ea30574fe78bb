"""nbgraft — B200-native batch predicates for the Kubeflow/ODH notebook controllers.

Layout (only what the hot path needs):
  csrc/cuda/   sm_100a kernels + the C ABI (libnbgraft.so, include/nbgraft.h)
  csrc/host/   C++ host side: Go-value object model, desired-object builders,
               packer, synthetic fleets (libnbgraft_host.so, include/nbgraft_host.h)
  engine.py          ctypes face of the C ABI (Context)
  host.py            objects / batches / fleets
  reconcilehelper.py Copy*Fields with the reference's names and side effects
  culler.py          the culler predicates with the reference's names

The directory name contains a hyphen, so import it with
``importlib.import_module("odh-kubeflow_b200")``.
"""
from . import _native, engine, host  # noqa: F401
from .engine import Context, NbgError  # noqa: F401

__all__ = ["Context", "NbgError", "engine", "host", "_native"]
