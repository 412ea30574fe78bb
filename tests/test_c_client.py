"""The C ABI driven from plain C99 (tests/c/abi_client.c): proves include/*.h are C headers a cgo shim
can include, that both libraries link from compiled code, and — on a GPU — that the reference's
known answers come out of the product path without any Python in between."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "odh-kubeflow_b200")


@pytest.fixture(scope="module")
def client(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("c_client") / "abi_client")
    cmd = ["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-O1", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "c", "abi_client.c"), "-o", exe, "-L", PKG, "-lnbgraft", "-lnbgraft_host",
           "-Wl,-rpath," + PKG]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def _has_gpu():
    import torch
    return torch.cuda.is_available()


def test_c_client_without_a_device_fails_loudly(client):
    """No CUDA device => nbg_init returns NBG_E_NODEVICE; nothing is computed on the CPU instead."""
    if _has_gpu():
        pytest.skip("this box has a GPU; see the gpu test")
    r = subprocess.run([client], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert r.stdout.startswith("nodevice"), r.stdout


@pytest.mark.gpu
def test_c_client_known_answers(client):
    r = subprocess.run([client], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("ok "), r.stdout
