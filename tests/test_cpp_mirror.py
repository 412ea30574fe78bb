"""include/nbgraft.hpp — the reference's function names in C++ over the two C ABIs — driven by a C++ restatement
of the reference's table tests (tests/cpp/reference_tables.cpp): compiled code all the way, no Python between the
test and the kernels."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "odh-kubeflow_b200")


@pytest.fixture(scope="module")
def exe(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("cpp_mirror") / "reference_tables")
    cmd = ["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-O1", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "reference_tables.cpp"), "-o", out, "-L", PKG, "-lnbgraft", "-lnbgraft_host",
           "-Wl,-rpath," + PKG]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


def test_cpp_mirror_builds_and_refuses_to_run_without_a_device(exe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("this box has a GPU; see the gpu test")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == "nodevice", r.stdout + r.stderr


@pytest.mark.gpu
def test_reference_tables_through_the_cpp_mirror(exe):
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("ok "), r.stdout
