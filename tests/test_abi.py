"""CPU tests of the boundary: the libraries load, export every symbol the headers
declare, the frozen schema constants match, and the product fails loudly (never
falls back) when there is no GPU."""
import ctypes as C
import importlib
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pkg = importlib.import_module("odh-kubeflow_b200")
N = pkg._native


def _declared(header, prefix):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(%s[a-z0-9_]+)\s*\(" % prefix, text)))


def test_every_declared_nbg_symbol_is_exported():
    lib = C.CDLL(os.path.join(ROOT, "odh-kubeflow_b200", "libnbgraft.so"))
    names = _declared("nbgraft.h", "nbg_")
    assert set(N.NBG_SYMBOLS) <= set(names)
    for n in names:
        assert hasattr(lib, n), n


def test_every_declared_nbh_symbol_is_exported():
    lib = C.CDLL(os.path.join(ROOT, "odh-kubeflow_b200", "libnbgraft_host.so"))
    names = _declared("nbgraft_host.h", "nbh_")
    for n in names:
        assert hasattr(lib, n), n
    assert set(N.NBH_SYMBOLS) <= set(names)


def test_schema_constants_match_header():
    text = open(os.path.join(ROOT, "include", "nbgraft.h")).read()
    def const(name):
        return int(re.search(r"#define\s+%s\s+\(?(-?\w+)u?\)?" % name, text).group(1).rstrip("u"), 0)
    assert const("NBG_POD_SLOTS") == N.NBG_POD_SLOTS == 192
    assert const("NBG_SVC_SLOTS") == N.NBG_SVC_SLOTS == 8
    assert const("NBG_VS_SLOTS") == N.NBG_VS_SLOTS == 8
    assert N.nbg().nbg_abi_version() == const("NBG_ABI_VERSION")
    assert C.sizeof(N.nbg_kv) == 16 and C.sizeof(N.nbg_cull_params) == 24


def test_ctypes_mirrors_have_the_c_compilers_layout(tmp_path):
    """Every struct of include/nbgraft.h: size and every field offset as gcc lays it out == the ctypes
    mirror in _native.py (a silent mismatch would shift every pointer after the first difference)."""
    import subprocess
    structs = [n for n in dir(N) if n.startswith("nbg_") and isinstance(getattr(N, n), type) and issubclass(getattr(N, n), C.Structure)]
    assert len(structs) >= 15
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "nbgraft.h"', 'int main(void) {']
    for n in structs:
        lines.append('  printf("%s size %%zu\\n", sizeof(%s));' % (n, n))
        for f, _ in getattr(N, n)._fields_:
            lines.append('  printf("%s %s %%zu\\n", offsetof(%s, %s));' % (n, f, n, f))
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = str(tmp_path / "layout")
    r = subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = subprocess.run([exe], capture_output=True, text=True).stdout.split("\n")
    seen = 0
    for line in out:
        if not line:
            continue
        n, f, v = line.split()
        cls = getattr(N, n)
        assert int(v) == (C.sizeof(cls) if f == "size" else getattr(cls, f).offset), line
        seen += 1
    assert seen > 100


def test_hash_contract():
    e = pkg.engine
    assert e.hash64("") == 0
    seen = {e.hash64("k%d" % i) for i in range(20000)}
    assert 0 not in seen and len(seen) == 20000
    assert e.hash64("abc") == e.hash64(b"abc") != e.hash64("abd")
    for n in (1, 2, 4, 8):
        shards = [e.shard_of(e.hash64("ns-%06d" % i), n) for i in range(4000)]
        assert set(shards) == set(range(n))
        assert max(shards.count(s) for s in range(n)) < 1.25 * 4000 / n


def test_strerror_covers_every_status():
    lib = N.nbg()
    msgs = {lib.nbg_strerror(s) for s in range(0, -8, -1)}
    assert len(msgs) == 8 and b"unknown status" not in msgs


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="this check is for the CPU-only box")
def test_no_device_is_an_error_not_a_fallback():
    with pytest.raises(pkg.NbgError) as ei:
        pkg.Context([0])
    assert ei.value.status == N.NBG_E_NODEVICE
    rh = importlib.import_module("odh-kubeflow_b200.reconcilehelper")
    with pytest.raises(pkg.NbgError):
        rh.CopyServiceFields({"spec": {}}, {"spec": {}})


def test_product_never_references_the_oracle():
    bad = []
    walk = list(os.walk(os.path.join(ROOT, "odh-kubeflow_b200"))) + list(os.walk(os.path.join(ROOT, "tools"))) + \
        list(os.walk(os.path.join(ROOT, "include")))
    for dirpath, _, files in walk:
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"oracle[/_.]|liboracle|import oracle", text):
                    bad.append(os.path.join(dirpath, f))
    assert not bad, bad


def test_libraries_stand_alone():
    """The product is two plain shared libraries: no Python, no torch, no CUDA runtime .so to find at load time
    (cudart is linked statically; the driver's libcuda is opened by it on first use)."""
    import subprocess
    allowed = ("linux-vdso", "libc.so", "libm.so", "libstdc++", "libgcc_s", "libdl", "librt", "libpthread", "ld-linux")
    for name in ("libnbgraft.so", "libnbgraft_host.so"):
        out = subprocess.run(["ldd", os.path.join(ROOT, "odh-kubeflow_b200", name)], capture_output=True, text=True).stdout
        deps = [line.split()[0] for line in out.splitlines() if line.strip()]
        assert deps and all(any(d.startswith(a) or ("/" + a) in d for a in allowed) for d in deps), (name, deps)
