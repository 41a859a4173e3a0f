"""CPU suite, part 2: the drop-in boundary. The C-ABI library must load on a box without a GPU, export every
symbol include/ekat_b200/ekat_b200.h declares, validate arguments the way the header promises (no compute is
attempted here), and the product must not reach into oracle/ or carry a CPU fallback."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_every_declared_symbol():
    from ekat_b200 import _lib
    L = _lib.lib()
    declared = _lib.declared_symbols()
    assert len(declared) >= 38
    for name in declared:
        assert hasattr(L, name), f"{name} declared in ekat_b200.h but not exported"
    # and nothing is bound from Python that the header does not declare
    assert set(_lib._SIGS) <= set(declared)
    assert L.ekat_b200_version() == 100
    # dynamic symbol table agrees (extern "C", unmangled)
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (ekat_b200_[a-z0-9_]+)", out))
    assert set(declared) <= exported


def test_argument_validation_returns_error_codes_without_touching_the_gpu():
    from ekat_b200 import _lib
    L = _lib.lib()
    p = ctypes.c_void_p(0x1000)  # never dereferenced: validation happens first
    f = L.ekat_b200_tridiag_bfb_f64
    assert f(p, p, p, p, 0, 72, 1, 1, 0, 0, 1, 0, None) == 0                      # ncol == 0: no-op
    assert f(p, p, p, p, 8, 72, 2, 3, 8, 8, 1, 0, None) == -1                     # nA not in {1, nrhs}
    assert f(p, p, p, p, 8, 0, 1, 1, 8, 8, 1, 0, None) == -1                      # nlev < 1
    assert f(p, p, p, p, 8, 72, 1, 1, 4, 8, 1, 0, None) == -1                     # pitch < ncol
    assert f(p, p, p, p, 8, 72, 1, 1, 71, 72, 0, 0, None) == -1                   # col stride < nlev
    assert f(p, p, p, p, 8, 72, 1, 1, 8, 8, 7, 0, None) == -1                     # unknown layout
    assert f(p, p, p, p, 8, 72, 1, 1, 8, 8, 1, 2, None) == -1                     # unknown flag
    assert f(None, p, p, p, 8, 72, 1, 1, 8, 8, 1, 0, None) == -1                  # null pointer
    g = L.ekat_b200_lin_interp_f64
    assert g(None, None, None, None, None, 0, 0, 0, 0, 0, 1, 0, 0, None) == 0     # empty problem (lin_interp_test.cpp:47-92)
    assert g(p, p, p, p, None, 10, 72, 0, 72, 0, 1, 0, 0, None) == 0              # km2 == 0
    assert g(p, p, p, p, None, 10, 1, 8, 1, 8, 1, 0, 0, None) == -1               # km1 < 2
    assert g(p, p, p, p, None, 10, 72, 128, 64, 128, 1, 0, 0, None) == -1         # pitch1 < km1
    assert L.ekat_b200_lin_interp_setup_f32(p, p, None, 10, 72, 128, 72, 128, None) == -1  # setup needs a map
    assert L.ekat_b200_tridiag_lin_interp_f64(p, p, p, p, p, p, p, None, 10, 1, 8, 2, 16, None) == -1  # nlev < 2


def test_product_never_touches_the_oracle_and_has_no_cpu_fallback():
    pkg = os.path.join(ROOT, "ekat_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if not fn.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                continue
            text = open(os.path.join(dp, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", text, re.M), f"{fn} imports oracle/"
            assert "libekat_oracle" not in text and "libekat_ref" not in text, f"{fn} loads a CPU checker"
    for fn in os.listdir(os.path.join(ROOT, "include", "ekat_b200")):
        text = open(os.path.join(ROOT, "include", "ekat_b200", fn)).read()
        assert "oracle" not in text.lower() or "no cpu" in text.lower()


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    from ekat_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError, match="no CPU or PyTorch fallback"):
        _lib.lib()


def test_host_adapters_compile_against_a_view_type():
    """include/ekat_b200/*.hpp are duck-typed on data()/extent_int()/stride(); compile them with g++ against the
    oracle's stand-in View (a Kokkos::View look-alike) and link against the library -- no GPU needed."""
    src = os.path.join(ROOT, "tests", "cpp", "adapter_compile.cpp")
    exe = os.path.join(ROOT, "tests", "cpp", "adapter_compile")
    cmd = ["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "oracle", "standin"),
           src, "-o", exe, "-L" + os.path.join(ROOT, "ekat_b200", "lib"), "-lekat_b200",
           "-Wl,-rpath," + os.path.join(ROOT, "ekat_b200", "lib")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe, "--no-gpu"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "adapter argument checks ok" in r.stdout
