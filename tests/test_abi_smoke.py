"""tests/abi_smoke.c: a plain-C client of libdynosam_b200.so on the reference's 3-frame graph.
CPU: it compiles as C against include/dynosam_b200.h, links, and without a device fails with the library's
"no CPU fallback" error (exit code 3). GPU: its chi^2 trace equals the CPU oracle's."""
import os
import subprocess

import numpy as np
import pytest

import scenes
from oracle_lib import OracleGraph

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "_build", "abi_smoke")


def build_abi_smoke():
    from dynosam_b200 import _lib, build
    _lib.load()   # builds the library when it is stale
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    src = os.path.join(ROOT, "tests", "abi_smoke.c")
    libdir = os.path.dirname(build.LIB_PATH)
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", src, "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "tests"),
           "-L" + libdir, "-ldynosam_b200", "-Wl,-rpath," + libdir, "-o", EXE]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    return EXE


CXX_EXE = os.path.join(ROOT, "tests", "_build", "adapter_smoke")


def build_adapter_smoke():
    """include/dynosam_b200/gtsam_adapter.hpp through a C++17 compiler (its GTSAM-free layer)."""
    from dynosam_b200 import _lib, build
    _lib.load()
    os.makedirs(os.path.dirname(CXX_EXE), exist_ok=True)
    libdir = os.path.dirname(build.LIB_PATH)
    cmd = ["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", os.path.join(ROOT, "tests", "adapter_smoke.cpp"), "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "tests"), "-L" + libdir, "-ldynosam_b200", "-Wl,-rpath," + libdir, "-o", CXX_EXE]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    return CXX_EXE


def _run():
    exe = build_abi_smoke() if not os.path.exists(EXE) or os.path.getmtime(EXE) < os.path.getmtime(os.path.join(ROOT, "tests", "abi_smoke.c")) else EXE
    return subprocess.run([exe], capture_output=True, text=True, timeout=300)


def test_c_client_compiles_and_reports_missing_device():
    build_abi_smoke()
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    res = _run()
    if has_gpu:
        assert res.returncode == 0, res.stderr
    else:
        assert res.returncode == 3 and "no CPU fallback" in res.stderr, (res.returncode, res.stderr)


@pytest.mark.gpu
def test_c_client_matches_oracle():
    res = _run()
    assert res.returncode == 0, res.stderr
    tok = res.stdout.split()
    iters, inner, n_trace = int(tok[0]), int(tok[1]), int(tok[2])
    chi2, trace = float(tok[3]), np.array([float(t) for t in tok[4:4 + n_trace]])
    g = scenes.three_frame("graph")
    ora = OracleGraph.from_graph(g)
    assert abs(chi2 - ora.error()) <= 1e-12 * ora.error()
    r = ora.optimize(max_iterations=4, relative_error_tol=0.0, absolute_error_tol=0.0)
    assert (iters, inner) == (r["iterations"], r["inner_iterations"])
    assert np.allclose(trace, r["error_trace"], rtol=1e-6, atol=1e-14)


def test_adapter_header_compiles_and_reports_missing_device():
    exe = build_adapter_smoke()
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    res = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    if has_gpu:
        assert res.returncode == 0, res.stderr
    else:
        assert res.returncode == 3 and "no CPU fallback" in res.stderr, (res.returncode, res.stderr)


@pytest.mark.gpu
def test_adapter_builder_matches_c_client():
    """Factor-by-factor hand-over in graph order (what the GTSAM shim does) gives the result of the batched C client."""
    a = _run()
    b = subprocess.run([build_adapter_smoke()], capture_output=True, text=True, timeout=300)
    assert a.returncode == 0 and b.returncode == 0, (a.stderr, b.stderr)
    ta, tb = a.stdout.split(), b.stdout.split()
    assert ta[:3] == tb[:3]
    assert np.allclose([float(x) for x in ta[3:]], [float(x) for x in tb[3:]], rtol=1e-9, atol=1e-15)
