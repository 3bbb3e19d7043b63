"""Bit equality of the device exp/log/log10/cbrt/pow (csrc/sgx_libm.cuh) with the host libm.

The reference calls std::pow/exp/log/log10/cbrt (eos/eos_davis.hpp:353-394,415,439;
eos/eos_helmholtz.hpp:750-769,1404-1477); the oracle therefore gets glibc's answers.  The
library runs the same algorithm on the device so that both sides see identical bits.

 * CPU: the header compiles as plain C++; tests/cpp/libm_check.cpp compares it against the
   live libm over ~4.6e7 seeded inputs (all special-case branches included).
 * GPU: >= 2^22 inputs per function through the C ABI (sgx_libm_eval) against the host libm of
   the GPU box, including the operand ranges the Davis EOS and Helmholtz produce.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "cpp", "_build")
GXX = "/usr/bin/g++"


def _build(src, out, flags):
    os.makedirs(BUILD, exist_ok=True)
    csrc = os.path.join(ROOT, "singularity-eos_b200", "csrc")
    deps = [src] + [os.path.join(csrc, f) for f in ("sgx_libm.cuh", "sgx_libm_tables.inc",
                                                     "sgx_linsolve.cuh", "sgx_math.cuh")]
    if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
        subprocess.check_call([GXX] + flags + [src, "-o", out])
    return out


def test_libm_restatement_matches_host_libm_cpu():
    """The shared header, compiled for the host, equals glibc bit for bit."""
    exe = _build(os.path.join(ROOT, "tests", "cpp", "libm_check.cpp"),
                 os.path.join(BUILD, "libm_check"),
                 ["-std=c++17", "-O2", "-mfma", "-ffp-contract=off"])
    out = subprocess.run([exe, "1000000"], capture_output=True, text=True)
    print(out.stdout)
    assert out.returncode == 0, out.stdout[-4000:]
    assert "total mismatches 0" in out.stdout


def test_table_extractor_is_reproducible():
    """The committed tables are what tools/extract_glibc_libm_tables.py reads from this image's
    libm (skipped where the libm build differs)."""
    inc = os.path.join(ROOT, "singularity-eos_b200", "csrc", "sgx_libm_tables.inc")
    committed = open(inc).read()
    bid = subprocess.run(["readelf", "-n", "/lib/x86_64-linux-gnu/libm.so.6"],
                         capture_output=True, text=True).stdout
    if "0d9969fe206760d250ec30a5a9be18aefbf84ea8" not in bid:
        pytest.skip("different libm build")
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "extract", os.path.join(ROOT, "tools", "extract_glibc_libm_tables.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        mod.OUT = os.path.join(td, "t.inc")
        mod.main()
        assert open(mod.OUT).read() == committed


def _host_libm():
    so = _build(os.path.join(ROOT, "tests", "cpp", "libm_host.c"),
                os.path.join(BUILD, "libm_host.so"),
                ["-x", "c", "-O2", "-fno-builtin", "-shared", "-fPIC", "-lm"])
    lib = C.CDLL(so)
    lib.host_libm_version.restype = C.c_char_p
    lib.host_libm_eval.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_long]
    return lib


def _families(rng, n):
    """(function id, name, x, y) input families."""
    u = lambda: rng.random(n)
    logu = lambda lo, hi: np.exp(np.log(lo) + (np.log(hi) - np.log(lo)) * u())
    bits = lambda: rng.integers(0, 2**64, n, dtype=np.uint64).view(np.float64)
    allnorm = lambda: np.ldexp(1.0 + u(), rng.integers(-1022, 1024, n).astype(np.int32))
    fam = [
        (0, "exp uniform[-750,715]", -750 + 1465 * u(), None),
        (0, "exp Davis b4y / -Zy range [-40,40]", -40 + 80 * u(), None),
        (0, "exp ln10*lT (Helmholtz pow10)", 2.30258509299405e+00 * (3 + 10 * u()), None),
        (0, "exp random bits", bits(), None),
        (1, "log all normals", allnorm(), None),
        (1, "log near 1", 0.9 + 0.2 * u(), None),
        (1, "log Helmholtz plasg/ywot range", logu(1e-30, 1e30), None),
        (1, "log random bits", bits(), None),
        (2, "log10 T in 10^[3,13]", logu(1e3, 1e13), None),
        (2, "log10 rho in 10^[-12,15]", logu(1e-12, 1e15), None),
        (2, "log10 all normals", allnorm(), None),
        (2, "log10 random bits", bits(), None),
        (3, "cbrt Helmholtz s range", logu(1e-40, 1e40), None),
        (3, "cbrt all normals +-", allnorm() * np.where(u() < 0.5, 1.0, -1.0), None),
        (3, "cbrt random bits", bits(), None),
        (4, "pow Davis vvc^(+-n), x in [0.05,20], y in [-8,8]", logu(0.05, 20), -8 + 16 * u()),
        (4, "pow Davis T/Ts ^(1+alpha)", logu(1e-3, 1e3), 1.0 + u()),
        (4, "pow Helmholtz plasg^b2", logu(1e-6, 1.0), 0.29561 + 0 * u()),
        (4, "pow x all normals, y in [-3,3]", allnorm(), -3 + 6 * u()),
        (4, "pow x<0, y integer", -logu(1e-2, 1e2), np.floor(-40 + 80 * u())),
        (4, "pow over/underflow edge", logu(1e-2, 1e2), (700 + 60 * u()) * np.where(u() < 0.5, 1, -1)),
        (4, "pow random bits", bits(), bits()),
    ]
    return fam


@pytest.mark.gpu
def test_device_libm_bit_identical_to_host_libm(pkg):
    import torch
    from helpers import assert_bits

    lib = pkg._lib.load()
    host = _host_libm()
    print("host glibc", host.host_libm_version().decode())
    rng = np.random.default_rng(20260101)
    n = 1 << 22
    total = 0
    for which, name, x, y in _families(rng, n):
        x = np.ascontiguousarray(x, dtype=np.float64)
        if which == 4:
            # the over/underflow family is centred on |y log x| ~ 700..760
            if "edge" in name:
                y = y / np.log(x)
            y = np.ascontiguousarray(y, dtype=np.float64)
        want = np.empty_like(x)
        host.host_libm_eval(which, x.ctypes.data, y.ctypes.data if y is not None else None,
                            want.ctypes.data, x.size)
        xd = torch.from_numpy(x).cuda()
        yd = torch.from_numpy(y).cuda() if y is not None else None
        od = torch.empty_like(xd)
        pkg._lib.check(lib.sgx_libm_eval(which, xd.data_ptr(), yd.data_ptr() if yd is not None else None,
                                         od.data_ptr(), x.size, None))
        torch.cuda.synchronize()
        assert_bits(od.cpu().numpy(), want, name)
        total += x.size
        print(f"  {name}: {x.size} inputs bit-identical")
    assert total >= 5 * (1 << 22)


def _linsolve_systems(n, batch=1 << 14):
    rng = np.random.default_rng(1234 + n)
    A = rng.standard_normal((batch, n, n))
    b = rng.standard_normal((batch, n))
    q = batch // 4
    A[q:2 * q] *= 10.0 ** rng.uniform(-12, 12, (q, n, 1))
    b[q:2 * q] *= 10.0 ** rng.uniform(-12, 12, (q, n))
    A[2 * q:3 * q] = 1.0 + 1e-10 * rng.standard_normal((q, n, n))
    if n > 1:
        A[3 * q:3 * q + q // 2, :, 1] = A[3 * q:3 * q + q // 2, :, 0]
    b[3 * q + q // 2:] = 0.0
    return A.reshape(batch, n * n), b


@pytest.mark.parametrize("n", [1, 2, 3, 4, 6])
def test_linsolve_restatement_bit_identical_to_reference_cpu(ref, n):
    """csrc/sgx_linsolve.cuh compiled for the host vs the reference's solve_Ax_b_wscr (Eigen
    colPivHouseholderQr for n >= 3, closure/mixed_cell_models.hpp:176-278): identical bits on
    random, badly scaled, nearly singular and rank-deficient systems."""
    from helpers import assert_bits

    so = _build(os.path.join(ROOT, "tests", "cpp", "linsolve_host.cpp"),
                os.path.join(BUILD, "linsolve_host.so"),
                ["-std=c++17", "-O2", "-ffp-contract=off", "-Wno-unknown-pragmas", "-shared", "-fPIC"])
    h = C.CDLL(so)
    h.host_solve_Ax_b_batch.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_long]
    A, b = _linsolve_systems(n)
    ok_r, x_r = ref.solve_Ax_b_batch(n, A, b)
    a2, x, ok = A.copy(), b.copy(), np.zeros(len(b), dtype=np.int32)
    h.host_solve_Ax_b_batch(n, a2.ctypes.data, x.ctypes.data, ok.ctypes.data, len(b))
    assert (ok == ok_r).all()
    assert_bits(x, x_r, f"solve_Ax_b n={n}")
