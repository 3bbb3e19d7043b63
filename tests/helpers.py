"""Shared test helpers (TEST INFRASTRUCTURE)."""
import numpy as np


def is_close(a, b, eps):
    """isClose of test/eos_unit_test_helpers.hpp:60-64: |b-a| / (|a+b| + 1e-20) <= eps."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.abs(b - a) / (np.abs(a + b) + 1e-20) <= eps


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    den = np.maximum(np.abs(b), 1e-300)
    return np.abs(a - b) / den


def assert_rel(a, b, tol, what=""):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    both_nan = np.isnan(a) & np.isnan(b)
    err = np.where(both_nan | (a == b), 0.0, rel_err(a, b))
    worst = int(np.argmax(err))
    assert err.flat[worst] <= tol, (
        f"{what}: max rel err {err.flat[worst]:.3e} > {tol:g} at {worst}: "
        f"got {a.flat[worst]!r} want {b.flat[worst]!r}")


def assert_bits(a, b, what=""):
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    same = a.view(np.uint64) == b.view(np.uint64)
    nz = np.isnan(a) & np.isnan(b)
    ok = same | nz
    if not ok.all():
        i = int(np.argmin(ok))
        raise AssertionError(f"{what}: {int((~ok).sum())}/{ok.size} values differ bitwise; first at "
                             f"{i}: got {a.flat[i]!r} want {b.flat[i]!r}")


def dev(t_np):
    import torch

    return torch.from_numpy(np.ascontiguousarray(t_np)).cuda()


def build_cpp_shim_test():
    """Compile tests/cpp/test_shim.cpp (the C++ host shim over the C ABI) with g++; returns the
    binary path and the environment it needs."""
    import os
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = os.path.join(root, "tests", "cpp", "test_shim.cpp")
    out_dir = os.path.join(root, "tests", "cpp", "_build")
    os.makedirs(out_dir, exist_ok=True)
    exe = os.path.join(out_dir, "test_shim")
    libdir = os.path.join(root, "singularity-eos_b200", "lib")
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    deps = [src, os.path.join(root, "include", "sgx_b200.h"),
            os.path.join(root, "include", "sgx", "eos.hpp"),
            os.path.join(root, "include", "sgx", "closure.hpp")]
    if not os.path.exists(exe) or any(os.path.getmtime(d) > os.path.getmtime(exe) for d in deps):
        subprocess.check_call(
            ["/usr/bin/g++", "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(root, "include"),
             "-I" + os.path.join(cuda, "include"), src, "-o", exe, "-L" + libdir, "-lsgx_b200",
             "-L" + os.path.join(cuda, "lib64"), "-lcudart", "-Wl,-rpath," + libdir,
             "-Wl,-rpath," + os.path.join(cuda, "lib64")])
    return exe


REF_ROOT = "/root/reference"


def build_cpp_plugin_test(force=False):
    """Compile tests/cpp/test_plugin.cpp -- the reference-side plugin (plugins/sgx) against the
    reference headers where they lie -- with the include paths of oracle/Makefile.  Needs
    /root/reference (this container); on the GPU box the prebuilt binary under tests/cpp/_build
    is used.  Returns the binary path, or None when neither is available."""
    import os
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out_dir = os.path.join(root, "tests", "cpp", "_build")
    exe = os.path.join(out_dir, "test_plugin")
    if not os.path.isdir(os.path.join(REF_ROOT, "singularity-eos")):
        return exe if os.path.exists(exe) else None
    os.makedirs(out_dir, exist_ok=True)
    src = os.path.join(root, "tests", "cpp", "test_plugin.cpp")
    deps = [src, os.path.join(root, "include", "sgx_b200.h"),
            os.path.join(root, "oracle", "sp5_tree.hpp")] + \
        [os.path.join(root, "plugins", "sgx", f) for f in os.listdir(os.path.join(root, "plugins", "sgx"))]
    if not force and os.path.exists(exe) and \
            all(os.path.getmtime(d) <= os.path.getmtime(exe) for d in deps):
        return exe
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    libdir = os.path.join(root, "singularity-eos_b200", "lib")
    ref_inc = ["-I" + os.path.join(root, "oracle", "hdf5_shim"),
               "-I" + os.path.join(root, "oracle", "ref_gen"), "-I" + REF_ROOT,
               "-I" + os.path.join(REF_ROOT, "utils", "ports-of-call"),
               "-I" + os.path.join(REF_ROOT, "utils", "spiner"),
               "-I" + os.path.join(REF_ROOT, "utils", "eigen"),
               "-I" + os.path.join(REF_ROOT, "utils", "variant", "include")]
    defs = ["-DSINGULARITY_USE_SPINER", "-DSINGULARITY_USE_SPINER_WITH_HDF5", "-DSPINER_USE_HDF",
            "-DSINGULARITY_USE_HELMHOLTZ", "-DPORTABLE_ALWAYS_THROW_OR_ABORT_THROWS"]
    subprocess.check_call(
        ["/usr/bin/g++", "-std=c++17", "-O1", "-w"] + defs + ref_inc +
        ["-I" + os.path.join(root, "include"), "-I" + os.path.join(cuda, "include"), src, "-o", exe,
         "-L" + libdir, "-lsgx_b200", "-L" + os.path.join(cuda, "lib64"), "-lcudart",
         "-Wl,-rpath,$ORIGIN/../../../singularity-eos_b200/lib", "-Wl,-rpath," + libdir,
         "-Wl,-rpath," + os.path.join(cuda, "lib64")])
    return exe
