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


EV2K = 1.160451930280894026e4  # eos/get_sg_eos.hpp:68


def get_sg_eos_composed(ref, L, r_l, eos_offsets, offsets, a, mass_frac_cutoff, prm):
    """The (rho, e) branch of the reference's get_sg_eos re-assembled from the reference's own
    pieces -- init_functor / final_functor arithmetic (get_sg_eos_functors.hpp:66-121,213-283)
    in numpy, PTESolverRhoT through the oracle harness with a SEPARATE CacheAccessor buffer,
    scalar lookups through the reference classes.  Needed because the reference driver itself
    places the lambda cache at solver_scratch + neq*(neq+4) + 2*npte (get_sg_eos_rho_e.cpp:49-52),
    INSIDE the 6*nmat doubles PTESolverRhoT keeps there for dpdv / dedv / dpdT / vtemp
    (mixed_cell_models.hpp:873-878,937-945): with materials that write their lambda
    (tabulated ones) every EOS call scribbles over the Jacobian's inputs and the solve returns
    garbage (T ~ 1e-11 K on the test mixture).  With analytic materials nothing is written and
    the real driver is the oracle.  Operates on the dict `a` in place like the driver."""
    import numpy as np

    nmat = len(eos_offsets)
    cd = a["press"].shape[0]
    fm, fv, fie = a["frac_mass"], a["frac_vol"], a["frac_ie"]
    for iloop in range(len(offsets)):
        i = int(offsets[iloop]) - 1
        mass_sum = 0.0
        for m in range(nmat):
            mass_sum += fm[i, m]
            fv[i, m] = fm[i, m]
            fie[i, m] = a["sie"][i] * fm[i, m] * 1.0
        for m in range(nmat):
            fm[i, m] /= mass_sum
        mats = [m for m in range(nmat) if fm[i, m] > mass_frac_cutoff]
        npte = len(mats)
        eos_p = [r_l[eos_offsets[m] - 1] for m in mats]
        vfrac = np.array([[fm[i, m] for m in mats]])
        vfrac_sum = 0.0
        for v in vfrac[0]:
            vfrac_sum += v
        rho = np.array([[fm[i, m] / a["spvol"][i] / fm[i, m] for m in mats]])
        sie = np.array([[a["sie"][i] * fm[i, m] * 1.0 for m in mats]])
        temp = np.zeros((1, npte))
        press = np.zeros((1, npte))
        cache = np.zeros((1, npte, 3))
        if npte > 1:
            ref.pte_solve(L.PTE_RHOT, eos_p, rho, vfrac, sie, temp, press, prm,
                          sie_tot=np.array([a["sie"][i]]), vfrac_tot=np.array([vfrac_sum]),
                          Tguess=np.array([0.0]), lambdas=cache)
        else:
            lam = cache[0, 0:1, :] if eos_p[0].nlambda() else None
            temp[0, 0] = eos_p[0].eval(L.T_FROM_RHO_E, rho[0, :1], sie[0, :1], lam)[0]
            press[0, 0] = eos_p[0].eval(L.P_FROM_RHO_T, rho[0, :1], temp[0, :1], lam)[0]
        a["temp"][i] = 0.0
        a["press"][i] = 0.0
        a["bmod"][i] = a["cv"][i] = a["dpde"][i] = 0.0
        for mp, m in enumerate(mats):
            lam = cache[0, mp:mp + 1, :] if eos_p[mp].nlambda() else None
            a["press"][i] += press[0, mp] * vfrac[0, mp] * 1.0
            a["temp"][i] += temp[0, mp] * vfrac[0, mp] * 1.0
            ie_m = sie[0, mp] * fm[i, m] * mass_sum
            a["sie"][i] += ie_m * 0.0
            fie[i, m] = ie_m
            fv[i, m] = vfrac[0, mp] * a["vol"][i]
            r1, t1 = rho[0, mp:mp + 1], temp[0, mp:mp + 1]
            bmod_m = eos_p[mp].eval(L.BMOD_FROM_RHO_T, r1, t1, lam)[0]
            a["bmod"][i] += bmod_m * vfrac[0, mp]
            cv_m = EV2K * eos_p[mp].eval(L.CV_FROM_RHO_T, r1, t1, lam)[0]
            a["cv"][i] += cv_m * fm[i, m]
            dpde_m = eos_p[mp].eval(L.GAMMA_FROM_RHO_T, r1, t1, lam)[0]
            a["dpde"][i] += dpde_m * vfrac[0, mp]
            a["frac_bmod"][i, m] = bmod_m
            a["frac_cv"][i, m] = cv_m
            a["frac_dpde"][i, m] = dpde_m
        a["temp"][i] /= EV2K
        for m in range(nmat):
            fm[i, m] *= mass_sum
        a["pmax"][i] = a["press"][i] if a["press"][i] > a["pmax"][i] else a["pmax"][i]
