"""The allocation-free entry point sgx_TemperatureFromDensityInternalEnergy_noalloc (no
counterpart in the reference): same kernels as
sgx_TemperatureFromDensityInternalEnergy on a SpinerEOSDependsRhoT handle, scratch in a static
device array instead of a stream-ordered allocation.  Collected last (file name) so that a
failure here cannot hide another test under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

from helpers import assert_rel, dev

pytestmark = pytest.mark.gpu
TOL = 1e-12  # the tolerance of the T(rho, sie) parity tests


@pytest.fixture(scope="module")
def torch_cuda():
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    torch.cuda.set_device(0)
    return torch


def _bind(lib):
    fn = lib.sgx_TemperatureFromDensityInternalEnergy_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                   C.c_int, C.c_void_p]
    lib.sgx_noalloc_max_cells.restype = C.c_int64
    lib.sgx_noalloc_max_cells.argtypes = []
    return fn


def test_no_alloc_root_find_matches_existing_entry_twice(pkg, torch_cuda):
    torch = torch_cuda
    L, W = pkg._lib, pkg.workloads
    lib = L.load()
    noalloc = _bind(lib)
    nmax = lib.sgx_noalloc_max_cells()
    assert nmax == 1 << 20
    eos = pkg.SpinerEOSDependsRhoT(pkg.tablegen.vinet_sp5_tables())
    # two calls in a row on the shared static scratch: the second with other cells and another
    # size, so stale counters or list entries of the first would show
    for call, (n, start) in enumerate(((1 << 20, 0), (300001, 1 << 21))):
        rho_h, T_h = W.spiner_inputs(n, start)
        rho_d, T_d = dev(rho_h), dev(T_h)
        sie_d = torch.empty_like(rho_d)
        eos.InternalEnergyFromDensityTemperature(rho_d, T_d, sie_d, n,
                                                 dev(np.full((n, 2), np.nan)))
        want = torch.empty_like(rho_d)
        lam_want = dev(np.full((n, 2), np.nan))
        eos.TemperatureFromDensityInternalEnergy(rho_d, sie_d, want, n, lam_want)
        got = torch.full_like(rho_d, float("nan"))
        lam_got = dev(np.full((n, 2), np.nan))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]  # cudaMemGetInfo: sees any allocator
        rc = noalloc(eos.handle, rho_d.data_ptr(), sie_d.data_ptr(), got.data_ptr(), n,
                     lam_got.data_ptr(), 2, None)
        L.check(rc)
        torch.cuda.synchronize()
        if call == 1:  # kernels loaded, static scratch resident: the call takes no device memory
            assert torch.cuda.mem_get_info()[0] == free_before
        g, w = got.cpu().numpy(), want.cpu().numpy()
        assert np.isfinite(g).all()
        assert_rel(g, w, TOL, f"no-alloc T(rho,sie), call {call}")
        assert_rel(lam_got.cpu().numpy(), lam_want.cpu().numpy(), TOL,
                   f"no-alloc lambda (lRho, lT), call {call}")
        # without a lambda array (null pointer), as the existing entry allows
        got.fill_(float("nan"))
        L.check(noalloc(eos.handle, rho_d.data_ptr(), sie_d.data_ptr(), got.data_ptr(), n, None,
                        0, None))
        eos.TemperatureFromDensityInternalEnergy(rho_d, sie_d, want, n)
        torch.cuda.synchronize()
        assert_rel(got.cpu().numpy(), want.cpu().numpy(), TOL, f"no-alloc, no lambda, call {call}")


def test_no_alloc_rejects_sizes_beyond_the_static_scratch(pkg, torch_cuda):
    torch = torch_cuda
    L = pkg._lib
    lib = L.load()
    noalloc = _bind(lib)
    eos = pkg.SpinerEOSDependsRhoT(pkg.tablegen.vinet_sp5_tables())
    n = lib.sgx_noalloc_max_cells() + 1
    a = torch.ones(8, dtype=torch.float64, device="cuda")  # never touched: the size check is first
    rc = noalloc(eos.handle, a.data_ptr(), a.data_ptr(), a.data_ptr(), n, None, 0, None)
    assert rc != 0
    msg = lib.sgx_last_error().decode()
    assert "exceed" in msg and str(1 << 20) in msg, msg
    ideal = pkg.IdealGas(0.4, 7.16e6)
    assert noalloc(ideal.handle, a.data_ptr(), a.data_ptr(), a.data_ptr(), 8, None, 0, None) != 0
