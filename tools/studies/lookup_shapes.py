"""Kernel-shape study of the headline lookup (run on a B200): times every shape of
csrc/sgx_tune_lookup.cu at 2^26 cells, with and without the lambda write-back, after checking
that each returns the same bits as the product entry point sgx_eos_eval.
usage: python tools/studies/lookup_shapes.py [log2 cells] > profiles/<tag>_lookup_shapes.json"""
import ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import __graft_entry__ as g

pkg = g.load_package(); L = pkg._lib; W = pkg.workloads
lib = L.load()
lib.sgx_tune_lookup.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_int64, C.c_void_p, C.c_int, C.c_void_p]
lib.sgx_tune_lookup.restype = C.c_int
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 26
n = 1 << lg
dev = torch.device("cuda", 0)
eos = pkg.SpinerEOSDependsRhoT(pkg.tablegen.vinet_sp5_tables())
rho, T = W.spiner_inputs(n, 0, xp=torch, device=dev)
want = torch.empty_like(rho); out = torch.empty_like(rho)
lam_want = torch.full((n, 2), float("nan"), dtype=torch.float64, device=dev)
lam = torch.empty_like(lam_want)
eos.PressureFromDensityTemperature(rho, T, want, n, lam_want)
torch.cuda.synchronize()


def run(variant, with_lambda, bps):
    rc = lib.sgx_tune_lookup(eos.handle, variant, rho.data_ptr(), T.data_ptr(), out.data_ptr(), n,
                             lam.data_ptr() if with_lambda else None, bps, None)
    L.check(rc)


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


res = {"cells": n, "shapes": {}}
res["product_lambda_ms"] = timed(lambda: eos.PressureFromDensityTemperature(rho, T, out, n, lam))
res["product_nolambda_ms"] = timed(lambda: eos.PressureFromDensityTemperature(rho, T, out, n))
names = {0: "4 cells/thread, 2 blocks, prefetch", 1: "2 cells, 4 blocks, prefetch",
         2: "2 cells, 4 blocks", 3: "4 cells, 3 blocks", 4: "2 cells, 3 blocks, prefetch",
         5: "1 cell, 4 blocks, prefetch", 6: "1 cell, 6 blocks"}
for base in names:
    for v in (base, base + 10):
        out.zero_(); lam.zero_()
        run(v, True, 0); torch.cuda.synchronize()
        same = bool((out.view(torch.int64) == want.view(torch.int64)).all()) and \
            bool((lam.view(torch.int64) == lam_want.view(torch.int64)).all())
        entry = {"shape": names[base] + (", single-patch assumed" if v >= 10 else ""),
                 "bit_identical_to_sgx_eos_eval": same}
        for bps in (0, 8, 16):
            t1 = timed(lambda: run(v, True, bps)); t0 = timed(lambda: run(v, False, bps))
            entry[f"grid_blocks_per_sm_{bps or 'resident'}"] = {
                "lambda_ms": t1, "lambda_Gcells_s": n / t1 / 1e6,
                "nolambda_ms": t0, "nolambda_Gcells_s": n / t0 / 1e6}
        res["shapes"][str(v)] = entry
        print(v, json.dumps(entry), file=sys.stderr, flush=True)

# What the table gather costs: the product kernel on the same number of cells whose (rho, T) are
# (a) the bench's random states, (b) all equal (every lane of a warp gathers the SAME 32-byte
# record: one L1TEX wavefront per warp instead of up to 32), (c) sorted by table cell.
if os.environ.get("SGX_STUDY_GATHER", "1") == "1":
    gather = {}
    rc, Tc = torch.full_like(rho, 8.9), torch.full_like(T, 650.0)
    key = (torch.log(rho) * 1e3).long() * 100000 + (torch.log(T) * 1e3).long()
    order = torch.argsort(key)
    rs, Ts = rho[order].contiguous(), T[order].contiguous()
    for nm, (r_, t_) in {"random": (rho, T), "all_equal": (rc, Tc), "sorted_by_cell": (rs, Ts)}.items():
        t1 = timed(lambda: eos.PressureFromDensityTemperature(r_, t_, out, n, lam))
        t0 = timed(lambda: eos.PressureFromDensityTemperature(r_, t_, out, n))
        gather[nm] = {"lambda_Gcells_s": n / t1 / 1e6, "nolambda_Gcells_s": n / t0 / 1e6}
    res["gather_cost_product_kernel"] = gather
print(json.dumps(res, indent=1))
