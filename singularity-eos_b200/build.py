"""Build libsgx_b200.so (sm_100a) in-tree with nvcc.  `python singularity-eos_b200/build.py`.

The library is plain CUDA C++ behind a C ABI (include/sgx_b200.h); no torch headers.  It is
written to  singularity-eos_b200/lib/libsgx_b200.so  (git-ignored, travels with gpurun).
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libsgx_b200.so")
SOURCES = ["sgx_api.cu", "sgx_spiner_host.cu", "sgx_pte.cu", "sgx_pte_rhot.cu", "sgx_helm.cu",
           "sgx_spiner_root.cu", "sgx_comm.cu", "sgx_pte_fixedt.cu", "sgx_pte_rhou.cu", "sgx_libm.cu",
           "sgx_api_extra.cu", "sgx_pte_tab.cu", "sgx_pte_tab_fixedt.cu", "sgx_pte_tab_pt.cu", "sgx_get_sg_eos.cu", "sgx_pte_fixedp.cu",
           "sgx_tune_lookup.cu", "sgx_modifiers.cu", "sgx_get_sg_eos_rho_e.cu",
           "sgx_get_sg_eos_rho_t.cu", "sgx_get_sg_eos_p_t.cu", "sgx_pte_tab_plain.cu",
           "sgx_pte_tab_sm.cu", "sgx_pte_tab_sm_mod.cu", "sgx_pte_tab_rhou.cu",
           "sgx_pte_tab_fixedp.cu", "sgx_pte_tab_pt4.cu", "sgx_pte_tab_rhou4.cu",
           "sgx_pte_tab_fixedp4.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
# -fmad=false: no FMA contraction, so device arithmetic rounds like the reference's x86-64
# build (parity contract: 1e-12 lookups, 1e-10 PTE states, identical flags).
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo",
    "-fmad=false", "-Xcompiler", "-fPIC,-O2,-ffp-contract=off", "-Xptxas", "-v",
    "--expt-relaxed-constexpr", "-cudart", "static",
]


def _includes(path, seen):
    """Transitive closure of the quoted #include files of one source (paths resolved)."""
    path = os.path.normpath(path)
    if path in seen or not os.path.exists(path):
        return
    seen.add(path)
    with open(path) as fh:
        for line in fh:
            line = line.strip()
            if line.startswith("#include \""):
                _includes(os.path.join(os.path.dirname(path), line.split('"')[1]), seen)


def _digest(src=None):
    """Hash of the flags and of one source with the headers it includes (per-object stamp),
    or of every source (library stamp)."""
    files = set()
    for s in ([src] if src else SOURCES):
        _includes(os.path.join(CSRC, s), files)
    h = hashlib.sha256()
    for f in sorted(files):
        with open(f, "rb") as fh:
            h.update(os.path.basename(f).encode())
            h.update(fh.read())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def _read(path):
    try:
        with open(path) as fh:
            return fh.read().strip()
    except OSError:
        return ""


def build(force=False, verbose=False):
    os.makedirs(LIBDIR, exist_ok=True)
    stamp = os.path.join(LIBDIR, ".stamp")
    dig = _digest()
    if not force and os.path.exists(LIB) and _read(stamp) == dig:
        return LIB
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(LIBDIR, src.replace(".cu", ".o"))
        objs.append(obj)
        odig = _digest(src)
        if not force and os.path.exists(obj) and _read(obj + ".stamp") == odig:
            continue  # this translation unit and the headers are unchanged
        cmd = [NVCC] + FLAGS + ["-c", os.path.join(CSRC, src), "-o", obj]
        log = open(obj + ".log", "w")
        procs.append((subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT), cmd, log,
                      obj, odig))
    for p, cmd, log, obj, odig in procs:
        rc = p.wait()
        log.close()
        if rc != 0:
            sys.stderr.write(open(obj + ".log").read())
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
        with open(obj + ".stamp", "w") as fh:
            fh.write(odig)
        if verbose:
            sys.stdout.write(open(obj + ".log").read())
    cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static",
           "-o", LIB] + objs
    subprocess.check_call(cmd)
    with open(stamp, "w") as fh:
        fh.write(dig)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
