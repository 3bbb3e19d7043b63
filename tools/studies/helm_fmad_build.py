"""Study for DESIGN.md 5.4 / SURVEY 7 hard part 1: what does FMA contraction buy on the Helmholtz
kernels and which parity bars survive it?  Builds singularity-eos_b200/lib_fmad/libsgx_b200.so:
every object of the product build, except sgx_helm.o recompiled with -fmad=true.
tools/studies/helm_fmad_run.sh swaps it in on the GPU box (scratch copy) and runs the Helmholtz
parity tests and timings.  Not part of the product build."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "singularity-eos_b200"))
import build as B

B.build()
out = os.path.join(ROOT, "singularity-eos_b200", "lib_fmad")
os.makedirs(out, exist_ok=True)
flags = [f if f != "-fmad=false" else "-fmad=true" for f in B.FLAGS]
obj = os.path.join(out, "sgx_helm.o")
with open(obj + ".log", "w") as log:
    subprocess.check_call([B.NVCC] + flags + ["-c", os.path.join(B.CSRC, "sgx_helm.cu"), "-o", obj],
                          stdout=log, stderr=subprocess.STDOUT)
objs = [obj if s == "sgx_helm.cu" else os.path.join(B.LIBDIR, s.replace(".cu", ".o"))
        for s in B.SOURCES]
lib = os.path.join(out, "libsgx_b200.so")
subprocess.check_call([B.NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-cudart",
                       "static", "-o", lib] + objs)
print(lib)
