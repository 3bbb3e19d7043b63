#!/bin/bash
# on the GPU box: Helmholtz parity + timing with the -fmad=true Helmholtz object swapped in
O=gpurun_out; mkdir -p $O; F=$O/${1:-r2af}_helm_fmad.txt
L=singularity-eos_b200/lib/libsgx_b200.so
echo "== product build (-fmad=false)" >> $F
python tests/prof_driver.py helm 28 2>&1 | grep helm_ >> $F
python -m pytest tests/test_gpu_parity.py -m gpu -q -k helm 2>&1 | tail -3 >> $F
python tests/diag_helm_bits.py 2>&1 | head -22 >> $F
cp $L /tmp/libsgx_keep.so; cp singularity-eos_b200/lib_fmad/libsgx_b200.so $L
echo "== sgx_helm.cu compiled with -fmad=true" >> $F
python tests/prof_driver.py helm 28 2>&1 | grep helm_ >> $F
python -m pytest tests/test_gpu_parity.py -m gpu -q -k helm 2>&1 | grep -v "^$" | tail -25 | cut -c1-220 >> $F
python tests/diag_helm_bits.py 2>&1 | head -22 >> $F
cp /tmp/libsgx_keep.so $L
cat $F
