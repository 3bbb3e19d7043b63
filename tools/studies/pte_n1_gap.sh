#!/bin/bash
# Why is PTE RhoT 4 % slower per GPU in a plain `python bench.py` than under torchrun?
O=gpurun_out; mkdir -p $O; F=$O/${1:-r2t}_pte_n1_gap.txt
echo "== plain python" >> $F
PTE_PARAMS=test python tests/prof_driver.py pte 31 >> $F 2>&1
echo "== OMP_NUM_THREADS=1" >> $F
OMP_NUM_THREADS=1 PTE_PARAMS=test python tests/prof_driver.py pte 31 >> $F 2>&1
echo "== torchrun --nproc-per-node 1" >> $F
PTE_PARAMS=test python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29520 tests/prof_driver.py pte 31 >> $F 2>&1
echo "== CUDA_DEVICE_MAX_CONNECTIONS=1" >> $F
CUDA_DEVICE_MAX_CONNECTIONS=1 PTE_PARAMS=test python tests/prof_driver.py pte 31 >> $F 2>&1
echo "== bench.py plain (pte entries)" >> $F
python bench.py --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['configs']['pte_rhot_3mat']['cell_solves_per_s'])" >> $F
echo "== bench.py OMP_NUM_THREADS=1 (pte entries)" >> $F
OMP_NUM_THREADS=1 python bench.py --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['configs']['pte_rhot_3mat']['cell_solves_per_s'])" >> $F
grep -v "^\*\*\*\|Setting OMP" $F
