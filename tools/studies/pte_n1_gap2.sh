#!/bin/bash
# second part: a live NCCL communicator in the process / GPU 0 of a multi-GPU box, single process
O=gpurun_out; mkdir -p $O; F=$O/${1:-r2u}_pte_n1_gap2.txt
nvidia-smi --query-gpu=index,name,power.limit,clocks.max.sm,clocks.max.mem,memory.total --format=csv >> $F
echo "== plain python, GPU 0" >> $F
PTE_PARAMS=test python tests/prof_driver.py pte 31 >> $F 2>&1
echo "== PROF_INIT_NCCL=1" >> $F
PROF_INIT_NCCL=1 PTE_PARAMS=test python tests/prof_driver.py pte 31 >> $F 2>&1
echo "== plain python again" >> $F
PTE_PARAMS=test python tests/prof_driver.py pte 31 >> $F 2>&1
cat $F
