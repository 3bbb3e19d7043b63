#!/bin/bash
O=gpurun_out; mkdir -p $O; F=$O/${1:-r2v}_pte_rank_times.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
echo "== 1 process, cells of rank 0" >> $F; python tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
echo "== 1 process, cells of rank 1" >> $F; START_OVERRIDE=1 python tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
echo "== 2 processes, own cells" >> $F; $TR --master-port 29541 tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
echo "== 2 processes, both the cells of rank 0" >> $F; START_OVERRIDE=0 $TR --master-port 29542 tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
cat $F
