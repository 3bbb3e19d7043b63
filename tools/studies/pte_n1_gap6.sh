#!/bin/bash
O=gpurun_out; mkdir -p $O; F=$O/${1:-r2z}_pte_alloc_order.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
echo "== one plain process" >> $F
python tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
echo "== two ranks, communicator first (as bench.py)" >> $F
$TR --master-port 29561 tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
echo "== two ranks, state arrays allocated BEFORE the communicator" >> $F
ALLOC_BEFORE_INIT=1 $TR --master-port 29562 tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
echo "== one process, PYTORCH_CUDA_ALLOC_CONF=expandable_segments:True" >> $F
PYTORCH_CUDA_ALLOC_CONF=expandable_segments:True python tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
cat $F
