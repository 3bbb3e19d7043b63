#!/bin/bash
O=gpurun_out; mkdir -p $O; F=$O/${1:-r2ab}_pte_stack_limit.txt
for sl in "" 1024 2048 4096 8192 16384 65536; do
  echo "== STACK_LIMIT=${sl:-default}" >> $F
  STACK_LIMIT=$sl LOG2_CELLS=24 python tools/studies/pte_rank_times.py 2>&1 | grep "ms/step\|StackSize\|SetLimit" >> $F
done
echo "== two-rank torchrun: what stack limit does the process end up with?" >> $F
if [ "$(nvidia-smi -L | wc -l)" -ge 2 ]; then
  LOG2_CELLS=24 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29581 tools/studies/pte_rank_times.py 2>&1 | grep "ms/step\|StackSize" >> $F
fi
cat $F
