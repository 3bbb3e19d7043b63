#!/bin/bash
O=gpurun_out; mkdir -p $O; F=$O/${1:-r2aa}_pte_one_chunk.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
smi() { nvidia-smi --query-gpu=index,clocks.sm,clocks.mem,power.draw,pstate,clocks_throttle_reasons.active --format=csv,noheader -lms 400 > $O/_smi.txt & echo $!; }
for lg in 23 27; do
  echo "== LOG2_CELLS=$lg one plain process" >> $F
  P=$(smi); LOG2_CELLS=$lg python tools/studies/pte_rank_times.py 2>&1 | grep "^rank.*ms/step" >> $F; kill $P; sort $O/_smi.txt | uniq -c | sort -rn | head -3 >> $F
  echo "== LOG2_CELLS=$lg two ranks" >> $F
  P=$(smi); LOG2_CELLS=$lg $TR --master-port 2957$((lg % 10)) tools/studies/pte_rank_times.py 2>&1 | grep "^rank.*ms/step" >> $F; kill $P; sort $O/_smi.txt | uniq -c | sort -rn | head -4 >> $F
done
rm -f $O/_smi.txt
cat $F
