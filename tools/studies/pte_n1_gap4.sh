#!/bin/bash
# two PLAIN processes (no NCCL, no torchrun), one per GPU, at the same time; then one alone
O=gpurun_out; mkdir -p $O; F=$O/${1:-r2x}_pte_two_plain.txt
echo "== two plain processes at once (GPU 0 and GPU 1), cells of rank 0" >> $F
LOCAL_RANK=0 python tools/studies/pte_rank_times.py > $O/_p0.txt 2>&1 &
P0=$!
LOCAL_RANK=1 python tools/studies/pte_rank_times.py > $O/_p1.txt 2>&1 &
P1=$!
wait $P0 $P1
grep "^rank" $O/_p0.txt | sed 's/^/gpu0: /' >> $F; grep "^rank" $O/_p1.txt | sed 's/^/gpu1: /' >> $F
echo "== one plain process on GPU 1 alone" >> $F
LOCAL_RANK=1 python tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
echo "== one plain process on GPU 0 while GPU 1 runs a plain copy loop" >> $F
python - > /dev/null 2>&1 <<PY &
import torch, time
torch.cuda.set_device(1)
a = torch.empty(1 << 28, dtype=torch.float64, device="cuda"); b = torch.empty_like(a)
t0 = time.time()
while time.time() - t0 < 45:
    for _ in range(50): b.copy_(a)
    torch.cuda.synchronize()
PY
BG=$!
sleep 8
LOCAL_RANK=0 python tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
wait $BG
rm -f $O/_p0.txt $O/_p1.txt
cat $F
