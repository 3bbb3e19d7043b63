#!/bin/bash
O=gpurun_out; mkdir -p $O; F=$O/${1:-r2y}_pte_peer_access.txt
echo "== one plain process" >> $F
python tools/studies/pte_rank_times.py 2>&1 | grep "^rank\|Peer" >> $F
echo "== one plain process, cudaDeviceEnablePeerAccess(1) before any allocation" >> $F
ENABLE_PEER=1 python tools/studies/pte_rank_times.py 2>&1 | grep "^rank\|Peer" >> $F
echo "== same, NCCL_P2P_DISABLE=1 two-rank torchrun (communicator without peer access)" >> $F
NCCL_P2P_DISABLE=1 NCCL_SHM_DISABLE=0 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
echo "== two-rank torchrun, defaults" >> $F
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 tools/studies/pte_rank_times.py 2>&1 | grep "^rank" >> $F
cat $F
