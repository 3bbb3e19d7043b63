#!/bin/bash
# N-GPU bench (launched the way the driver launches it), PCIe bandwidth at N ranks, the NCCL
# stats all-reduce check.  usage (under gpurun --gpus N): tools/multi_gpu.sh N TAG
N=${1:-2}; TAG=${2:-r2x}
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > $O/${TAG}_bench_${N}gpu.json 2> $O/${TAG}_bench_${N}gpu.err
tail -2 $O/${TAG}_bench_${N}gpu.err | cut -c1-300
$TR --master-port 29512 tools/pcie_bw.py > $O/${TAG}_pcie_bw_n${N}.json 2> $O/${TAG}_pcie_bw_n${N}.err
cut -c1-400 $O/${TAG}_pcie_bw_n${N}.json
$TR --master-port 29513 tests/run_stats_allreduce.py > $O/${TAG}_stats_allreduce_n${N}.txt 2>&1
tail -2 $O/${TAG}_stats_allreduce_n${N}.txt | cut -c1-300
$TR --master-port 29514 bench.py --impl reference --gpus $N --steps 5 --warmup 1 > $O/${TAG}_bench_reference_${N}gpu.json 2> $O/${TAG}_bench_reference_${N}gpu.err
cut -c1-200 $O/${TAG}_bench_reference_${N}gpu.json
python - <<PY
import json
line = [l for l in open("$O/${TAG}_bench_${N}gpu.json") if l.startswith("{")][-1]
open("$O/${TAG}_bench_${N}gpu.json", "w").write(line)  # NCCL prints its version to stdout first
d = json.loads(line)
print("N=$N value", d["value"], "e2e", d["e2e"]["value"], "pte", d["configs"]["pte_rhot_3mat"]["cell_solves_per_s"], "clocks", d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
PY
