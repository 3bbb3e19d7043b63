#!/bin/bash
# Helmholtz register-budget sweep, PTE timings (both MixParams sets, three mappings) and the ncu
# launch list of the bench command.  Evidence goes to .txt / .csv (*.log is git-ignored).
TAG=${1:-r2x}
O=gpurun_out
mkdir -p $O
for b in 3 4 5; do
  echo "== SGX_HELM_MINB=$b" >> $O/${TAG}_helm_minb.txt
  SGX_HELM_MINB=$b python tests/prof_driver.py helm 26 >> $O/${TAG}_helm_minb.txt 2>&1
done
cat $O/${TAG}_helm_minb.txt
for p in test default; do
  echo "== PTE_PARAMS=$p" >> $O/${TAG}_pte_timing.txt
  PTE_PARAMS=$p python tests/prof_driver.py pte 24 >> $O/${TAG}_pte_timing.txt 2>&1
  PTE_PARAMS=$p python tests/prof_driver.py pte 28 >> $O/${TAG}_pte_timing.txt 2>&1
done
for v in percell warp; do
  echo "== SGX_PTE_KERNEL=$v PTE_PARAMS=test" >> $O/${TAG}_pte_timing.txt
  SGX_PTE_KERNEL=$v PTE_PARAMS=test python tests/prof_driver.py pte 24 >> $O/${TAG}_pte_timing.txt 2>&1
done
cat $O/${TAG}_pte_timing.txt
python bench.py --steps 2 --warmup 1 > $O/${TAG}_bench_short.json 2> $O/${TAG}_bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 1 > $O/${TAG}_ncu_bench.out 2>&1
wc -l $O/${TAG}_launches.csv
python tools/pcie_bw.py > $O/${TAG}_pcie_bw_n1.json 2>&1; tail -3 $O/${TAG}_pcie_bw_n1.json
