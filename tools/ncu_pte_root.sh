#!/bin/bash
# ncu --set full of the PTE state-machine kernel and of the two T(rho,sie) kernels; only text /
# CSV pages travel back (the reports embed the 80 MB module).
O=gpurun_out; T=/tmp/ncu_pr; mkdir -p $O $T
TAG=${1:-r2x}
PTE_PARAMS=test ncu --set full --clock-control none --import-source on -k regex:pte_rhot_sm_kernel --launch-skip 1 -c 1 -o $T/pte python tests/prof_driver.py pte 24 > $O/${TAG}_ncu_pte.out 2>&1
ncu --set full --clock-control none --import-source on -k regex:rhot_T_ --launch-skip 2 -c 2 -o $T/root python tests/prof_driver.py root 24 > $O/${TAG}_ncu_root.out 2>&1
for k in pte root; do
  ncu -i $T/$k.ncu-rep --page details > $O/${TAG}_${k}_details.txt
  ncu -i $T/$k.ncu-rep --page raw --csv > $O/${TAG}_${k}_raw.csv
done
grep -n "Duration\|Issue Slots Busy\|highest-utilized\|Avg. Active Threads\|Registers Per\|Local" $O/${TAG}_pte_details.txt $O/${TAG}_root_details.txt | cut -c1-160
