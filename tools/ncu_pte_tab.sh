#!/bin/bash
# timing (state machine vs thread-per-cell) and ncu --set full of the tabulated-mixture PTE kernel
O=gpurun_out; T=/tmp/ncu_t; mkdir -p $O $T; TAG=${1:-r2x}
for v in sm percell; do SGX_PTE_KERNEL=$v python tests/prof_driver.py ptetab 26 2>&1 | grep pte_rhot >> $O/${TAG}_pte_tab_timing.txt; done
cat $O/${TAG}_pte_tab_timing.txt
ncu --set full --clock-control none -k regex:pte_rhot_sm_kernel --launch-skip 1 -c 1 -o $T/tab python tests/prof_driver.py ptetab 24 > $O/${TAG}_ncu_pte_tab.out 2>&1
ncu -i $T/tab.ncu-rep --page details > $O/${TAG}_pte_tab_details.txt
grep -n "Duration\|Issue Slots Busy\|highest-utilized\|Avg. Active Threads\|Registers Per\|L1/TEX Hit\|L2 Hit" $O/${TAG}_pte_tab_details.txt | cut -c1-150
