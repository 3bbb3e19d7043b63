#!/bin/bash
# ncu --set full of the warp-per-cell PTE comparison variant (SGX_PTE_KERNEL=warp), to sit next to
# the state machine's capture (tools/ncu_pte_root.sh).  2^18 cells: the variant is ~36x slower.
O=gpurun_out; T=/tmp/ncu_w; mkdir -p $O $T; TAG=${1:-r2x}
SGX_PTE_KERNEL=warp PTE_PARAMS=test ncu --set full --clock-control none -k regex:pte_rhot_warp_kernel --launch-skip 1 -c 1 -o $T/warp python tests/prof_driver.py pte 22 > $O/${TAG}_ncu_pte_warp.out 2>&1
ncu -i $T/warp.ncu-rep --page details > $O/${TAG}_pte_warp_details.txt
grep -n "Duration\|Issue Slots Busy\|highest-utilized\|Avg. Active Threads\|Avg. Not Predicated\|Registers Per\|Achieved Occupancy" $O/${TAG}_pte_warp_details.txt | cut -c1-150
