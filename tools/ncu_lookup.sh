#!/bin/bash
# ncu --set full on the headline lookup with and without lambda; the reports embed the whole
# 60 MB module, so only the text/CSV pages travel back.
O=gpurun_out; T=/tmp/ncu_lookup; mkdir -p $O $T
TAG=${1:-r2m}
ncu --set full --clock-control none --import-source on -k regex:eval_kernel --launch-skip 5 -c 1 -o $T/nolam python tests/prof_driver.py lookup 26 > $O/${TAG}_ncu_lookup.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:eval_kernel --launch-skip 1 -c 1 -o $T/lam python tests/prof_driver.py lookup 26 >> $O/${TAG}_ncu_lookup.log 2>&1
for k in lam nolam; do
  ncu -i $T/$k.ncu-rep --page details > $O/${TAG}_lookup_${k}_details.txt
  ncu -i $T/$k.ncu-rep --page raw --csv > $O/${TAG}_lookup_${k}_raw.csv
  ncu -i $T/$k.ncu-rep --page source --csv > $O/${TAG}_lookup_${k}_source.csv
done
ls -la $O
