#!/bin/bash
# Runs ON THE GPU BOX (under gpurun): one `ncu --set full` capture per hot kernel, exported to
# CSV/text so that only small files travel back in gpurun_out/.
# usage: tools/ncu_capture.sh <tag> [targets...]   targets: lookup root rhosie ideal pte helm
set -u
TAG=${1:-r1}; shift
TARGETS=${@:-lookup root rhosie ideal pte}
OUT=gpurun_out
mkdir -p $OUT
for t in $TARGETS; do
  case $t in
    # prof_driver launches one e(rho,T) first, then the target kernel 4x: -s 1 skips the former
    lookup) KRE='eval_kernel'; LG=${LOOKUP_LG:-24};;
    root)   KRE='eval_kernel|rhot_T_from'; LG=24;;
    rhosie) KRE='eval_kernel'; LG=24;;
    ideal)  KRE='eval_kernel'; LG=24;;
    pte)    KRE='pte_'; LG=22;;
    helm)   KRE='helm_'; LG=22;;
  esac
  python tests/prof_driver.py $t $LG > $OUT/prof_${t}_${TAG}_plain.log 2>&1 || { echo "plain run of $t failed"; tail -5 $OUT/prof_${t}_${TAG}_plain.log; continue; }
  ncu --set full --clock-control none --import-source on -k regex:"$KRE" -s 1 -c 1 -f \
      -o $OUT/prof_${t}_${TAG} python tests/prof_driver.py $t $LG > $OUT/prof_${t}_${TAG}_ncu.log 2>&1
  REP=$OUT/prof_${t}_${TAG}.ncu-rep
  if [ -f $REP ]; then
    ncu -i $REP --page raw --csv > $OUT/prof_${t}_${TAG}_raw.csv 2>/dev/null
    ncu -i $REP --page details > $OUT/prof_${t}_${TAG}_details.txt 2>/dev/null
    ncu -i $REP --page source --csv > $OUT/prof_${t}_${TAG}_source.csv 2>/dev/null
    SZ=$(stat -c %s $REP)
    if [ $SZ -gt 12000000 ]; then rm -f $REP; echo "$t: report $SZ bytes dropped (CSV kept)"; fi
  else
    echo "$t: no report"; tail -5 $OUT/prof_${t}_${TAG}_ncu.log
  fi
  cat $OUT/prof_${t}_${TAG}_plain.log
done
du -sh $OUT
