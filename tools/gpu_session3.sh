#!/bin/bash
# smoke + parity suite + PTE / Helmholtz register-budget timings + both bench arms.
# usage: tools/gpu_session3.sh TAG
TAG=${1:-r2x}
O=gpurun_out
mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> $O/${TAG}_smoke.log
tail -3 $O/${TAG}_smoke.log
python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log
tail -8 $O/${TAG}_pytest.log
for p in test default; do
  PTE_PARAMS=$p python tests/prof_driver.py pte 24 >> $O/${TAG}_pte_timing.log 2>&1
  PTE_PARAMS=$p python tests/prof_driver.py pte 28 >> $O/${TAG}_pte_timing.log 2>&1
done
cat $O/${TAG}_pte_timing.log
for b in 3 4 5; do
  echo "== SGX_HELM_MINB=$b" >> $O/${TAG}_helm_minb.log
  SGX_HELM_MINB=$b python tests/prof_driver.py helm 26 >> $O/${TAG}_helm_minb.log 2>&1
done
cat $O/${TAG}_helm_minb.log
python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; tail -3 $O/${TAG}_bench.err
python bench.py --impl reference > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; tail -3 $O/${TAG}_bench_reference.err
cat $O/${TAG}_bench_reference.json
