#!/bin/bash
# parity suite + PTE variant timings + bench.  usage: tools/gpu_session2.sh TAG
TAG=${1:-r2x}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log
tail -5 $O/${TAG}_pytest.log
for v in sm percell warp; do
  SGX_PTE_KERNEL=$v PTE_PARAMS=test python tests/prof_driver.py pte 24 >> $O/${TAG}_pte_variants.log 2>&1
done
cat $O/${TAG}_pte_variants.log
for mode in exact hybrid direct; do
  echo "== SGX_ROOT_MODE=$mode" >> $O/${TAG}_root_timing.log
  SGX_ROOT_MODE=$mode python tests/prof_driver.py root 26 >> $O/${TAG}_root_timing.log 2>&1
done
cat $O/${TAG}_root_timing.log
python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; tail -3 $O/${TAG}_bench.err
python tools/pcie_bw.py > $O/${TAG}_pcie_bw_n1.json 2>&1; cat $O/${TAG}_pcie_bw_n1.json
