#!/bin/bash
# One GPU-box session: parity suite, root-find timings per mode, ncu captures.  usage: tools/gpu_session.sh TAG
TAG=${1:-r2x}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log
tail -5 $O/${TAG}_pytest.log
for mode in exact hybrid direct; do
  echo "== SGX_ROOT_MODE=$mode" >> $O/${TAG}_root_timing.log
  SGX_ROOT_MODE=$mode python tests/prof_driver.py root 26 >> $O/${TAG}_root_timing.log 2>&1
done
cat $O/${TAG}_root_timing.log
