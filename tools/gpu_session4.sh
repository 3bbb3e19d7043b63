#!/bin/bash
# parity suite (+ the no-alloc test alone in a fresh process) + lookup shape study + per-kernel
# timings + ncu of the lookups + both bench arms.  usage: tools/gpu_session4.sh TAG
TAG=${1:-r2x}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log
tail -4 $O/${TAG}_pytest.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/${TAG}_pytest_noalloc_alone.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_noalloc_alone.log
tail -3 $O/${TAG}_pytest_noalloc_alone.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> $O/${TAG}_smoke.log
tail -2 $O/${TAG}_smoke.log
SGX_STUDY_GATHER=1 python tools/studies/lookup_shapes.py 26 > $O/${TAG}_lookup_shapes.json 2> $O/${TAG}_shapes.err
python - <<PY
import json
d = json.load(open("$O/${TAG}_lookup_shapes.json"))
print("product ms", d["product_lambda_ms"], d["product_nolambda_ms"])
for k, v in d["shapes"].items():
    print(k, v["bit_identical_to_sgx_eos_eval"], [(round(x["lambda_Gcells_s"], 1), round(x["nolambda_Gcells_s"], 1)) for kk, x in v.items() if kk.startswith("grid")])
print(d.get("gather_cost_product_kernel"))
PY
python tests/prof_driver.py all 26 > $O/${TAG}_prof_all.log 2>&1; cat $O/${TAG}_prof_all.log
python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; tail -3 $O/${TAG}_bench.err
python bench.py --impl reference > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; tail -2 $O/${TAG}_bench_reference.err
bash tools/ncu_lookup.sh ${TAG} > $O/${TAG}_ncu_ls.log 2>&1; tail -3 $O/${TAG}_ncu_ls.log
