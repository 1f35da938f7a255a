# round 2, call b: pair-order A/B on the C5 shard + full GPU test-suite (comm tests included)
set -x
for PO in 0 1; do
  FCM_PAIR_ORDER=$PO python bench.py --steps 5 --warmup 3 --no-recount --no-cpu --no-clique-mix > gpurun_out/r02_b_pairorder_$PO.json 2> gpurun_out/r02_b_pairorder_$PO.err; echo "bench PO=$PO rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r02_b_pairorder_$PO.json")); print("PO=$PO", d["value"], d["ms_per_step"], d["e2e"]["value"])
PY
done
python -m pytest tests -m gpu -x -q > gpurun_out/r02_b_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r02_b_pytest_gpu.log
