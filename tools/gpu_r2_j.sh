set -x
for CV in 100 default 70; do
  if [ "$CV" = default ]; then unset FCM_CARVEOUT; else export FCM_CARVEOUT=$CV; fi
  timeout 300 python bench.py --steps 5 --warmup 3 --no-recount --no-cpu --no-clique-mix --no-configs > gpurun_out/cv_$CV.json 2> gpurun_out/cv_$CV.err; echo "bench rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/cv_$CV.json")); print("CARVEOUT $CV: value %.1f M/s  ms/step %.3f  instr %.1f M/s" % (d["value"]/1e6, d["ms_per_step"], d["value_instrumented"]/1e6))
PY
done
unset FCM_CARVEOUT
timeout 300 python -m pytest tests/test_gpu_chain.py -x -q 2>&1 | tail -2
