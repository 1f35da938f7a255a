# round 2, call g: the new index hot path (split flip/move, cp.async staging, exact look-ahead): parity tests, then speed
set -x
timeout 600 python -m pytest tests/test_gpu_chain.py tests/test_gpu_edge_cases.py -x -q > gpurun_out/r02_g_pytest_chain.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/r02_g_pytest_chain.log
for C in 4096 2048; do
timeout 300 python bench.py --steps 5 --warmup 3 --chains $C --no-recount --no-cpu --no-clique-mix --no-configs > gpurun_out/r02_g_chain_$C.json 2> gpurun_out/r02_g_chain_$C.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r02_g_chain_$C.json")); print("chains $C: value %.1f M/s  ms/step %.3f  e2e %.1f M/s instr %.1f spot %s" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["value_instrumented"]/1e6, d.get("parity_spot_check")))
PY
done
