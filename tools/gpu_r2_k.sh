set -x
timeout 300 python bench.py --steps 5 --warmup 3 --no-recount --no-cpu --no-clique-mix --no-configs > gpurun_out/r02_k.json 2> gpurun_out/r02_k.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r02_k.json")); print("RESULT: value %.1f M/s  ms/step %.3f  instr %.1f M/s" % (d["value"]/1e6, d["ms_per_step"], d["value_instrumented"]/1e6))
PY
timeout 300 ncu --set full --clock-control none --import-source on -k regex:chain_step_kernel -s 5 -c 1 -o gpurun_out/r02_k_chain -f python bench.py --steps 2 --warmup 3 --chains 4096 --proposals 500 --no-recount --no-cpu --no-clique-mix --no-configs > gpurun_out/ncu_chain_k.log 2>&1; echo "ncu rc=$?"
