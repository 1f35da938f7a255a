set -x
timeout 900 python -m pytest tests/test_gpu_chain.py tests/test_gpu_edge_cases.py tests/test_gpu_limits_threads.py tests/test_cpp_host.py -x -q -m gpu > gpurun_out/r02_l_pytest.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/r02_l_pytest.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-recount --no-cpu --no-configs > gpurun_out/r02_l.json 2> gpurun_out/r02_l.err; echo "bench rc=$?"
tail -3 gpurun_out/r02_l.err
python - <<PY
import json
d=json.load(open("gpurun_out/r02_l.json")); print("RESULT: value %.1f M/s  ms/step %.3f  instr %.1f M/s default_mix %.1f M/s bytes/chain %d" % (d["value"]/1e6, d["ms_per_step"], d["value_instrumented"]/1e6, d["default_mix"]["value"]/1e6, d["bytes_per_chain"]))
PY
timeout 300 ncu --set full --clock-control none --import-source on -k regex:chain_step_kernel -s 5 -c 1 -o gpurun_out/r02_l_chain -f python bench.py --steps 2 --warmup 3 --chains 4096 --proposals 500 --no-recount --no-cpu --no-clique-mix --no-configs > gpurun_out/ncu_chain_l.log 2>&1; echo "ncu rc=$?"
