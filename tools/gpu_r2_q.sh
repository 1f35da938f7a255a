set -x
timeout 900 python -m pytest tests/test_gpu_chain.py tests/test_gpu_edge_cases.py tests/test_gpu_limits_threads.py -x -q -k "not c4_sized" > gpurun_out/r02_q_pytest.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/r02_q_pytest.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-recount --no-cpu --no-configs > gpurun_out/r02_q.json 2> gpurun_out/r02_q.err; echo "bench rc=$?"
tail -3 gpurun_out/r02_q.err
python - <<PY
import json
d=json.load(open("gpurun_out/r02_q.json")); print("RESULT: value %.1f M/s  default_mix %.1f M/s (%.3f ms/step, create %.2f s)" % (d["value"]/1e6, d["default_mix"]["value"]/1e6, d["default_mix"]["ms_per_step"], d["default_mix"]["create_s"]))
PY
