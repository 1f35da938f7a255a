set -x
timeout 900 python -m pytest tests/test_gpu_chain.py tests/test_gpu_edge_cases.py tests/test_gpu_limits_threads.py -x -q -k "not c4_sized" > gpurun_out/r02_r_pytest.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/r02_r_pytest.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-recount > gpurun_out/r02_r.json 2> gpurun_out/r02_r.err; echo "bench rc=$?"
tail -3 gpurun_out/r02_r.err
python - <<PY
import json
d=json.load(open("gpurun_out/r02_r.json")); print("RESULT: value %.1f M/s  default_mix %.1f M/s spot %s" % (d["value"]/1e6, d["default_mix"]["value"]/1e6, d.get("parity_spot_check")))
for k,v in d.get("configs",{}).items(): print(k, {m: (round(v[m]["value"]), round(v[m].get("cpu_baseline",{}).get("value",0))) for m in ("simple","default_mix")})
PY
