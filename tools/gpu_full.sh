# the round's standard check: GPU test-suite, smoke, default bench line, reference arm; $1 = tag for the output names
set -x
T=${1:-x}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/${T}_pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench rc=$?"
tail -2 gpurun_out/${T}_bench_n1.err
python - <<PY
import json
d=json.load(open("gpurun_out/${T}_bench_n1.json"))
print("BENCH value %.1f M/s e2e %.1f M/s frac %.3f rand %.3f spot %s | default_mix %.1f M/s create %.2fs | recount %.3f ms (kernel %.3f) e2e %.2f ms" % (d["value"]/1e6, d["e2e"]["value"]/1e6, d["roofline"]["frac"], (d["roofline"]["random_access"] or {}).get("frac", 0), d.get("parity_spot_check"), d["default_mix"]["value"]/1e6, d["default_mix"]["create_s"], d["recount"]["ms_per_count"], d["recount"]["ms_per_count_kernel_only"], d["recount"]["e2e"]["seconds"]*1e3))
for k,v in d.get("configs",{}).items(): print(k, {m: (round(v[m]["value"]), round(v[m].get("cpu_baseline",{}).get("value",0))) for m in ("simple","default_mix")})
PY
