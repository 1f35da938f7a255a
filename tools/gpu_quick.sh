# quick A/B point: chain leg (simple + default mix) of the bench, no CPU legs; $1 = tag
T=${1:-q}
timeout 300 python bench.py --steps 5 --warmup 3 --no-recount --no-cpu --no-configs > gpurun_out/${T}.json 2> gpurun_out/${T}.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/${T}.json")); print("RESULT ${T}: simple %.1f M/s (%.3f ms/step)  default_mix %.1f M/s (%.3f ms/step)" % (d["value"]/1e6, d["ms_per_step"], d["default_mix"]["value"]/1e6, d["default_mix"]["ms_per_step"]))
PY
