set -x
for GB in 0.5 1 2 4 8 32 64; do ./tools/randsector $GB > gpurun_out/rs_$GB.json; python - <<PY
import json
d=json.load(open("gpurun_out/rs_$GB.json")); print("RANDSECTOR $GB GB: max %.1f G sectors/s; chase 28 warps %.0f ns, 1 warp %.0f ns" % (max(r["gsectors_per_s"] for r in d["rows"]), d["chase_ns_per_load_28warps"], d["chase_ns_per_load_1warp"]))
PY
done
