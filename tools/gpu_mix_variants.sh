# A/B of experiment builds on BOTH chain legs (simple moves and the default move mix): args = variant names ("base" = libfcm.so)
for V in "$@"; do
  if [ "$V" = base ]; then L=""; else L="$PWD/flag_complex_mcmc_b200/libfcm_$V.so"; fi
  FCM_LIB=$L timeout 300 python bench.py --steps 5 --warmup 3 --no-recount --no-cpu --no-configs > gpurun_out/mix_${V}.json 2> gpurun_out/mix_${V}.err; echo "bench $V rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/mix_${V}.json")); print("VARIANT ${V}: simple %.1f M/s (%.3f ms/step)  default_mix %.1f M/s (%.3f ms/step)" % (d["value"]/1e6, d["ms_per_step"], d["default_mix"]["value"]/1e6, d["default_mix"]["ms_per_step"]))
PY
done
