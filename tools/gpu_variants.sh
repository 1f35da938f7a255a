# A/B of experiment builds (csrc/Makefile VARIANT=...): $1 = chains, rest = variant names ("base" = libfcm.so)
set -x
C=$1; shift
for V in "$@"; do
  if [ "$V" = base ]; then L=""; else L="$PWD/flag_complex_mcmc_b200/libfcm_$V.so"; fi
  FCM_LIB=$L timeout 300 python bench.py --steps 5 --warmup 3 --chains $C --no-recount --no-cpu --no-clique-mix --no-configs > gpurun_out/var_${V}_$C.json 2> gpurun_out/var_${V}_$C.err; echo "bench $V rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/var_${V}_$C.json")); print("VARIANT $V chains $C: value %.1f M/s  ms/step %.3f  e2e %.1f M/s" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6))
PY
done
