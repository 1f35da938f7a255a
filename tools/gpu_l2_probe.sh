# Does the clique index stay in L2?  Simple-move rate and L2 / DRAM sector counts of chain_step_kernel (4096 chains x 500
# proposals) for: the default build, the access-policy window off (FCM_L2_PERSIST=0), and experiment builds given as args.
M=dram__sectors_read.sum,dram__sectors_write.sum,lts__t_sectors_srcunit_tex_op_read_lookup_hit.sum,lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum,lts__t_sectors_evict_last_lookup_hit.sum,lts__t_sectors_evict_last_lookup_miss.sum,lts__t_sectors_evict_first_lookup_hit.sum,lts__t_sectors_evict_first_lookup_miss.sum,gpu__time_duration.sum
run() {   # $1 = tag, $2 = lib ("" = default), $3 = FCM_L2_PERSIST value ("" = unset)
  local L=""; [ -n "$2" ] && L="$PWD/flag_complex_mcmc_b200/libfcm_$2.so"
  local P=""; [ -n "$3" ] && P="FCM_L2_PERSIST=$3"
  env FCM_LIB=$L $P timeout 200 python bench.py --steps 5 --warmup 3 --no-recount --no-cpu --no-configs --no-clique-mix > gpurun_out/l2_$1.json 2> gpurun_out/l2_$1.err
  python -c "
import json
d=json.load(open('gpurun_out/l2_$1.json')); print('RATE $1: %.1f M/s (%.3f ms/step)' % (d['value']/1e6, d['ms_per_step']))"
  env FCM_LIB=$L $P timeout 200 ncu --metrics $M --clock-control none -k regex:chain_step_kernel -s 5 -c 1 --csv --log-file gpurun_out/l2_$1.csv python bench.py --steps 2 --warmup 3 --chains 4096 --proposals 500 --no-recount --no-cpu --no-clique-mix --no-configs > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/l2_$1.csv")) if len(r)>5]
h=rows[0]; i=h.index("Metric Name"); v=h.index("Metric Value")
P=4096*500
print("NCU $1: " + "  ".join("%s=%.1f" % (r[i].replace("lts__t_sectors_","").replace(".sum",""), float(r[v].replace(",",""))/(1 if "time" in r[i] else P)) for r in rows[1:]))
PY
}
run base "" ""
run nowin "" 0
for V in "$@"; do run $V $V ""; done
