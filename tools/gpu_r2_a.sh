# round 2, call a: baseline data before any change (chain occupancy sweep, random-sector ceiling, ncu reports with source)
set -x
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv
./tools/randsector 16 > gpurun_out/r02_a_randsector.json 2> gpurun_out/randsector.err; echo "randsector rc=$?"
cat gpurun_out/r02_a_randsector.json
for C in 1024 2048 4096; do
  python bench.py --steps 5 --warmup 3 --chains $C --no-recount --no-cpu --no-clique-mix > gpurun_out/r02_a_chain_$C.json 2> gpurun_out/r02_a_chain_$C.err; echo "bench $C rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r02_a_chain_$C.json")); print($C, d["value"], d["ms_per_step"], d["e2e"]["value"])
PY
done
ncu --set full --clock-control none --import-source on -k regex:chain_step_kernel -s 5 -c 1 -o gpurun_out/r02_a_chain_full -f python bench.py --steps 2 --warmup 3 --chains 4096 --proposals 500 --no-recount --no-cpu --no-clique-mix > gpurun_out/ncu_chain.log 2>&1; echo "ncu chain rc=$?"
ncu --set full --clock-control none --import-source on -k regex:count_local_kernel -s 2 -c 1 -o gpurun_out/r02_a_count_full -f python bench.py --steps 1 --warmup 1 --chains 64 --proposals 10 --no-cpu --no-clique-mix > gpurun_out/ncu_count.log 2>&1; echo "ncu count rc=$?"
ls -la gpurun_out
