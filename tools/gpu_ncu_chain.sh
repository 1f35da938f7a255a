set -x
ncu --set full --clock-control none --import-source on -k regex:"chain_step_kernel<1, 0>" -s 1 -c 1 -o gpurun_out/chain_full -f python bench.py --steps 2 --warmup 3 --chains 4096 --proposals 500 --no-recount --no-cpu --no-clique-mix > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/ncu_full.log | cut -c1-200
