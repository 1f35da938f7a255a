# ncu --set full of the uninstrumented chain kernel (launches 0-3 of the bench are the instrumented <1,1>)
set -x
ncu --set full --clock-control none --import-source on -k regex:chain_step_kernel -s 5 -c 1 -o gpurun_out/chain_full -f python bench.py --steps 2 --warmup 3 --chains 4096 --proposals 500 --no-recount --no-cpu --no-clique-mix > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
grep -i "profiling\|error" gpurun_out/ncu_full.log | head -3
