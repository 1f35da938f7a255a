set -x
bash tools/gpu_variants.sh 4096 base nola nola2
timeout 300 ncu --set full --clock-control none --import-source on -k regex:chain_step_kernel -s 5 -c 1 -o gpurun_out/r02_h_chain -f python bench.py --steps 2 --warmup 3 --chains 4096 --proposals 500 --no-recount --no-cpu --no-clique-mix --no-configs > gpurun_out/ncu_chain_h.log 2>&1; echo "ncu rc=$?"
