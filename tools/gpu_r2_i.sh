set -x
timeout 600 python -m pytest tests/test_gpu_chain.py -x -q > gpurun_out/r02_i_pytest_chain.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r02_i_pytest_chain.log
bash tools/gpu_variants.sh 4096 base la1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:chain_step_kernel -s 5 -c 1 -o gpurun_out/r02_i_chain -f python bench.py --steps 2 --warmup 3 --chains 4096 --proposals 500 --no-recount --no-cpu --no-clique-mix --no-configs > gpurun_out/ncu_chain_i.log 2>&1; echo "ncu rc=$?"
