# round 2, call c: GPU test-suite again (comm fixes) + ncu of the chain kernel under both pair orders (DRAM bytes per proposal)
set -x
python -m pytest tests/test_gpu_comm.py -x -q > gpurun_out/r02_c_pytest_comm.log 2>&1; echo "pytest comm rc=$?"
tail -5 gpurun_out/r02_c_pytest_comm.log
for PO in 0 1; do
  FCM_PAIR_ORDER=$PO ncu --set full --clock-control none --import-source on -k regex:chain_step_kernel -s 5 -c 1 -o gpurun_out/r02_c_chain_po$PO -f python bench.py --steps 2 --warmup 3 --chains 4096 --proposals 500 --no-recount --no-cpu --no-clique-mix > gpurun_out/ncu_chain_po$PO.log 2>&1; echo "ncu rc=$?"
done
python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_comm.py > gpurun_out/r02_c_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r02_c_pytest_gpu.log
