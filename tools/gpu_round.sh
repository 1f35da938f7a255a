# one GPU round: tests, bench, reference arm, ncu launch list + full capture of the chain kernel
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 --no-recount --no-cpu > gpurun_out/bench_chain.json 2> gpurun_out/bench_chain.err; echo "bench rc=$?"
cat gpurun_out/bench_chain.json
if [ "$1" = "ncu" ]; then
ncu --set full --clock-control none --import-source on -k regex:chain_step -s 3 -c 1 -o gpurun_out/chain_full -f python bench.py --steps 2 --warmup 3 --chains 4096 --proposals 500 --no-recount --no-cpu > gpurun_out/ncu_full.log 2>&1
fi
