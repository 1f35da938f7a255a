set -x
python -m pytest tests -m gpu -x -q -k "edgebits" 2>&1 | tail -8
ncu --set full --clock-control none --import-source on -k regex:count_local_kernel -s 3 -c 1 -o gpurun_out/count_full -f python bench.py --steps 1 --warmup 1 --chains 64 --proposals 10 --no-cpu > gpurun_out/ncu_count.log 2>&1; echo "ncu rc=$?"
