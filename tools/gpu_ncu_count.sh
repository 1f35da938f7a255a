set -x
ncu --set full --clock-control none --import-source on -k regex:count_local -s 1 -c 1 -o gpurun_out/count_full -f python bench.py --steps 1 --warmup 1 --chains 64 --proposals 10 --no-cpu > gpurun_out/ncu_count.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_count.log
