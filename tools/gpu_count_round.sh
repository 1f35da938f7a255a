set -x
python -m pytest tests/test_gpu_count.py -m gpu -x -q > gpurun_out/pytest_count.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_count.log
python bench.py --steps 3 --warmup 2 --chains 256 --proposals 100 --no-cpu > gpurun_out/bench_count.json 2> gpurun_out/bench_count.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_count.err
python -c "
import json; d=json.load(open('gpurun_out/bench_count.json'))['recount']; print(d['value'], d['ms_per_count'], d['flag_count'], d['e2e'])"
if [ "$1" = "ncu" ]; then
ncu --set full --clock-control none --import-source on -k regex:count_local_kernel -s 3 -c 1 -o gpurun_out/count_full -f python bench.py --steps 1 --warmup 1 --chains 64 --proposals 10 --no-cpu > gpurun_out/ncu_count.log 2>&1; echo "ncu rc=$?"
grep -i "count_local" gpurun_out/ncu_count.log | head
fi
