set -x
timeout 300 python tools/dense_probe.py 2>&1 | tail -3
timeout 200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,lts__t_sector_hit_rate.pct -k regex:count_dense -c 1 python tools/dense_probe.py > gpurun_out/r02_s_dense_ncu.log 2>&1; echo "ncu rc=$?"
grep -E "count_dense|gpu__time|dram__bytes|lts__t|l1tex__t|issue_active|inst_executed" gpurun_out/r02_s_dense_ncu.log | head -12
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_s_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r02_s_pytest_gpu.log
