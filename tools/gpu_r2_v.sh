set -x
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r02_v_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r02_v_pytest_gpu.log
