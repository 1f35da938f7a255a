set -x
FCM_TIMING=1 timeout 300 python tools/create_probe.py 2>&1 | grep -E "ensemble_create|simple=" | tail -12
timeout 900 python -m pytest tests/test_gpu_chain.py tests/test_gpu_edge_cases.py tests/test_gpu_limits_threads.py -x -q -k "not c4_sized" 2>&1 | tail -3
