# clique-move parity tests, then the default-mix probes
set -x
timeout 900 python -m pytest tests/test_gpu_chain.py tests/test_gpu_edge_cases.py -m gpu -x -q -k "clique or default_move_mix or tiny or checkpoint or edgebits" 2>&1 | tail -5
timeout 300 python tools/clique_probe_nostats.py c3_er_2000 4096 200 2>&1 | tail -1
timeout 300 python tools/clique_probe_nostats.py c2_chunglu_279 4096 200 2>&1 | tail -1
