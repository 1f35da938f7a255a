set -x
./tools/randsector 16 32 > gpurun_out/r02_m_randsector_32.json; tail -c 300 gpurun_out/r02_m_randsector_32.json
./tools/randsector 16 64 > gpurun_out/r02_m_randsector_64.json; tail -c 300 gpurun_out/r02_m_randsector_64.json
for FG in 64 32; do for LP in 1 0; do
  FCM_L2_FETCH=$FG FCM_L2_PERSIST=$LP timeout 300 python bench.py --steps 5 --warmup 3 --no-recount --no-cpu --no-clique-mix --no-configs > gpurun_out/fg_${FG}_$LP.json 2> gpurun_out/fg_${FG}_$LP.err; echo "bench rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/fg_${FG}_$LP.json")); print("FETCH $FG PERSIST $LP: value %.1f M/s  ms/step %.3f  instr %.1f M/s" % (d["value"]/1e6, d["ms_per_step"], d["value_instrumented"]/1e6))
PY
done; done
timeout 600 python -m pytest tests/test_gpu_chain.py -x -q 2>&1 | tail -2
timeout 300 python -m pytest tests/test_gpu_count.py tests/test_gpu_comm.py -x -q 2>&1 | tail -2
FCM_TIMING=1 timeout 120 python tools/count_e2e_probe.py 2>&1 | tail -12
