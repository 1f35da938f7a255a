# bench at N GPUs of one box (torchrun, one rank per GPU); $1 = N; the 2-device comm tests ride along
set -x
N=${1:-2}
timeout 300 python -m pytest tests/test_gpu_comm.py -x -q > gpurun_out/pytest_comm_n$N.log 2>&1; echo "pytest comm rc=$?"
tail -4 gpurun_out/pytest_comm_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_n$N.err
python - <<PY
import json
d=json.load(open("gpurun_out/bench_n$N.json"))
print("value", d["value"], "e2e", d["e2e"]["value"], "default_mix", d.get("default_mix",{}).get("value"))
r=d["recount"]; print({k:r[k] for k in ("ms_per_count","ms_per_count_kernel_only","ms_per_count_nccl","nccl_error","value")}, r["e2e"])
PY
