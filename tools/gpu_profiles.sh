# ncu evidence for profiles/: launch list of the default bench, ncu --set full of the timed kernels; $1 = tag
set -x
T=${1:-x}
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_default.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/${T}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:chain_step_kernel -s 5 -c 1 -o gpurun_out/${T}_chain_step -f python bench.py --steps 2 --warmup 3 --chains 4096 --proposals 500 --no-recount --no-cpu --no-clique-mix --no-configs > gpurun_out/${T}_ncu_chain.log 2>&1; echo "ncu chain rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:count_local_kernel -s 2 -c 1 -o gpurun_out/${T}_count_local -f python bench.py --steps 1 --warmup 1 --chains 64 --proposals 10 --no-cpu --no-clique-mix --no-configs > gpurun_out/${T}_ncu_count.log 2>&1; echo "ncu count rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:chain_step_kernel -s 1 -c 1 -o gpurun_out/${T}_chain_default_mix -f python tools/clique_mix_probe.py c3_er_2000 4096 100 > gpurun_out/${T}_ncu_mix.log 2>&1; echo "ncu mix rc=$?"
ls -la gpurun_out | grep $T
