# ncu --set full of the chain kernel under the CLI's default move mix (clique moves on the cold path);
# launch 0 = warm-up step, launch 1 = the timed one
set -x
timeout 300 python tools/clique_probe_nostats.py c3_er_2000 4096 200 > gpurun_out/clique_plain.log 2>&1; echo "plain rc=$?"
tail -1 gpurun_out/clique_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:chain_step_kernel -s 1 -c 1 -o gpurun_out/clique_full -f python tools/clique_probe_nostats.py c3_er_2000 4096 200 > gpurun_out/ncu_clique.log 2>&1; echo "ncu rc=$?"
grep -i "profiling\|error" gpurun_out/ncu_clique.log | head -3
