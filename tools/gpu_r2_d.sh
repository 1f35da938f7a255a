# round 2, call d: new tests (limits, threads, C4 chains, dense for_each_cell, checkpoint refusal) + the full default bench line
set -x
timeout 400 python -m pytest tests/test_gpu_count.py tests/test_gpu_comm.py "tests/test_gpu_chain.py::test_checkpoint_restore_resumes_exactly" -x -q > gpurun_out/r02_d_pytest_new.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r02_d_pytest_new.log
timeout 600 python bench.py > gpurun_out/r02_d_bench_n1.json 2> gpurun_out/r02_d_bench_n1.err; echo "bench rc=$?"
tail -5 gpurun_out/r02_d_bench_n1.err
timeout 120 ncu --metrics dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum -k regex:rand_kernel -c 3 ./tools/randsector 4 > gpurun_out/r02_d_randsector_ncu.log 2>&1; echo "ncu rc=$?"
