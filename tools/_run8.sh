cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "stored_probabilities" > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/r2h_pytest.log
timeout 200 python tools/gpu_time_bwd.py --flags 0x100 > gpurun_out/r2h_time_keep.log 2>&1; echo rc=$?; tail -25 gpurun_out/r2h_time_keep.log
timeout 200 python tools/gpu_time_bwd.py > gpurun_out/r2h_time_default.log 2>&1; echo rc=$?; tail -25 gpurun_out/r2h_time_default.log
timeout 300 python bench.py --kernel-flags 0x100 --no-cpu-baseline > gpurun_out/r2h_bench_keep.log 2>&1; echo rc=$?; tail -3 gpurun_out/r2h_bench_keep.log
