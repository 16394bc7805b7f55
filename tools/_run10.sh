cd $GRAFT_REPO_ROOT
export ARCB_DEFAULT_FLAGS=0x100
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2l_pytest.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2l_bench_cfg4.log 2>&1; echo rc=$?; tail -1 gpurun_out/r2l_bench_cfg4.log | cut -c1-300
timeout 300 python bench.py --no-cpu-baseline --config cfg3 > gpurun_out/r2l_bench_cfg3.log 2>&1; echo rc=$?; tail -1 gpurun_out/r2l_bench_cfg3.log | cut -c1-300
unset ARCB_DEFAULT_FLAGS
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2l_bench_cfg4_default.log 2>&1; echo rc=$?; tail -1 gpurun_out/r2l_bench_cfg4_default.log | cut -c1-300
timeout 300 python bench.py --no-cpu-baseline --config cfg3 > gpurun_out/r2l_bench_cfg3_default.log 2>&1; echo rc=$?; tail -1 gpurun_out/r2l_bench_cfg3_default.log | cut -c1-300
