set -x
cd $GRAFT_REPO_ROOT
timeout 1200 python -m pytest tests -q -m gpu > gpurun_out/run_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/run_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/run_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/run_smoke.log
timeout 400 python bench.py > gpurun_out/run_bench_cfg4.log 2>&1; echo rc=$?; tail -1 gpurun_out/run_bench_cfg4.log | cut -c1-300
timeout 300 python bench.py --no-cpu-baseline --config cfg3 > gpurun_out/run_bench_cfg3.log 2>&1; echo rc=$?; tail -1 gpurun_out/run_bench_cfg3.log | cut -c1-300
timeout 300 python bench.py --no-cpu-baseline --config cfg2 > gpurun_out/run_bench_cfg2.log 2>&1; echo rc=$?; tail -1 gpurun_out/run_bench_cfg2.log | cut -c1-300
BARGS="--steps 2 --warmup 3 --no-cpu-baseline --no-parity --sustained-seconds 0 --kernel-flags 0x80"
timeout 200 python bench.py $BARGS > gpurun_out/run_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"bwd_|fwd_|l2norm|loss_|merge_|target_prep|gemm|probs_" -c 120 --csv --log-file gpurun_out/run_launches.csv python bench.py $BARGS > gpurun_out/run_ncu1.log 2>&1
echo "ncu1 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gemm2_sm100|gemm_sm100" -s 9 -c 3 -o gpurun_out/run_full python bench.py $BARGS > gpurun_out/run_ncu2.log 2>&1
echo "ncu2 rc=$?"
ls -la gpurun_out | tail -12
