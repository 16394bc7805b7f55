set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2e_pytest.log
timeout 300 python -m pytest tests/test_gpu_fullsize.py -q -s -k "eager_fp32_port" > gpurun_out/r2e_fullsize.log 2>&1; grep "fullsize" gpurun_out/r2e_fullsize.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2e_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2e_smoke.log
timeout 400 python bench.py > gpurun_out/r2e_bench_cfg4.log 2>&1; echo "bench rc=$?"
timeout 200 python bench.py --config cfg3 > gpurun_out/r2e_bench_cfg3.log 2>&1; echo "rc=$?"
timeout 200 python bench.py --config cfg2 > gpurun_out/r2e_bench_cfg2.log 2>&1; echo "rc=$?"
BARGS="--steps 2 --warmup 3 --no-cpu-baseline --no-parity --sustained-seconds 0 --kernel-flags 0x80"
timeout 200 python bench.py $BARGS > gpurun_out/r2e_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:arcb -c 120 --csv --log-file gpurun_out/r2e_launches.csv python bench.py $BARGS > gpurun_out/r2e_ncu1.log 2>&1
echo "ncu1 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"bwd_persistent|fwd_resident" -s 6 -c 2 -o gpurun_out/r2e_full python bench.py $BARGS > gpurun_out/r2e_ncu2.log 2>&1
echo "ncu2 rc=$?"
