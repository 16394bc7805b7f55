set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,memory.total --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
timeout 400 python bench.py > gpurun_out/r2a_bench_cfg4.log 2>&1; echo "rc=$?"
timeout 200 python bench.py --config cfg3 > gpurun_out/r2a_bench_cfg3.log 2>&1; echo "rc=$?"
timeout 200 python bench.py --config cfg2 > gpurun_out/r2a_bench_cfg2.log 2>&1; echo "rc=$?"
BARGS="--steps 2 --warmup 3 --no-cpu-baseline --no-parity --sustained-seconds 0 --kernel-flags 0x80"
timeout 200 python bench.py $BARGS > gpurun_out/r2a_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2a_launches.csv python bench.py $BARGS > gpurun_out/r2a_ncu1.log 2>&1
echo "ncu1 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"bwd_persistent|gemm_sm100" -s 6 -c 2 -o gpurun_out/r2a_full python bench.py $BARGS > gpurun_out/r2a_ncu2.log 2>&1
echo "ncu2 rc=$?"
ls -la gpurun_out
