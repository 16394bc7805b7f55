set -x
cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_gpu_sharded.py -x -q -W default -k "nccl and (2 or 8)" > gpurun_out/r2m8b_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2m8b_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
P=29800
for N in 8 4 2; do
  P=$((P+1)); timeout 300 $TR --nproc-per-node $N --master-port $P bench.py --gpus $N --no-cpu-baseline > gpurun_out/r2m8b_bench_cfg4_n$N.log 2>&1; echo "cfg4 n$N rc=$?"
  P=$((P+1)); timeout 300 $TR --nproc-per-node $N --master-port $P bench.py --gpus $N --config cfg3 --no-cpu-baseline > gpurun_out/r2m8b_bench_cfg3_n$N.log 2>&1; echo "cfg3 n$N rc=$?"
done
grep -h -o '"value": [0-9.]*, "unit": "samples/s", "n_gpus": [0-9]*' gpurun_out/r2m8b_bench_*.log
