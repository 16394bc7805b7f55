set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 600 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_trainer.py -x -q -W default -k "nccl or trainers" > gpurun_out/r2m8_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2m8_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
P=29700
for N in 8 4; do
  P=$((P+1)); timeout 300 $TR --nproc-per-node $N --master-port $P bench.py --gpus $N --no-cpu-baseline > gpurun_out/r2m8_bench_cfg4_n$N.log 2>&1; echo "cfg4 n$N rc=$?"
  P=$((P+1)); timeout 300 $TR --nproc-per-node $N --master-port $P bench.py --gpus $N --config cfg3 --no-cpu-baseline > gpurun_out/r2m8_bench_cfg3_n$N.log 2>&1; echo "cfg3 n$N rc=$?"
done
P=$((P+1)); ARCB_PEER_EXCHANGE=0 timeout 300 $TR --nproc-per-node 8 --master-port $P bench.py --gpus 8 --no-cpu-baseline --sustained-seconds 0 --no-parity > gpurun_out/r2m8_bench_cfg4_n8_nccl.log 2>&1; echo "cfg4 n8 nccl rc=$?"
for N in 8 4; do
  for head in replicated sharded; do
    P=$((P+1)); timeout 300 $TR --nproc-per-node $N --master-port $P examples/e2e_train.py --head $head --classes 10000 --steps 8 --warmup 3 --out gpurun_out/r2m8_e2e.jsonl > gpurun_out/r2m8_e2e_${N}_$head.log 2>&1; echo "e2e $N $head rc=$?"
  done
done
cat gpurun_out/r2m8_e2e.jsonl
grep -h -o '"value": [0-9.]*, "unit": "samples/s", "n_gpus": [0-9]*' gpurun_out/r2m8_bench_*.log
