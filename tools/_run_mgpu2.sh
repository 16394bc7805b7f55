set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=index,name --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2m2_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2m2_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 2 --master-port 29611 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/r2m2_bench_cfg4_n2.log 2>&1; echo "rc=$?"; tail -c 1500 gpurun_out/r2m2_bench_cfg4_n2.log
timeout 300 $TR --nproc-per-node 2 --master-port 29612 bench.py --gpus 2 --config cfg3 --no-cpu-baseline > gpurun_out/r2m2_bench_cfg3_n2.log 2>&1; echo "rc=$?"
for head in replicated sharded; do
  timeout 300 python examples/e2e_train.py --head $head --classes 10000 --steps 10 --warmup 3 --out gpurun_out/r2m2_e2e.jsonl > gpurun_out/r2m2_e2e_1_$head.log 2>&1; echo "rc=$?"
  timeout 300 $TR --nproc-per-node 2 --master-port 29613 examples/e2e_train.py --head $head --classes 10000 --steps 10 --warmup 3 --out gpurun_out/r2m2_e2e.jsonl > gpurun_out/r2m2_e2e_2_$head.log 2>&1; echo "rc=$?"
done
cat gpurun_out/r2m2_e2e.jsonl
