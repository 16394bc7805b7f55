set -x
cd $GRAFT_REPO_ROOT
timeout 200 python -m pytest tests/test_gpu_sharded.py -x -q -W default -k "nccl and 8" > gpurun_out/run_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/run_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 8 --master-port 29801 bench.py --gpus 8 --no-cpu-baseline > gpurun_out/run_bench_cfg4_n8.log 2>&1; echo "cfg4 n8 rc=$?"
timeout 200 $TR --nproc-per-node 8 --master-port 29802 bench.py --gpus 8 --config cfg3 --no-cpu-baseline > gpurun_out/run_bench_cfg3_n8.log 2>&1; echo "cfg3 n8 rc=$?"
timeout 200 $TR --nproc-per-node 4 --master-port 29803 bench.py --gpus 4 --no-cpu-baseline > gpurun_out/run_bench_cfg4_n4.log 2>&1; echo "cfg4 n4 rc=$?"
timeout 200 $TR --nproc-per-node 4 --master-port 29804 bench.py --gpus 4 --config cfg3 --no-cpu-baseline > gpurun_out/run_bench_cfg3_n4.log 2>&1; echo "cfg3 n4 rc=$?"
grep -h -o '"value": [0-9.]*, "unit": "samples/s", "n_gpus": [0-9]*' gpurun_out/run_bench_*.log
