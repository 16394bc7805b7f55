set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_trainer.py -x -q -W default > gpurun_out/r2m2b_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r2m2b_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 2 --master-port 29611 bench.py --gpus 2 --no-cpu-baseline --classes 250000 > gpurun_out/r2m2b_bench_peer.log 2>&1; echo "rc=$?"; tail -c 600 gpurun_out/r2m2b_bench_peer.log
ARCB_PEER_EXCHANGE=0 timeout 300 $TR --nproc-per-node 2 --master-port 29612 bench.py --gpus 2 --no-cpu-baseline --classes 250000 > gpurun_out/r2m2b_bench_nccl.log 2>&1; echo "rc=$?"
grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2m2b_bench_peer.log | head -1; grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2m2b_bench_nccl.log | head -1
