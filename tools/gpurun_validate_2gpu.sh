set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_trainer.py -x -q -W default > gpurun_out/run_pytest2.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/run_pytest2.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 2 --master-port 29611 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/run_bench_cfg4_n2.log 2>&1; echo "rc=$?"
timeout 300 $TR --nproc-per-node 2 --master-port 29613 bench.py --gpus 2 --no-cpu-baseline --config cfg3 > gpurun_out/run_bench_cfg3_n2.log 2>&1; echo "rc=$?"
for f in gpurun_out/run_bench_cfg4_n2.log gpurun_out/run_bench_cfg3_n2.log; do echo $f; tail -1 $f | grep -o '"value": [0-9.]*, "unit": "samples/s", "n_gpus": [0-9]*\|"ms_per_step": [0-9.]*\|"loss_rel": [0-9.e-]*\|"cos_dx": [0-9.]*\|"kernels_ms_per_step": {[^}]*}' | head -6; done
