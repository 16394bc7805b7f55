cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "stored_probabilities" > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2k_pytest.log
for f in 0x100; do for m in 0 8 16 24 0 24; do echo "flags $f debug $m"; ARCB_DEBUG_EPI=$m timeout 200 python tools/gpu_time_bwd.py --flags $f 2>&1 | tail -1 | cut -c1-420; done; done > gpurun_out/r2k_dw_debug.log 2>&1
cat gpurun_out/r2k_dw_debug.log
