set -x
cd $GRAFT_REPO_ROOT
timeout 400 python -m pytest tests/test_gpu_parity.py -x -q -k "stored_probabilities or resident or inside_the_gemm" > gpurun_out/r2c_pytest_new.log 2>&1; echo "new tests rc=$?"; tail -15 gpurun_out/r2c_pytest_new.log
T="timeout 120 python tools/gpu_time_bwd.py"
{
$T --tag default
$T --tag streaming --flags 0x8
$T --tag keep_probs --flags 0x100
$T --tag keep_probs_dx1cta --flags 0x120
$T --tag keep_probs_norm_separate --flags 0x101
$T --tag c125k --classes 125000
$T --tag c125k_keep --classes 125000 --flags 0x100
$T --tag cfg3 --classes 100000 --batch 512
} > gpurun_out/r2c_timing.log 2>&1
cat gpurun_out/r2c_timing.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2c_pytest.log
