set -x
cd $GRAFT_REPO_ROOT
timeout 120 python tools/gpu_debug_probs.py 1000 512 20000 > gpurun_out/r2d_debug.log 2>&1; cat gpurun_out/r2d_debug.log
timeout 120 python tools/gpu_debug_probs.py 1024 512 20480 >> gpurun_out/r2d_debug.log 2>&1; tail -12 gpurun_out/r2d_debug.log
T="timeout 120 python tools/gpu_time_bwd.py"
{
$T --tag keep_probs --flags 0x100
ARCB_DEBUG_EPI=1 $T --tag keep_dbg1_no_stores --flags 0x100
ARCB_DEBUG_EPI=2 $T --tag keep_dbg2_no_pass1 --flags 0x100
ARCB_DEBUG_EPI=3 $T --tag keep_dbg3_no_epilogue --flags 0x100
} > gpurun_out/r2d_timing.log 2>&1
cat gpurun_out/r2d_timing.log
