set -x
cd $GRAFT_REPO_ROOT
timeout 120 python tools/gpu_debug_probs.py 1000 512 20000 > gpurun_out/r2f_debug.log 2>&1; cat gpurun_out/r2f_debug.log
ARCB_DEBUG_EPI=4 timeout 120 python tools/gpu_debug_probs.py 1000 512 20000 > gpurun_out/r2f_debug4.log 2>&1; tail -4 gpurun_out/r2f_debug4.log
T="timeout 120 python tools/gpu_time_bwd.py"
{
$T --tag default
$T --tag keep_probs --flags 0x100
$T --tag keep_probs_dx1cta --flags 0x120
ARCB_DEBUG_EPI=1 $T --tag keep_dbg1_no_stores --flags 0x100
ARCB_DEBUG_EPI=3 $T --tag keep_dbg3_no_epilogue --flags 0x100
$T --tag c125k --classes 125000
$T --tag c125k_keep --classes 125000 --flags 0x100
} > gpurun_out/r2f_timing.log 2>&1
cat gpurun_out/r2f_timing.log
BARGS="--steps 2 --warmup 3 --no-cpu-baseline --no-parity --sustained-seconds 0 --kernel-flags 0x80"
timeout 200 python bench.py $BARGS > gpurun_out/r2f_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"bwd_|fwd_|l2norm|loss_|merge_|target_prep|gemm_" -c 120 --csv --log-file gpurun_out/r2f_launches.csv python bench.py $BARGS > gpurun_out/r2f_ncu1.log 2>&1
echo "ncu1 rc=$?"
