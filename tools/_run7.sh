cd $GRAFT_REPO_ROOT
for m in 6 8; do ARCB_DEBUG_EPI=$m timeout 120 python tools/gpu_debug_probs2.py >> gpurun_out/r2g_debug.log 2>&1; done
cat gpurun_out/r2g_debug.log
