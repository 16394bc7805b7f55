cd $GRAFT_REPO_ROOT
for m in 0 256 512 0 256 512; do echo "debug $m"; ARCB_DEBUG_EPI=$m timeout 200 python tools/gpu_time_bwd.py 2>&1 | tail -1 | cut -c150-420; done
