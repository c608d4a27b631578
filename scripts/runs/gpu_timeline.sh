mkdir -p gpurun_out
HOP_SM100_LIB=$PWD/hop-vae_b200/libhop_sm100_trace.so timeout 300 python scripts/tc_timeline.py ${1:-sweep_p64k_d256} > gpurun_out/timeline.log 2>&1; echo rc=$?
tail -5 gpurun_out/timeline.log
