mkdir -p gpurun_out
CMD="python scripts/prof_small.py fwdbwd 2 MNIST_28"
$CMD > gpurun_out/warm_plain.log 2>&1 && ncu --set full --clock-control none --cache-control none --import-source on -k regex:"hop_fwd_simt|bwd_core_q|bwd_core_kv|wgrad_kernel" -s 20 -c 4 -f -o gpurun_out/prof_warm_mnist $CMD > gpurun_out/warm_ncu.log 2>&1
tail -3 gpurun_out/warm_ncu.log
