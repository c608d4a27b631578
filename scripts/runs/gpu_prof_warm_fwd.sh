mkdir -p gpurun_out
CMD="python scripts/prof_small.py fwd 2 MNIST_28"
$CMD > gpurun_out/warmf_plain.log 2>&1 && ncu --set full --clock-control none --cache-control none --import-source on -k regex:"hop_fwd_simt" -s 6 -c 1 -f -o gpurun_out/prof_warm_fwd $CMD > gpurun_out/warmf_ncu.log 2>&1
tail -2 gpurun_out/warmf_ncu.log
