mkdir -p gpurun_out
CMD="python scripts/prof_small.py fwd 2 MNIST_28"
$CMD > gpurun_out/simt_mnist_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:hop_fwd_simt -s 6 -c 1 -f -o gpurun_out/prof_simt_mnist $CMD > gpurun_out/simt_mnist_ncu.log 2>&1
tail -3 gpurun_out/simt_mnist_ncu.log
timeout 600 python scripts/bench_model_configs.py > gpurun_out/models_cpu.log 2>&1; echo rc=$? >> gpurun_out/models_cpu.log
cut -c1-300 gpurun_out/models_cpu.log | tail -6
