mkdir -p gpurun_out
timeout 300 python scripts/prof_small.py fwdbwd 200 MNIST_28 > gpurun_out/hostprof_fwdbwd.log 2>&1
head -60 gpurun_out/hostprof_fwdbwd.log | cut -c1-150
