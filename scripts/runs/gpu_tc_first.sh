mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "tensor_core" -s > gpurun_out/tc_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tc_tests.log
tail -25 gpurun_out/tc_tests.log
timeout 300 python bench.py --workload sweep_p64k_d256 --path f16 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tc_d256.log 2>&1; echo "rc=$?" >> gpurun_out/bench_tc_d256.log
tail -5 gpurun_out/bench_tc_d256.log
