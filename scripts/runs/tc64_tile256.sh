mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_parity_gpu.py -x -q -k "d512 or d1024 or 128_row or bf16 or single_gpu_tensor" --timeout 60 > gpurun_out/tile256_tests.log 2>&1; echo "rc=$?" >> gpurun_out/tile256_tests.log
tail -6 gpurun_out/tile256_tests.log | cut -c1-250
P='import sys,json; l=json.loads(sys.stdin.read()); print(sys.argv[1], l["config"]["workload"], "ms", round(l["ms_per_step"],3), "TF", round(l["roofline"]["achieved"],1), "frac", round(l["roofline"]["frac"],3), l["clocks"]["sm_mhz"], l["clocks"]["reasons"])'
for t in 256 128 256 128; do
HOP_TC64_TILE=$t timeout 200 python bench.py --workload sweep_p64k_d512 --no-cpu-baseline 2>&1 | grep '^{' | python -c "$P" tile$t
done
