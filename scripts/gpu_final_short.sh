# Short re-verification for a small GPU budget: smoke, one bench line, then the GPU tests with what is left.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/short_gpu.txt 2>&1
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/short_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/short_smoke.log
tail -3 gpurun_out/short_smoke.log
timeout 240 python bench.py > gpurun_out/short_bench.log 2>&1; echo "rc=$?" >> gpurun_out/short_bench.log
tail -2 gpurun_out/short_bench.log
timeout ${1:-600} python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/short_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/short_tests.log
tail -25 gpurun_out/short_tests.log
