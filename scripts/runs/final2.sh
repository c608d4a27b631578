mkdir -p gpurun_out
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final2_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/final2_smoke.log
tail -4 gpurun_out/final2_smoke.log
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/final2_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final2_tests.log
tail -4 gpurun_out/final2_tests.log
timeout 240 python bench.py > gpurun_out/final2_bench.log 2>&1; echo "rc=$?" >> gpurun_out/final2_bench.log
tail -2 gpurun_out/final2_bench.log | cut -c1-900
