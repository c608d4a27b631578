mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/re_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/re_tests.log
tail -4 gpurun_out/re_tests.log
timeout 300 python scripts/bench_ffhq64_ddp.py > gpurun_out/re_ffhq64.log 2>&1; echo "rc=$?" >> gpurun_out/re_ffhq64.log
grep -E '^\{|rc=' gpurun_out/re_ffhq64.log | cut -c1-500
