# first light of the 64-rows-per-CTA pair kernel: D = 512 / 1024 parity cases under a short timeout
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_parity_gpu.py -x -q -k "d512 or d1024" --timeout 60 > gpurun_out/halfm_tests.log 2>&1; echo "rc=$?" >> gpurun_out/halfm_tests.log
tail -25 gpurun_out/halfm_tests.log | cut -c1-250
