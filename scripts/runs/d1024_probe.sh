# first light of the D = 1024 tensor-core form: targeted parity cases under a short timeout, then timing
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_parity_gpu.py -x -q -k "d1024" --timeout 60 > gpurun_out/d1024_tests.log 2>&1; echo "rc=$?" >> gpurun_out/d1024_tests.log
tail -15 gpurun_out/d1024_tests.log
