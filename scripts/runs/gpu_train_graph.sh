mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "under_cuda_graph or training_step" > gpurun_out/tg_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tg_tests.log
tail -15 gpurun_out/tg_tests.log
timeout 600 python scripts/bench_model_configs.py > gpurun_out/tg_models.log 2>&1; echo "rc=$?" >> gpurun_out/tg_models.log
cut -c1-1200 gpurun_out/tg_models.log
