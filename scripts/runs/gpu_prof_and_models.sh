bash scripts/gpu_profile_tc.sh
timeout 600 python scripts/bench_model_configs.py > gpurun_out/model_configs.log 2>&1; echo "models rc=$?"; grep '^{' gpurun_out/model_configs.log | cut -c1-700; tail -3 gpurun_out/model_configs.log | grep -v '^{'
