mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python scripts/prof_small.py fwd 200 > gpurun_out/small_fwd_host.log 2>&1
python scripts/prof_small.py fwdbwd 100 > gpurun_out/small_fwdbwd_host.log 2>&1
python scripts/bench_model_configs.py > gpurun_out/model_configs.jsonl 2> gpurun_out/model_configs.err
cut -c1-640 gpurun_out/model_configs.jsonl
