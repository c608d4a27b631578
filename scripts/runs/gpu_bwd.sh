mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q 2>&1 | tail -3
python scripts/prof_small.py fwdbwd 2 FFHQ_64 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/small_fwdbwd_launches_ffhq64_v2.csv python scripts/prof_small.py fwdbwd 2 FFHQ_64 > gpurun_out/ncu_s8.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/small_fwdbwd_launches_v2.csv python scripts/prof_small.py fwdbwd 2 > gpurun_out/ncu_s9.log 2>&1
python scripts/bench_model_configs.py > gpurun_out/model_configs.jsonl 2> gpurun_out/model_configs.err
python - <<'PY'
import json
for ln in open("gpurun_out/model_configs.jsonl"):
    j=json.loads(ln); print(j["config"], {k: round(v,1) for k,v in j.items() if k.startswith("us_")})
PY
timeout 300 python scripts/bench_ffhq64_ddp.py 20 2>&1 | grep -E '^\{|Error' | cut -c1-400
