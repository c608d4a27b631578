mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q 2>&1 | tail -3
python scripts/bench_model_configs.py > gpurun_out/model_configs.jsonl 2> gpurun_out/model_configs.err; tail -2 gpurun_out/model_configs.err
python - <<'PY'
import json
for ln in open("gpurun_out/model_configs.jsonl"):
    j=json.loads(ln); print(j["config"], {k: (round(v,1) if v>1 else v) for k,v in j.items() if k.startswith("us_") or k.startswith("parity")})
PY
