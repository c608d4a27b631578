mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "backward or training" > gpurun_out/kv_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/kv_tests.log
tail -4 gpurun_out/kv_tests.log
for W in 3 2 1; do
HOP_BWD_KV_WAVES=$W timeout 600 python scripts/bench_model_configs.py > gpurun_out/kv_models_w$W.log 2>&1
python - <<PY
import json
for ln in open('gpurun_out/kv_models_w$W.log'):
    if ln.startswith('{'):
        j=json.loads(ln)
        if 'us_fwdbwd_ours_graph' in j: print('waves $W', j['config'], 'fwdbwd graph', round(j['us_fwdbwd_ours_graph'],1), 'fwd graph', round(j['us_fwd_ours_graph'],1), 'eager-call', round(j['us_fwdbwd_ours'],1))
PY
done
