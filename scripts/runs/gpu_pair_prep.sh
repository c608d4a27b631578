mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pp_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pp_tests.log
tail -4 gpurun_out/pp_tests.log
timeout 600 python scripts/bench_model_configs.py > gpurun_out/pp_models.log 2>&1
python - <<PY
import json
for ln in open('gpurun_out/pp_models.log'):
    if ln.startswith('{'):
        j=json.loads(ln)
        if 'us_fwdbwd_ours_graph' in j: print(j['config'], 'fwd graph', round(j['us_fwd_ours_graph'],1), 'fwdbwd graph', round(j['us_fwdbwd_ours_graph'],1), 'fwd eager-call', round(j['us_fwd_ours'],1), 'fwdbwd eager-call', round(j['us_fwdbwd_ours'],1))
        else: print(j['config'], 'i2e graph', round(j['us_index_to_embedding_ours_graph'],1), 'cpu', round(j['us_lookup_plus_decoder_cpu_upstream_shape'],1))
PY
