mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "not tensor_core and not full_size" > gpurun_out/pro_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pro_tests.log
tail -4 gpurun_out/pro_tests.log
i=0
for K in "A=1" "HOP_TINY_LPR=4" "HOP_TINY_LPR=8" "HOP_TINY_LPR=16" "HOP_TINY_LPR=32"; do
i=$((i+1))
env $K timeout 600 python scripts/bench_model_configs.py > gpurun_out/pro_$i.log 2>&1
python - <<PY
import json
for ln in open('gpurun_out/pro_$i.log'):
    if ln.startswith('{'):
        j=json.loads(ln)
        if 'us_fwdbwd_ours_graph' in j: print('$K', j['config'], 'fwd graph', round(j['us_fwd_ours_graph'],1), 'fwdbwd graph', round(j['us_fwdbwd_ours_graph'],1))
        else: print('$K', j['config'], 'i2e graph', round(j['us_index_to_embedding_ours_graph'],1), 'eagercall', round(j['us_index_to_embedding_ours'],1))
PY
done
