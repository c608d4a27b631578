mkdir -p gpurun_out
i=0
for K in "A=1" "HOP_SIMT_SPLITS=8" "HOP_SIMT_RM=2" "HOP_SIMT_RM=2 HOP_SIMT_SPLITS=8" "HOP_SIMT_RM=2 HOP_SIMT_SPLITS=4" "HOP_SIMT_RM=1 HOP_SIMT_SPLITS=4"; do
i=$((i+1))
env $K timeout 600 python scripts/bench_model_configs.py > gpurun_out/knob_$i.log 2>&1
python - <<PY
import json
for ln in open('gpurun_out/knob_$i.log'):
    if ln.startswith('{'):
        j=json.loads(ln)
        if 'us_fwdbwd_ours_graph' in j: print('$K', j['config'], 'fwd graph', round(j['us_fwd_ours_graph'],1), 'fwdbwd graph', round(j['us_fwdbwd_ours_graph'],1))
PY
done
