timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "forward_host" 2>&1 | tail -15
for i in 1 2 3; do
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/graph_$i.log 2>&1
python - gpurun_out/graph_$i.log "run $i" <<'PY'
import json,sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        j=json.loads(ln); print(sys.argv[2], "value %.4g e2e %.4g"%(j["value"], j["e2e"]["value"]))
PY
grep -E "Error|error" gpurun_out/graph_$i.log | head -3
done
python scripts/e2e_steps.py 2>&1 | tail -4 | cut -c1-330
