mkdir -p gpurun_out
for c in 1,3,3,1 1,6,1 1,2,4,1 1,5,2 2,5,1 1,4,2,1 1,10,4,1; do
  timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-chunks $c > gpurun_out/e2e_$c.log 2>&1
  python - gpurun_out/e2e_$c.log "chunks $c" <<'PY'
import json,sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        j=json.loads(ln); print(sys.argv[2], "value %.4g e2e %.4g"%(j["value"], j["e2e"]["value"]))
PY
done
