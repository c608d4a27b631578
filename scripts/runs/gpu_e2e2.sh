for sk in 0 1; do for c in 1,2,4,1 1,3,3,1; do
  if [ $sk = 0 ]; then export HOP_TC_STREAMK=0; else unset HOP_TC_STREAMK; fi
  timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-chunks $c > gpurun_out/e2e2_${sk}_$c.log 2>&1
  python - gpurun_out/e2e2_${sk}_$c.log "streamk=$sk chunks $c" <<'PY'
import json,sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        j=json.loads(ln); print(sys.argv[2], "value %.4g e2e %.4g"%(j["value"], j["e2e"]["value"]))
PY
done; done
