# the driver's scaling run: bench.py at N = 8 and 4 (batch-parallel, default workload), as torchrun launches it
mkdir -p gpurun_out
for G in 8 4; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $G --steps 20 --warmup 3 > gpurun_out/scale_g$G.log 2>&1; echo "rc=$?" >> gpurun_out/scale_g$G.log
  grep -E '^\{|rc=' gpurun_out/scale_g$G.log | python -c "
import sys, json
for ln in sys.stdin:
    if ln.startswith('{'):
        j=json.loads(ln); print('g$G: value %.4g ms/step %.3f e2e %.4g numa %s clocks %s'%(j['value'], j['ms_per_step'], j['e2e']['value'], j['config'].get('host_numa_cpus'), j['clocks']))
    else: print('g$G', ln.strip())
"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 5 --warmup 3 --impl reference 2>&1 | grep -E '^\{' | cut -c1-400
