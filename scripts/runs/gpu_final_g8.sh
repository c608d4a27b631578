# final 8-GPU check of the driver's scaling run (batch-parallel default bench) + one pattern-sharded line
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519"
timeout 300 $TR bench.py --gpus 8 --no-cpu-baseline > gpurun_out/final_g8_batch.log 2>&1; echo "rc=$?" >> gpurun_out/final_g8_batch.log
timeout 300 $TR bench.py --gpus 8 --mode sharded --workload sweep_p256k_d256 --no-cpu-baseline > gpurun_out/final_g8_sharded.log 2>&1; echo "rc=$?" >> gpurun_out/final_g8_sharded.log
for f in batch sharded; do grep -E '^\{|rc=' gpurun_out/final_g8_$f.log | cut -c1-900; done
