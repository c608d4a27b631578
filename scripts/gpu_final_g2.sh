# 2-GPU re-check after the stream-K / graph changes: sharded + DDP tests, bench in both modes and the reference arm
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q > gpurun_out/final_g2_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final_g2_tests.log
tail -4 gpurun_out/final_g2_tests.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 300 $TR bench.py --gpus 2 > gpurun_out/final_g2_batch.log 2>&1; echo "rc=$?" >> gpurun_out/final_g2_batch.log
timeout 300 $TR bench.py --gpus 2 --mode sharded > gpurun_out/final_g2_sharded.log 2>&1; echo "rc=$?" >> gpurun_out/final_g2_sharded.log
timeout 300 $TR bench.py --gpus 2 --impl reference --steps 3 > gpurun_out/final_g2_ref.log 2>&1; echo "rc=$?" >> gpurun_out/final_g2_ref.log
for f in batch sharded ref; do grep -E '^\{|rc=' gpurun_out/final_g2_$f.log | cut -c1-700; done
