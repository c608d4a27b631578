mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "not tensor_core and not full_size" > gpurun_out/tb_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tb_tests.log
tail -6 gpurun_out/tb_tests.log
timeout 600 python scripts/bench_model_configs.py > gpurun_out/tb_models.log 2>&1
python - <<PY
import json
for ln in open('gpurun_out/tb_models.log'):
    if ln.startswith('{'):
        j=json.loads(ln)
        if 'us_fwdbwd_ours_graph' in j: print(j['config'], 'fwd graph', round(j['us_fwd_ours_graph'],1), 'fwdbwd graph', round(j['us_fwdbwd_ours_graph'],1), 'fwdbwd eager-call', round(j['us_fwdbwd_ours'],1), 'torch graph', round(j['us_fwdbwd_eager_upstream_shape_graph'],1))
PY
timeout 300 python scripts/prof_train_graph.py MNIST_28 > gpurun_out/tb_prof_mnist.log 2>&1
head -14 gpurun_out/tb_prof_mnist.log | cut -c1-120
