mkdir -p gpurun_out
timeout 300 python scripts/prof_train_graph.py MNIST_28 > gpurun_out/prof_train_mnist.log 2>&1; echo rc=$? >> gpurun_out/prof_train_mnist.log
timeout 300 python scripts/prof_train_graph.py FFHQ_64 > gpurun_out/prof_train_ffhq64.log 2>&1; echo rc=$? >> gpurun_out/prof_train_ffhq64.log
head -30 gpurun_out/prof_train_mnist.log; tail -3 gpurun_out/prof_train_mnist.log
