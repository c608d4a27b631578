mkdir -p gpurun_out
timeout 600 python scripts/e2e_steps.py 1,3,3,1 1,2,3,3,2,1 1,2,4,4,2,1 1,3,4,3,1 1,2,5,5,2,1 2,5,5,2 1,2,4,6,4,2,1 1,4,4,1 1,3,3,2,1 > gpurun_out/e2e_sched.log 2>&1
cut -c1-110 gpurun_out/e2e_sched.log
