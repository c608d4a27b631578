mkdir -p gpurun_out
show() { python - "$1" "$2" <<'PY'
import json,sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        j=json.loads(ln); print(sys.argv[2], "ms/step %.3f kernel_ms %.3f TF %.1f frac %.3f e2e %.4g clocks %s"%(j["ms_per_step"], j["roofline"]["kernel_ms"], j["roofline"]["achieved"], j["roofline"]["frac"], j["e2e"]["value"], j["clocks"]["sm_mhz"]))
PY
}
for tag in base p2 p3; do
  if [ "$tag" = base ]; then unset HOP_SM100_LIB; else export HOP_SM100_LIB=$PWD/hop-vae_b200/libhop_sm100_$tag.so; fi
  for wl in sweep_p64k_d64 sweep_p64k_d128 sweep_p64k_d256; do
    timeout 300 python bench.py --workload $wl --path f16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/tune_${tag}_$wl.log 2>&1
    show gpurun_out/tune_${tag}_$wl.log "$tag $wl"
  done
done
unset HOP_SM100_LIB
for c in 2 4 8 16; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-chunks $c > gpurun_out/tune_chunks_$c.log 2>&1
  show gpurun_out/tune_chunks_$c.log "e2e-chunks $c"
done
# ncu: D = 512 fused forward (launch list + full capture), current fp32 forward at the FFHQ-64 shape
CMD="python bench.py --workload sweep_p64k_d512 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_tc_d512.csv $CMD > gpurun_out/ncu_d512_1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:hop_fwd_tc -s 4 -c 1 -o gpurun_out/prof_tc_d512 $CMD > gpurun_out/ncu_d512_2.log 2>&1
python scripts/prof_small.py fwd 2 FFHQ_64 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/small_fwd_launches_ffhq64_v2.csv python scripts/prof_small.py fwd 2 FFHQ_64 > gpurun_out/ncu_s6.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:hop_fwd_simt -s 9 -c 3 -o gpurun_out/prof_simt_fwd_v2 python scripts/prof_small.py fwd 2 FFHQ_64 > gpurun_out/ncu_s7.log 2>&1
ls -la gpurun_out/*.ncu-rep
