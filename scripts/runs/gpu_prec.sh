mkdir -p gpurun_out
echo skip-tests
PTS="--points=262144,256,1;262144,256,3;1048576,256,0;1048576,256,3;65536,256,3;262144,512,3"
timeout 900 python scripts/sweep_config5.py "$PTS" 2>gpurun_out/prec.err | python -c "
import sys,json
for ln in sys.stdin:
    j=json.loads(ln); print('P=%7d D=%4d T=%d  %.3f ms frac %.3f  rel-L2 %.2e  argmax %d/%d'%(j['patterns'],j['dim'],j['update_steps'],j['ms'],j['frac_of_measured_peak'],j['rel_l2'],j['argmax_mismatch'],j['argmax_decided_rows']))"
tail -3 gpurun_out/prec.err
