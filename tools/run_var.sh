# A/B runs of the update step on the GPU box: tail kernel with / without a CTA barrier per trip
E=/root/repo/mixed-fem_b200/exp
MFB_LIB=$E/libmfb_TS.so timeout 200 python -m pytest tests/test_gpu_parity.py -x -q -k "variants or indefinite or deterministic" 2>&1 | tail -2
for v in "" TS; do
  if [ -n "$v" ]; then export MFB_LIB=$E/libmfb_$v.so; fi
  python bench.py --steps 5 --warmup 3 --no-cpu --no-newton --no-newton-c4 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); t=d['roofline_tail']
print('$v', 'ms_step', round(d['ms_per_step'],3), 'tail ms_step', round(t['ms_per_step'],3), 'tail kernel_ms', round(t['kernel_ms'],3), 'slow', round(t['psd_slow_fraction'],3), d['parity']['ok'])
"
done
