# A/B runs of the update step on the GPU box: each line = one configuration (env / library options) of bench.py's value leg
run() {
  label=$1; shift
  env "$@" python bench.py --steps 20 --warmup 3 --no-cpu --no-newton --no-newton-c4 --no-tail $OPTS 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$label', 'ms_step', round(d['ms_per_step'],3), 'kernel_ms', round(d['roofline']['kernel_ms'],3), 'frac', round(d['roofline']['frac'],4), 'e2e_ms', round(d['e2e']['ms_per_step'],3), 'parity', d['parity']['ok'])
"
}
OPTS="" run zero_overlap X=1
OPTS="--opt zero_overlap=0" run inline_memset X=1
OPTS="" run zero_overlap X=1
