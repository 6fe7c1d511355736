# 2-GPU check of the peer-mailbox CG: the test script with its prints, then the Newton legs of bench.py with and without it
N=${N:-2}
[ -n "$SKIPTEST" ] || MFB_REQUIRE_PEER=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 tests/dist/run_dist_pcg.py 2>&1 | grep -v "^W\|^\*\*\*\|OMP_NUM" | tail -12
for o in ${OPTSETS:-"peer=1" "peer=1,cg_persist=0" "peer=0"}; do
  oo=""; for x in $(echo $o | tr "," " "); do oo="$oo --opt $x"; done
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu --no-tail $oo 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if not l.startswith('{'): continue
    d=json.loads(l)
    n3=d.get('newton_iteration') or {}; n4=d.get('newton_iteration_C4') or {}
    print('$o', 'N', d['n_gpus'], 'value', round(d['value']/1e9,3), 'e9 tets/s; C3 ms/newton', round(n3.get('ms_per_newton_iteration',0),2), 'us/cg', round(1e3*n3.get('ms_per_cg_iteration',0),1), '; C4 ms/newton', round(n4.get('ms_per_newton_iteration',0),2), 'us/cg', round(1e3*n4.get('ms_per_cg_iteration',0),1), 'parity', d['parity']['ok']); print(l) if '$SAVE' else None; print('   C3 kcycles/iter', n3.get('cg_persistent_kcycles_per_iteration')); print('   C4 kcycles/iter', n4.get('cg_persistent_kcycles_per_iteration'))
"
done
