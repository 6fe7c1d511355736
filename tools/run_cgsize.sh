# persistent CG kernel on one GPU at the per-rank sizes of C3 over 8 / 4 GPUs: cooperative grid size sweep
for G in 0 148 296 444; do
for g in "7 55 55" "14 55 55"; do
  MFB_PCG_G=$G python bench.py --steps 3 --warmup 3 --no-cpu --no-tail --no-newton-c4 --newton-grid $g --grid 30 30 30 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if not l.startswith('{'): continue
    d=json.loads(l); n3=d['newton_iteration']
    print('G=$G', '$g', 'us/cg', round(1e3*n3['ms_per_cg_iteration'],1), n3.get('cg_persistent_kcycles_per_iteration'))
"
done
done
