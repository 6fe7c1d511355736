# single-GPU CG time per iteration at the per-rank sizes of C3 split over 1, 2, 4, 8 GPUs (no exchange at all), persistent
# cooperative kernel against three launches per iteration
for o in "cg_persist=1" "cg_persist=0"; do
for g in "55 55 55" "28 55 55" "14 55 55" "7 55 55"; do
  python bench.py --steps 3 --warmup 3 --no-cpu --no-tail --no-newton-c4 --newton-grid $g --grid 30 30 30 --opt $o 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if not l.startswith('{'): continue
    d=json.loads(l); n3=d['newton_iteration']
    print('$o', '$g', 'ms/newton', round(n3['ms_per_newton_iteration'],2), 'us/cg', round(1e3*n3['ms_per_cg_iteration'],1), 'cg iters', n3['cg_iterations'][:3], 'frac', round(n3['roofline_cg']['frac'],3))
"
done
done
