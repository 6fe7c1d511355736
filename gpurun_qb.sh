python -m pytest tests -x -q -m gpu 2>&1 | grep -E "^E|passed|failed" | head -12 | cut -c1-300
for v in 0 1 2; do
python bench.py --steps 10 --warmup 3 --grid 55 55 55 --no-cpu --variant $v 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('variant $v 1M: tets/s %.4g ms %.3f kernel_ms %.3f e2e %.4g launches %d'%(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['value'], d['gpu_launches']))"
done
python bench.py --steps 10 --warmup 3 --no-cpu --variant 2 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('variant 2 10M: tets/s %.4g ms %.3f kernel_ms %.3f e2e %.4g'%(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['value']))"
python bench.py --steps 10 --warmup 3 --no-cpu --variant 1 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('variant 1 10M: tets/s %.4g ms %.3f kernel_ms %.3f e2e %.4g'%(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['value']))"
