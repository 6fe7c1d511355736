#!/bin/bash
# bench every variant library under mixed-fem_b200/exp (see build_variant.sh) plus the default one
for lib in "" mixed-fem_b200/exp/libmfb_*.so; do
  [ -n "$lib" ] && [ ! -e "$lib" ] && continue
  if [ -n "$lib" ]; then export MFB_LIB=$PWD/$lib; else unset MFB_LIB; fi
  python bench.py --no-cpu --no-newton --steps 30 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('${lib:-default}', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['e2e']['ms_per_step'])"
done
