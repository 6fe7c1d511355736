#!/bin/bash
# one GPU call: timings of the §8(f) paths, A/B of the variant libraries, parity of the update kernel
python tools/bench_paths.py 2>&1 | tail -1
bash tools/run_variants.sh; bash tools/run_variants.sh
for lib in mixed-fem_b200/exp/libmfb_*.so; do
  [ -e "$lib" ] || continue
  MFB_LIB=$PWD/$lib python -m pytest tests/test_gpu_parity.py tests/test_gpu_newton.py -m gpu -x -q 2>&1 | tail -1
done
