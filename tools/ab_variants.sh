#!/bin/bash
# A/B of library variants built into build_variants/ (dev tool): parity subset + fusion timing.
WL=${1:-cfg3}
for lib in build_variants/lib_*.so; do
  export VACANCY_B200_LIB=$PWD/$lib
  ok=$(timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "bunny or options_matrix or more_views" 2>&1 | tail -1)
  out=$(timeout 600 python bench.py --workload $WL --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1)
  echo "$lib | $ok | $(python -c "
import json,sys
d=json.loads(sys.argv[1]); print('step %.1f ms fusion %.1f ms sdf %.2f mc %.2f'%(d['ms_per_step'], d['stages_ms_per_step']['fusion'], d['stages_ms_per_step']['sdf2d'], d['stages_ms_per_step']['marching_cubes']))" "$out")"
done
