#!/bin/bash
# tcgen05 contraction: parity (every test that goes through a base conversion) then A/B timing against the mma.sync kernel
timeout 600 python -m pytest tests/test_bconv_gpu.py tests/test_keyswitch_gpu.py tests/test_baseline_shapes_gpu.py tests/test_modup_sharded_gpu.py tests/test_zz_no_alloc.py tests/test_evaluator_gpu.py -x -q 2>&1 | tail -8
for v in tc5 mma; do
  if [ $v = mma ]; then export B200FHE_BCONV_NO_TC5=1; fi
  echo "== $v"
  timeout 300 python bench_ops.py 2>/dev/null | grep -i "bconv\|contract" | cut -c1-400
  timeout 300 python bench_keyswitch.py --steps 20 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print({k:d[k] for k in d if k in ('value','ms_per_step','ms','parts_ms','cuda_graph')})"
done
