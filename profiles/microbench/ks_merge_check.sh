#!/bin/bash
# ModUp's forward transforms as one scatter launch: parity, then A/B against the per-range launches
timeout 900 python -m pytest tests/test_keyswitch_gpu.py tests/test_ntt_gpu.py tests/test_baseline_shapes_gpu.py tests/test_bgv_shim_gpu.py tests/test_bgv_dotprod_gpu.py tests/test_modup_sharded_gpu.py tests/test_protected_gpu.py -x -q 2>&1 | tail -3
for v in merged ranges; do
  if [ $v = ranges ]; then export B200FHE_KS_NO_MERGE=1; fi
  echo "== $v"
  timeout 400 python bench_keyswitch.py --steps 20 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print({k:d[k] for k in ('ms_eager','ms_graph','pieces_ms','gpu_launches_per_step')})"
done
unset B200FHE_KS_NO_MERGE
timeout 300 python bench.py --no-sub --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['roofline']['kernel_ms'])"
