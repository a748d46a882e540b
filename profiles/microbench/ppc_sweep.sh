#!/bin/bash
# Row-pass "polynomials per CTA" sweep (B200FHE_ROW_PPC) on bench.py's C2 batch; run on the GPU box from the repo root.
python -m pytest tests/test_ntt_gpu.py tests/test_protected_gpu.py tests/test_baseline_shapes_gpu.py tests/test_keyswitch_gpu.py -x -q 2>&1 | tail -3
for p in 1 2 4 8 0; do
  B200FHE_ROW_PPC=$p python bench.py --steps 100 --warmup 5 --no-sub --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ppc=$p (0 = automatic)', round(d['value']), {k:round(v*1e3,1) for k,v in d['roofline']['kernel_ms'].items()}, 'fwd', round(d['abft']['forward_ms'],4), 'prot', round(d['abft']['forward_protected_ms'],4), 'abft %', round(d['abft']['overhead_pct'],1))"
done
