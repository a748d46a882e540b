#!/bin/bash
# protected-NTT parity + ABFT overhead on bench.py's C2 batch (GPU box, repo root)
python -m pytest tests/test_protected_gpu.py tests/test_protected_inverse_gpu.py tests/test_baseline_shapes_gpu.py tests/test_prototypes_gpu.py tests/test_bgv_shim_gpu.py -x -q 2>&1 | tail -15
for i in 1 2; do python bench.py --steps 100 --warmup 5 --no-sub --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(round(d['value']), {k:round(v*1e3,1) for k,v in d['roofline']['kernel_ms'].items()}, 'fwd', round(d['abft']['forward_ms'],4), 'prot', round(d['abft']['forward_protected_ms'],4), 'abft %', round(d['abft']['overhead_pct'],1))"; done
