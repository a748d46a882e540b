#!/bin/bash
# k_pmult with the short-sum fold: parity of everything that multiplies-and-accumulates, then the key-switch pieces and C3
timeout 900 python -m pytest tests/test_pmult_gpu.py tests/test_keyswitch_gpu.py tests/test_bgv_shim_gpu.py tests/test_bgv_dotprod_gpu.py tests/test_evaluator_gpu.py tests/test_baseline_shapes_gpu.py tests/test_harness_gpu.py -x -q 2>&1 | tail -3
timeout 400 python bench_keyswitch.py --steps 20 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print({k:d[k] for k in ('ms_eager','ms_graph','pieces_ms')})"
timeout 300 python bench_ops.py 2>/dev/null | grep -o '"metric": "ct\*pt[^}]*' | cut -c1-300
