#!/bin/bash
# k_ct_tensor: pairs of coefficients per thread (loads issued up front) -- B200FHE_CT_E = 1, 2, 4
timeout 300 python -m pytest tests/test_evaluator_gpu.py -x -q 2>&1 | tail -2
for e in 1 2 4; do
  echo "== E=$e"
  B200FHE_CT_E=$e timeout 300 python -m pytest tests/test_evaluator_gpu.py -x -q 2>&1 | tail -1
  B200FHE_CT_E=$e timeout 300 python bench_ops.py 2>/dev/null | grep -o '"ct_tensor": {[^}]*}'
done
