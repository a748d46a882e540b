#!/bin/bash
# protected inverse transform: parity, then the overhead numbers bench.py reports (abft.inverse) and the fully protected key-switch
timeout 600 python -m pytest tests/test_protected_inverse_gpu.py tests/test_protected_gpu.py tests/test_keyswitch_gpu.py -x -q 2>&1 | tail -3
timeout 300 python bench.py --no-sub --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(json.dumps(d['abft']))"
