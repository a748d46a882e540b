#!/bin/bash
# checksum-carrying tcgen05 contraction: parity + fault detection, then the overhead numbers (per digit and per key-switch)
timeout 600 python -m pytest tests/test_bconv_protected_gpu.py tests/test_keyswitch_gpu.py tests/test_bconv_gpu.py tests/test_baseline_shapes_gpu.py -x -q 2>&1 | tail -6
timeout 300 python bench_ops.py > gpurun_out/r3a_ops.jsonl 2>gpurun_out/r3a_ops.err
grep -i "bconv\|contract\|checksum" gpurun_out/r3a_ops.jsonl | cut -c1-600
timeout 400 python bench_keyswitch.py --steps 20 > gpurun_out/r3a_keyswitch.json 2>gpurun_out/r3a_ks.err
cut -c1-1500 gpurun_out/r3a_keyswitch.json
