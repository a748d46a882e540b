#!/bin/bash
# multi-GPU evidence (box with $1 GPUs, repo root): sharded ModUp through the C ABI + the driver's bench command under torchrun
n=$1; tag=$2
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29551"
python -m pytest tests/test_modup_sharded_gpu.py -x -q 2>&1 | tail -2
python bench_modup.py > gpurun_out/${tag}_modup_1gpu.json 2>/dev/null; cat gpurun_out/${tag}_modup_1gpu.json | cut -c1-300
timeout 300 $T bench_modup.py > gpurun_out/${tag}_modup_${n}gpu.json 2> gpurun_out/${tag}_modup_${n}gpu.err; echo "rc=$?"; cat gpurun_out/${tag}_modup_${n}gpu.json | cut -c1-400
timeout 900 $T bench.py --gpus $n --steps 50 --warmup 5 > gpurun_out/${tag}_bench_${n}gpu.json 2> gpurun_out/${tag}_bench_${n}gpu.err; echo "rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/${tag}_bench_${n}gpu.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['value'], d['e2e']['value'], d.get('modup_sharded'))"
