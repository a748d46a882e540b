#!/bin/bash
# ncu evidence for round 2 (GPU box, repo root): launch list of the default bench command, and one --set full capture of a warm
# iteration of the four transforms (unprotected + protected) with the summary columns DESIGN.md quotes.
set -x
python profiles/microbench/prot_driver.py 6 > gpurun_out/$1_prot_driver.txt 2>&1 || exit 1
python bench.py --steps 2 --warmup 1 --no-sub --no-cpu-baseline > gpurun_out/$1_bench_short.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/$1_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-sub --no-cpu-baseline > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_fwd|k_inv|k_abft' --launch-skip 10 --launch-count 10 -f -o gpurun_out/$1_ntt \
    python profiles/microbench/prot_driver.py 2 > gpurun_out/$1_ncu.log 2>&1
python profiles/ncu_summary.py gpurun_out/$1_ntt.ncu-rep > gpurun_out/$1_ncu_full_summary.txt 2>&1
ls -la gpurun_out/
