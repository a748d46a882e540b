# usage (on the GPU box): bash profiles/microbench/runexp.sh <tag> [<tag> ...]   -- one bench line per experiment build
# (fhe_reliability_gpu_b200/fastbuild.sh <tag> <flags> builds lib_exp/<tag>/libb200fhe.so here; env B200FHE_* passes through)
for t in "$@"; do
  B200_BENCH_NOCHECK=1 B200FHE_LIB=$PWD/fhe_reliability_gpu_b200/lib_exp/$t/libb200fhe.so timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-sub > gpurun_out/exp_$t.json 2> gpurun_out/exp_$t.err || tail -3 gpurun_out/exp_$t.err
  python - <<E
import json
try:
    d=json.load(open("gpurun_out/exp_$t.json")); k=d["roofline"]["kernel_ms"]
    print("$t", round(d["value"]), "ms/step %.4f"%d["ms_per_step"], {a:round(b*1000,1) for a,b in k.items()}, "abft %.1f%%"%d["abft"]["overhead_pct"])
except Exception as e: print("$t FAILED", e)
E
done
