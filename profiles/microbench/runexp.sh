for t in "$@"; do
  B200FHE_LIB=$PWD/fhe_reliability_gpu_b200/lib_exp/$t/libb200fhe.so timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/exp_$t.json 2> gpurun_out/exp_$t.err || tail -3 gpurun_out/exp_$t.err
  python - <<E
import json
try:
    d=json.load(open("gpurun_out/exp_$t.json")); k=d["roofline"]["kernel_ms"]
    print("$t", round(d["value"]), "ms/step %.4f"%d["ms_per_step"], {a:round(b,4) for a,b in k.items()}, "abft %.1f%%"%d["abft"]["overhead_pct"])
except Exception as e: print("$t FAILED", e)
E
done
