for mb in 2 4 8 16 36; do
B200FHE_HOST_CHUNK_MB=$mb B200FHE_LIB=$PWD/fhe_reliability_gpu_b200/lib_exp/hostpipe/libb200fhe.so python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/t.json 2>gpurun_out/t.err || tail -3 gpurun_out/t.err
python -c "
import json; d=json.load(open('gpurun_out/t.json')); print($mb, 'e2e', round(d['e2e']['value']), 'dev', round(d['value']))"
done
