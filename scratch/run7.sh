python bench.py --steps 20 --warmup 5 > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo rc=$?; tail -5 gpurun_out/r2g_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2g_bench.json'))
for k in ['value','ms_per_step','roofline','e2e','h2d_probe','push','push_fp16_nhwc','push_coco','coco','nhwc16','stage2','dense_mode','cpu_baseline','c1','fp32_reroutes']:
    print(k, json.dumps(d.get(k))[:900])
PY
