python -m pytest tests -q -m gpu --timeout 180 --tb=short 2>&1 | tail -12 > gpurun_out/r2h_tests.log; cat gpurun_out/r2h_tests.log
python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err; echo rc=$?; tail -3 gpurun_out/r2h_bench.err
python -c "
import json
d=json.load(open('gpurun_out/r2h_bench.json'))
print(json.dumps(d.get('stage2'))[:700])"
