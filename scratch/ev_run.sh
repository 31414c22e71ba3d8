set -x
python -m pytest tests/test_gpu_evalmap.py -q > gpurun_out/r1e_tests.log 2>&1; tail -5 gpurun_out/r1e_tests.log
python bench.py --steps 20 --warmup 5 --no-push --no-cpu > gpurun_out/r1e_bench.json 2> gpurun_out/r1e_bench.err; tail -2 gpurun_out/r1e_bench.err
python -c "import json;d=json.load(open('gpurun_out/r1e_bench.json'));print(json.dumps(d.get('eval_epilogue'),indent=1))"
