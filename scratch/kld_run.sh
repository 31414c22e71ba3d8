set -x
python -m pytest tests/test_gpu_kld.py -x -q > gpurun_out/r1k_tests.log 2>&1
tail -25 gpurun_out/r1k_tests.log
python bench.py --steps 20 --warmup 5 --no-push > gpurun_out/r1k_bench.json 2> gpurun_out/r1k_bench.err; tail -3 gpurun_out/r1k_bench.err
python -c "import json;d=json.load(open('gpurun_out/r1k_bench.json'));print(d.get('kld'))"
