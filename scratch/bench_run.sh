set -x
python bench.py > gpurun_out/r1h_bench.json 2> gpurun_out/r1h_bench.err; echo rc=$?; tail -3 gpurun_out/r1h_bench.err
python -c "import json;d=json.load(open('gpurun_out/r1h_bench.json'));print(d['value'], d['ms_per_step']);print(json.dumps(d.get('eval_epilogue'),indent=1));print(json.dumps(d.get('topk'),indent=1))"
