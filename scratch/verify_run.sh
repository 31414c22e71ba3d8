set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r1g_tests.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1g_smoke.log 2>&1; echo smoke=$? >> gpurun_out/r1g_smoke.log
python bench.py > gpurun_out/r1g_bench.json 2> gpurun_out/r1g_bench.err
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r1g_bench_ref.json 2> gpurun_out/r1g_bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1g_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu --push-maps 64 > gpurun_out/r1g_ncu.log 2>&1
tail -3 gpurun_out/r1g_tests.log; tail -2 gpurun_out/r1g_smoke.log; cat gpurun_out/r1g_bench.json | head -c 600
