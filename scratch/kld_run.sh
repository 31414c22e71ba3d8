set -x
python -m pytest tests/test_gpu_kld.py tests/test_gpu_evalmap.py -q > gpurun_out/r1k_tests.log 2>&1
tail -5 gpurun_out/r1k_tests.log
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_bytes.sum --clock-control none --csv --log-file gpurun_out/r1k_kld_launches.csv python scratch/kld_quick.py > gpurun_out/r1k_ncu.log 2>&1
tail -3 gpurun_out/r1k_ncu.log
python bench.py --steps 50 --warmup 5 --no-push > gpurun_out/r1k_bench.json 2> gpurun_out/r1k_bench.err; tail -3 gpurun_out/r1k_bench.err
python -c "import json;d=json.load(open('gpurun_out/r1k_bench.json'));print(json.dumps(d.get('kld'),indent=1))"
