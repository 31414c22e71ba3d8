set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r1f_tests.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1f_smoke.log 2>&1
python bench.py > gpurun_out/r1f_bench.json 2> gpurun_out/r1f_bench.err
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r1f_bench_ref.json 2> gpurun_out/r1f_bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1f_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu --push-maps 64 > gpurun_out/r1f_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:head_bwd_tc_kernel -s 3 -c 1 -f -o gpurun_out/r1f_bwd_full python scratch/quick.py 4 bwd > gpurun_out/r1f_ncubwd.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:head_fwd_tc_kernel -s 3 -c 1 -f -o gpurun_out/r1f_fwd_full python scratch/quick.py 4 fwd > gpurun_out/r1f_ncufwd.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:push_scan_tc_kernel -s 3 -c 1 -f -o gpurun_out/r1f_push_full python scratch/quick.py 4 push > gpurun_out/r1f_ncupush.log 2>&1
tail -2 gpurun_out/r1f_tests.log; tail -2 gpurun_out/r1f_smoke.log; ls -la gpurun_out/*.ncu-rep
timeout 60 scratch/mma_rate > gpurun_out/r1f_probe_mma_rate.txt 2>&1
timeout 60 scratch/xu_probe > gpurun_out/r1f_probe_xu.txt 2>&1
