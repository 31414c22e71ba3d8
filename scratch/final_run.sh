set -x
python -m pytest tests -m gpu -q > gpurun_out/r1f_tests.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1f_smoke.log 2>&1; echo smoke=$? >> gpurun_out/r1f_smoke.log
python bench.py > gpurun_out/r1f_bench.json 2> gpurun_out/r1f_bench.err
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r1f_bench_ref.json 2> gpurun_out/r1f_bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1f_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu --push-maps 64 > gpurun_out/r1f_ncu.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_bytes.sum --clock-control none --csv --log-file gpurun_out/r1f_kld_launches.csv python scratch/kld_quick.py > gpurun_out/r1f_ncukld.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:kld_feat_dz_kernel -s 2 -c 1 -f -o gpurun_out/r1f_kld_dz_full python scratch/kld_quick.py > gpurun_out/r1f_ncudz.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:kld_feat_dp_kernel -s 2 -c 1 -f -o gpurun_out/r1f_kld_dp_full python scratch/kld_quick.py > gpurun_out/r1f_ncudp.log 2>&1
tail -4 gpurun_out/r1f_tests.log; tail -2 gpurun_out/r1f_smoke.log; ls -la gpurun_out/*.ncu-rep
