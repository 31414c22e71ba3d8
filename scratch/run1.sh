mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tma.py -q --timeout 90 2>&1 | tail -40 > gpurun_out/r2a_tma_tests.log; tail -25 gpurun_out/r2a_tma_tests.log
timeout 300 python scratch/bench_modes.py > gpurun_out/r2a_modes.log 2>&1; cat gpurun_out/r2a_modes.log
