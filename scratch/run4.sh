python -m pytest tests -q -m gpu --timeout 120 --tb=line 2>&1 | tail -15 > gpurun_out/r2d_tests.log; cat gpurun_out/r2d_tests.log
