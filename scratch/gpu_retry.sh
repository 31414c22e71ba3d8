#!/bin/bash
# usage: gpu_retry.sh <timeout> <script> ; retries while the pod answers busy (rc 3)
for i in 1 2 3 4 5 6 7 8 9 10; do
  /usr/local/graft/bin/gpurun --timeout $1 -- "bash $2" > gpurun_out/.retry_$$.log 2>&1
  rc=$?
  if grep -q "status=transient\|rc=3" gpurun_out/.retry_$$.log && ! grep -q "status=ok" gpurun_out/.retry_$$.log; then sleep 90; continue; fi
  break
done
cat gpurun_out/.retry_$$.log | tail -40
rm -f gpurun_out/.retry_$$.log
