set -x
python -m pytest tests/test_gpu_kld.py tests/test_gpu_dropin.py -x -q > gpurun_out/r1k_tests.log 2>&1
tail -25 gpurun_out/r1k_tests.log
python -m pytest tests/test_gpu_parity.py -x -q -k single_scale > gpurun_out/r1k_tests2.log 2>&1
tail -25 gpurun_out/r1k_tests2.log
cat > /tmp/kq.py <<'PY'
import sys, torch
sys.path.insert(0, '.')
import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
from scaleprotoseg_b200.kld import KldTables, kld_loss
B,H,W,P,K,S=16,65,65,760,19,4
dev='cuda:0'
tables=KldTables(ho.make_identity(P,K,S), ho.uniform_ranges(P,S), dev)
labels=torch.stack([ho.make_labels(i,H,W,K) for i in range(B)]).to(dev)
dist=(torch.rand(B,P,H,W,device=dev)*6).requires_grad_(True)
for i in range(4):
    dist.grad=None
    kld_loss(dist,labels,tables).backward()
torch.cuda.synchronize()
PY
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed --clock-control none --csv --log-file gpurun_out/r1k_kld_launches.csv python /tmp/kq.py > gpurun_out/r1k_ncu.log 2>&1
tail -3 gpurun_out/r1k_ncu.log
