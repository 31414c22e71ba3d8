for sh in one small c2; do for m in f16 f16x3; do for w in fwd fwdd bwd; do
  echo "== $w $m $sh"; timeout 40 python scratch/dbg_tma.py $w $m $sh 2>&1 | tail -2; echo "rc=$?"
done; done; done
