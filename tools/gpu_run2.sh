#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_field_gpu.py -q -m gpu 2>&1 | tail -40
cat > /tmp/small.py <<'PY'
import sys; sys.path.insert(0,'.')
import libsnark_b200 as b
from oracle import refo
b.init(0)
ref=refo.RefOracle()
for curve,group,logn in [("alt_bn128",refo.G1,20),("alt_bn128",refo.G2,16),("bls12_377",refo.G1,16)]:
    n=1<<logn
    bases=ref.make_bases(curve,group,n,1); sc=ref.field_random(curve,refo.FR,0,n)
    rb=b.ResidentBases(curve,group,bases)
    for it in range(2): rb.multi_exp(sc)
    print(curve,group,logn,b.last_msm_timing())
PY
python /tmp/small.py && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_r1_msm_probe.csv python /tmp/small.py > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
