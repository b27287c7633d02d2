#!/bin/bash
mkdir -p gpurun_out
cat > /tmp/one.py <<'PY'
import sys; sys.path.insert(0,'.')
import libsnark_b200 as b
from oracle import refo
b.init(0)
ref=refo.RefOracle()
n=1<<16
bases=ref.make_bases("alt_bn128",refo.G2,n,1); sc=ref.field_random("alt_bn128",refo.FR,0,n)
rb=b.ResidentBases("alt_bn128",refo.G2,bases)
for it in range(2): rb.multi_exp(sc)
print(b.last_msm_timing())
PY
python /tmp/one.py && ncu --set full --clock-control none --import-source on -k regex:k_reduce_chunks -s 1 -c 1 -o gpurun_out/prof_reduce_g2_r1a python /tmp/one.py > gpurun_out/ncu_full2.log 2>&1
tail -3 gpurun_out/ncu_full2.log
