#!/bin/bash
# Round 2, call Z9: L2::64B gather loads only (ownership and AFF_KB as before): parity subset, stage times, prove, traffic.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_msm_gpu.py tests/test_kat_gpu.py -q -m gpu -x 2>&1 | tail -2
{
python tools/affine_probe.py --check-log 12 2>/dev/null | tail -1
python tools/affine_probe.py --check-log 12 --group 2 2>/dev/null | tail -1
python tools/affine_probe.py --check-log 12 --curve bls12_377 2>/dev/null | tail -1
python tools/affine_probe.py --check-log 12 --curve bw6_761 --log-n 18 2>/dev/null | tail -1
B200_AFFINE_ROUNDS=0 python tools/affine_probe.py --check-log 12 2>/dev/null | tail -1
python tools/prove_quick.py --steps 10 2>/dev/null | tail -1
} | tee gpurun_out/r2_aff_l2_64b.jsonl
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --kernel-name regex:'k_aff_up_round|k_aff_round|k_accumulate' --launch-skip 14 --launch-count 7 --csv --log-file gpurun_out/r2_aff_l2_64b_ncu.csv python tools/affine_probe.py --check-log 10 > /dev/null 2>&1
B200_AFFINE_ROUNDS=0 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --kernel-name regex:'k_accumulate' --launch-skip 2 --launch-count 1 --csv --log-file gpurun_out/r2_acc_l2_64b_ncu.csv python tools/affine_probe.py --check-log 10 > /dev/null 2>&1
python - <<'PY'
import csv
for f in ['gpurun_out/r2_aff_l2_64b_ncu.csv','gpurun_out/r2_acc_l2_64b_ncu.csv']:
    rows=list(csv.reader(l for l in open(f) if l.startswith('"')))
    hdr=rows[0]; k=hdr.index("Kernel Name"); mn=hdr.index("Metric Name"); mv=hdr.index("Metric Value"); idc=hdr.index("ID")
    out={}
    for r in rows[1:]:
        out.setdefault((r[idc], r[k].split('<')[0][-18:]),{})[r[mn].split('__')[1][:18]]=r[mv]
    for (i,kn),v in out.items(): print(i, kn, v)
PY
