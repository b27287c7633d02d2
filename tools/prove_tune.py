"""Slot-plan experiment for b200_groth16_prove (B200_MSM_SLOTS / B200_WM_FIRST, csrc/groth16.cu): one key, several plans."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import libsnark_b200 as b  # noqa: E402

b.init(0)
PLANS = [("2,1,2,0,3", "1"), ("2,1,2,0,3", "0"), ("0,1,0,0,0", "1"), ("2,1,2,2,2", "1"), ("2,1,3,0,3", "1"), ("1,1,1,1,1", "1"),
         ("0,0,0,0,0", "1"), ("2,1,2,3,0", "1")]
log_m = int(sys.argv[1]) if len(sys.argv) > 1 else 20
for plan, wm in PLANS:
    os.environ["B200_MSM_SLOTS"], os.environ["B200_WM_FIRST"] = plan, wm
    r = bench.bench_groth16(b, log_m, 6, 3)
    print(json.dumps({"slots(A,Bg2,Bg1,H,L)": plan, "wm_first": wm, "ms": round(r["value"], 3), "min_ms": round(r["min_ms"], 3),
                      "stage_ms": {k: round(v, 2) for k, v in r["stage_ms"].items()}}), flush=True)
