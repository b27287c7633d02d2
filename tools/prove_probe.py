"""One-off probe: b200_groth16_prove on the bench's synthetic instance (for ncu launch lists)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import libsnark_b200 as b  # noqa: E402
import bench  # noqa: E402

log_m = int(sys.argv[1]) if len(sys.argv) > 1 else 20
b.init(0)
print(bench.bench_groth16(b, log_m, 3, 3))
