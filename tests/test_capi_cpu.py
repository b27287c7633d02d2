"""CPU tier: the C-ABI library loads, exports every symbol include/libsnark_b200.h declares, validates
arguments, and refuses to compute without a GPU (no CPU fallback). Plus the host-side tail of the
multi-GPU path: per-shard window sums gathered with torch.distributed (gloo, world_size 2) and folded."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "libsnark_b200.h")


@pytest.fixture(scope="module")
def b():
    import libsnark_b200 as m
    if not os.path.exists(os.path.join(ROOT, "libsnark_b200", "libsnark_b200.so")):
        import __graft_entry__
        __graft_entry__.build()
    return m


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_A-Z]+)\s*\(", text)))


def test_exports_every_declared_symbol(b):
    L = b.lib()
    names = declared_symbols()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert set(b.EXPORTS) <= set(names), sorted(set(b.EXPORTS) - set(names))


def test_sizes_and_argument_validation(b):
    assert b.sizeof_fr("alt_bn128") == 32 and b.sizeof_fr("bw6_761") == 48
    assert b.sizeof_group("alt_bn128", b.G1) == 96 and b.sizeof_group("alt_bn128", b.G2) == 192
    assert b.sizeof_group("bls12_377", b.G1) == 144 and b.sizeof_group("bls12_377", b.G2) == 288
    assert b.sizeof_group("bw6_761", b.G1) == 288 and b.sizeof_group("bw6_761", b.G2) == 288
    L = b.lib()
    assert L.b200_msm(7, 1, None, 96, 0, None, None) == b.ERR_ARG
    assert L.b200_sizeof_fr(9) == 0
    assert L.b200_msm_num_windows(0, 1 << 20) * L.b200_msm_window_bits(0, 1 << 20) >= 255
    assert L.b200_bases_free(12345) == b.ERR_HANDLE


def test_no_cpu_fallback(b):
    """Without a CUDA device every compute entry point fails loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    a = np.zeros((8, 4), dtype=np.uint64)
    with pytest.raises(b.B200Error) as e:
        b.get_evaluation_domain("alt_bn128", 8).FFT(a)
    assert e.value.code == b.ERR_CUDA and "no CPU fallback" in str(e.value)
    with pytest.raises(b.B200Error):
        b.multi_exp("alt_bn128", b.G1, np.zeros((2, 12), np.uint64), np.zeros((2, 4), np.uint64))


def _fake_windows(P, curve, group, W, shard, glimbs):
    """Window sums of a pretend shard: affine points in the library's X|Y|ZZ|ZZZ form (ZZ = ZZZ = one)."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "msm.npz"))
    pts = g["family.bases"][shard * W:(shard + 1) * W]        # alt_bn128 G1: X | Y | Z(one)
    one = pts[:, 8:12]
    return np.concatenate([pts[:, 0:8], one, one], axis=1), pts


def test_fold_windows_host(b):
    """b200_msm_fold_windows (host_fold.hpp) == sum_w 2^(c w) sum_shards R_{s,w}, checked with the C oracle."""
    from oracle import port
    P = port.OraclePort()
    n_per_shard = 1 << 10
    L = b.lib()
    W, c = L.b200_msm_num_windows(0, n_per_shard), L.b200_msm_window_bits(0, n_per_shard)
    shards = 2
    wins, pts = zip(*[_fake_windows(P, "alt_bn128", 1, W, s, 12) for s in range(shards)])
    got = b.fold_windows("alt_bn128", b.G1, n_per_shard, np.concatenate(wins), shards)
    acc = None
    for w in range(W - 1, -1, -1):
        if acc is not None:
            for _ in range(c):
                acc = P.group_op("alt_bn128", 1, 2, acc)
        for s in range(shards):
            p = pts[s][w:w + 1]
            acc = p if acc is None else P.group_op("alt_bn128", 1, 0, acc, p)
    assert np.array_equal(got, acc[0])


WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
import libsnark_b200 as b
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
g = np.load(os.path.join(sys.argv[1], "tests", "golden", "msm.npz"))
n = 1 << 10
W = b.lib().b200_msm_num_windows(0, n)
pts = g["family.bases"][rank * W:(rank + 1) * W]
one = pts[:, 8:12]
mine = torch.from_numpy(np.concatenate([pts[:, 0:8], one, one], axis=1).astype(np.int64).reshape(-1))
gathered = [torch.empty_like(mine) for _ in range(world)]
dist.all_gather(gathered, mine)                      # the NVLink/NCCL gather of bench.py, on gloo here
if rank == 0:
    allw = torch.cat(gathered).numpy().view(np.uint64)
    out = b.fold_windows("alt_bn128", b.G1, n, allw, world)
    np.save(sys.argv[2], out)
dist.barrier()
dist.destroy_process_group()
'''


def test_sharded_fold_gloo_world2(b, tmp_path):
    """world_size-2 run of the multi-GPU host logic: gather per-rank window sums, fold on rank 0."""
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    out = tmp_path / "out.npy"
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
    subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                    "--master-addr", "127.0.0.1", "--master-port", "29533", str(script), ROOT, str(out)],
                   check=True, env=env, timeout=300)
    from oracle import port
    P = port.OraclePort()
    n = 1 << 10
    L = b.lib()
    W, c = L.b200_msm_num_windows(0, n), L.b200_msm_window_bits(0, n)
    wins = [_fake_windows(P, "alt_bn128", 1, W, s, 12)[0] for s in range(2)]
    want = b.fold_windows("alt_bn128", b.G1, n, np.concatenate(wins), 2)
    assert np.array_equal(np.load(out), want)


def test_product_does_not_touch_the_oracle():
    """The shipped path (package + shared library) never imports, links or opens anything under oracle/."""
    import subprocess
    pkg = os.path.join(ROOT, "libsnark_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")) and "overlay" not in dirpath:
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                for needle in ("import oracle", "from oracle", "oracle/", "libref_", "oracle_port", "refo"):
                    assert needle not in text, f"{f} refers to {needle}"
    needed = subprocess.run(["readelf", "-d", os.path.join(pkg, "libsnark_b200.so")], capture_output=True, text=True).stdout
    assert "libref" not in needed and "oracle" not in needed


def test_synthetic_r1cs_is_satisfied():
    """tools/synth_r1cs.py (the bench's stand-in for generate_r1cs_example_with_field_input): <a,w> * <b,w> = <c,w>."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import synth_r1cs
    for curve in ("alt_bn128", "bw6_761"):
        s = synth_r1cs.make(curve, 50, 4)
        p, Rinv = s["modulus"], pow(s["R"], -1, s["modulus"])
        w = np.concatenate([s["one"][None, :], s["primary"], s["auxiliary"]])
        ints = [sum(int(x) << (64 * i) for i, x in enumerate(row)) * Rinv % p for row in w]
        assert ints[0] == 1 and len(ints) == s["nv"] + 1
        nc, rp, col = s["nc"], s["row_ptr"], s["col"]
        ev = [sum(ints[c] for c in col[rp[r]:rp[r + 1]]) % p for r in range(3 * nc)]
        for i in range(nc):
            assert ev[i] * ev[nc + i] % p == ev[2 * nc + i], (curve, i)
