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


def test_domain_kind_follows_the_factory(b):
    """b200_domain_kind = valid_for_size of basic / extended / step in the order get_evaluation_domain.tcc:37-58 tries them
    (host logic, no GPU); b200_set_two_adicity moves the boundaries the way lowering FieldT::s does."""
    K = b.domain_kind
    for curve, s in [("alt_bn128", 28), ("bls12_377", 47), ("bw6_761", 46)]:
        assert K(curve, 1) == b.DOMAIN_NONE and K(curve, 2) == b.DOMAIN_BASIC
        assert K(curve, 1 << 20) == b.DOMAIN_BASIC and K(curve, 1 << min(s, 40)) == b.DOMAIN_BASIC
        assert K(curve, (1 << 20) + 16) == b.DOMAIN_STEP and K(curve, 3) == b.DOMAIN_STEP
        assert K(curve, (1 << 20) + 11) == b.DOMAIN_NONE          # the factory rounds this one up first
    assert K("alt_bn128", 1 << 29) == b.DOMAIN_EXTENDED and K("alt_bn128", 1 << 30) == b.DOMAIN_NONE
    assert K("alt_bn128", (1 << 28) + 4) == b.DOMAIN_NONE        # step needs k + 1 <= s
    assert K("alt_bn128", (1 << 27) + 4) == b.DOMAIN_STEP
    b.set_two_adicity("alt_bn128", 9)
    try:
        assert K("alt_bn128", 512) == b.DOMAIN_BASIC and K("alt_bn128", 1024) == b.DOMAIN_EXTENDED
        assert K("alt_bn128", 2048) == b.DOMAIN_NONE and K("alt_bn128", 256 + 8) == b.DOMAIN_STEP
        assert K("alt_bn128", 512 + 8) == b.DOMAIN_NONE
        assert K("bls12_377", 2048) == b.DOMAIN_BASIC             # per field
    finally:
        b.set_two_adicity("alt_bn128", 0)
    assert K("alt_bn128", 1024) == b.DOMAIN_BASIC
    with pytest.raises(b.B200Error):
        b.set_two_adicity("alt_bn128", 29)


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
    with pytest.raises(b.B200Error) as e:
        b.batch_exp("alt_bn128", b.G1, np.zeros(12, np.uint64), np.zeros((3, 4), np.uint64))
    assert e.value.code == b.ERR_CUDA
    L = b.lib()
    assert L.b200_bases_precompute(4242) == b.ERR_HANDLE and L.b200_bases_table_rows(4242) == 0
    assert L.b200_groth16_shard_msm(4242, 31, None) == b.ERR_HANDLE


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


# ---- sharded prover: host logic (b200_shard_range, blob layout, b200_groth16_assemble) -----------------

QUERY_SIZES = [2001, 1900, 2047, 1991]          # A, B, H, L totals of a pretend key (alt_bn128)


def _fake_blob(b, world, rank):
    """One rank's blob: window sums of its five pretend partial MSMs = distinct affine golden points
    (X | Y | ZZ = ZZZ = one). Returns (blob bytes, {msm k: affine points in the reference layout})."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "msm.npz"))
    g1, g2 = g["rand.alt_bn128.g1.bases"], g["rand.alt_bn128.g2.bases"]
    L = b.lib()
    blob = np.zeros(b.groth16_shard_bytes("alt_bn128", QUERY_SIZES, world), dtype=np.uint8)
    pts, at = {}, 0
    for k, (q, grp) in enumerate(zip([0, 1, 1, 2, 3], [1, 2, 1, 1, 1])):
        wmax = max(L.b200_msm_num_windows(0, b.shard_range(QUERY_SIZES[q], world, r)[1]) for r in range(world))
        W = L.b200_msm_num_windows(0, b.shard_range(QUERY_SIZES[q], world, rank)[1])
        src = g1 if grp == 1 else g2
        fl = 4 if grp == 1 else 8                  # u64 limbs per coordinate
        p = src[(rank * 5 + k) * 3:(rank * 5 + k) * 3 + W] if W <= 3 else np.resize(src[(rank * 5 + k):], (W, src.shape[1]))
        one = p[:, 2 * fl:3 * fl]
        win = np.ascontiguousarray(np.concatenate([p[:, :2 * fl], one, one], axis=1))
        blob[at:at + win.nbytes] = win.view(np.uint8).reshape(-1)
        at += wmax * L.b200_msm_window_bytes(0, grp)
        pts[k] = p
    assert at == blob.shape[0]
    return blob, pts


def _expected_proof(b, P, world, all_pts, r=None, s=None):
    """The proof of r1cs_gg_ppzksnark.tcc:583-595 from every rank's windows (Horner) with the C oracle's group law:
    g_A = alpha + A + r delta, g_B = beta + B + s delta (G2 and G1), g_C = H + L + s g_A + r g_B(G1) - r s delta."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "msm.npz"))
    L = b.lib()

    def fold(k, q, grp):
        total = None
        for r in range(world):
            n = b.shard_range(QUERY_SIZES[q], world, r)[1]
            W, c = L.b200_msm_num_windows(0, n), L.b200_msm_window_bits(0, n)
            acc = None
            for w in range(W - 1, -1, -1):
                if acc is not None:
                    for _ in range(c):
                        acc = P.group_op("alt_bn128", grp, 2, acc)
                p = all_pts[r][k][w:w + 1]
                acc = p if acc is None else P.group_op("alt_bn128", grp, 0, acc, p)
            total = acc if total is None else P.group_op("alt_bn128", grp, 0, total, acc)
        return total
    key = _key_points(g)
    add = lambda grp, x, y: P.group_op("alt_bn128", grp, 0, x, y)
    gA = add(1, key[0][None, :], fold(0, 0, 1))
    gB = add(2, key[2][None, :], fold(1, 1, 2))
    gC = add(1, fold(3, 2, 1), fold(4, 3, 1))
    if r is not None:
        mul = lambda grp, k, pt: P.msm("alt_bn128", grp, np.ascontiguousarray(pt).reshape(1, -1), k[None, :])[None, :]
        gB1 = add(1, add(1, key[1][None, :], fold(2, 1, 1)), mul(1, s, key[3]))
        gA = add(1, gA, mul(1, r, key[3]))
        gB = add(2, gB, mul(2, s, key[4]))
        rs = P.field_op("alt_bn128", 0, 2, r[None, :], s[None, :])[0]
        neg = P.group_op("alt_bn128", 1, 3, mul(1, rs, key[3]))
        gC = add(1, add(1, add(1, gC, mul(1, s, gA[0])), mul(1, r, gB1[0])), neg)
    return np.concatenate([gA[0], gB[0], gC[0]])


def _key_points(g):
    g1, g2 = g["rand.alt_bn128.g1.bases"], g["rand.alt_bn128.g2.bases"]
    return [g1[90], g1[91], g2[90], g1[92], g2[91]]      # alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2


def test_shard_range_covers_every_query(b):
    for total in (0, 1, 7, 8, 9, 1000, (1 << 20) - 1):
        for world in (1, 2, 3, 8):
            at = 0
            for r in range(world):
                f, c = b.shard_range(total, world, r)
                assert f == at
                at += c
            assert at == total


def test_sharded_assemble_host(b):
    """b200_groth16_assemble on blobs of 1, 2 and 3 pretend ranks (ragged shards, different window geometry per
    query) against the C oracle's group law."""
    from oracle import port
    P = port.OraclePort()
    g = np.load(os.path.join(ROOT, "tests", "golden", "msm.npz"))
    zero = np.zeros(4, dtype=np.uint64)
    for world in (1, 2, 3):
        blobs, pts = zip(*[_fake_blob(b, world, r) for r in range(world)])
        got = b.groth16_assemble("alt_bn128", *_key_points(g), QUERY_SIZES, world, np.concatenate(blobs), zero, zero)
        assert np.array_equal(got, _expected_proof(b, P, world, pts)), world
        r, s = g["rand.alt_bn128.g1.scalars"][40 + world], g["rand.alt_bn128.g1.scalars"][50 + world]
        got = b.groth16_assemble("alt_bn128", *_key_points(g), QUERY_SIZES, world, np.concatenate(blobs), r, s)
        assert np.array_equal(got, _expected_proof(b, P, world, pts, r, s)), world


SHARD_WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import libsnark_b200 as b
import test_capi_cpu as T
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
blob, _ = T._fake_blob(b, world, rank)
mine = torch.from_numpy(blob)
gathered = torch.empty(world * mine.numel(), dtype=torch.uint8)
dist.all_gather_into_tensor(gathered, mine)          # DistributedProver's NCCL all-gather, on gloo here
if rank == 0:
    g = np.load(os.path.join(sys.argv[1], "tests", "golden", "msm.npz"))
    zero = np.zeros(4, dtype=np.uint64)
    np.save(sys.argv[2], b.groth16_assemble("alt_bn128", *T._key_points(g), T.QUERY_SIZES, world, gathered.numpy(), zero, zero))
dist.barrier()
dist.destroy_process_group()
'''


def test_sharded_prover_gloo_world2(b, tmp_path):
    """world_size-2 run of the sharded prover's host side: every rank's blob all-gathered, proof assembled on rank 0."""
    script = tmp_path / "worker.py"
    script.write_text(SHARD_WORKER)
    out = tmp_path / "proof.npy"
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29537")
    subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                    "--master-addr", "127.0.0.1", "--master-port", "29537", str(script), ROOT, str(out)],
                   check=True, env=env, timeout=300)
    from oracle import port
    P = port.OraclePort()
    pts = [_fake_blob(b, 2, r)[1] for r in range(2)]
    assert np.array_equal(np.load(out), _expected_proof(b, P, 2, pts))


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


@pytest.mark.parametrize("curve,nc", [("alt_bn128", 90), ("bls12_377", 59), ("bw6_761", 27)])
def test_assemble_reproduces_reference_proofs(b, curve, nc):
    """b200_groth16_assemble (host fold + r1cs_gg_ppzksnark.tcc:583-595, incl. the G2 arithmetic of each curve) on the
    committed fixtures: the five MSM results — computed here by the C oracle — enter as one-shard blobs with the whole
    sum in window 0, and the assembled proof must be the reference prover's (tests/golden/groth16.npz)."""
    from oracle import port
    from test_oracle_cpu import groth16_fixture
    P = port.OraclePort()
    key, r1cs, w, r, s, proof = groth16_fixture(curve, nc)
    ncons, ni, nv, m = r1cs["nc"], r1cs["ni"], r1cs["nv"], r1cs["m"]
    aA, aB, aC = P.r1cs_evaluate(curve, ncons, ni, r1cs["row_ptr"], r1cs["col"], r1cs["val"], w, m)
    Hc = P.witness_map_h(curve, aA, aB, aC)
    Bg2, Bg1 = P.kc_msm(curve, key["B_indices"], key["B_values"], 0, nv + 1, w)
    sums = [P.msm(curve, 1, key["A"], w, method=2), Bg2, Bg1, P.msm(curve, 1, key["H"][:m - 1], Hc[:m - 1]),
            P.msm(curve, 1, key["L"], w[ni + 1:], method=2)]
    sizes = [nv + 1, key["B_indices"].shape[0], m - 1, nv - ni]
    L = b.lib()
    cid = b.CURVES[curve]
    blob = np.zeros(b.groth16_shard_bytes(curve, sizes, 1), dtype=np.uint8)
    at = 0
    for k, (q, grp) in enumerate(zip([0, 1, 1, 2, 3], [1, 2, 1, 1, 1])):
        pt = sums[k]
        fl = pt.shape[0] // 3                       # u64 limbs per coordinate
        assert pt[2 * fl:].any(), "an MSM of the fixture is the point at infinity"
        one = pt[2 * fl:]                           # Z of a normalised point = one
        win = np.ascontiguousarray(np.concatenate([pt[:2 * fl], one, one]))       # X | Y | ZZ | ZZZ
        assert win.nbytes == L.b200_msm_window_bytes(cid, grp)
        blob[at:at + win.nbytes] = win.view(np.uint8)
        at += L.b200_msm_num_windows(cid, sizes[q]) * win.nbytes
    assert at == blob.shape[0]
    got = b.groth16_assemble(curve, key["alpha_g1"], key["beta_g1"], key["beta_g2"], key["delta_g1"], key["delta_g2"], sizes, 1,
                             blob, r, s)
    assert np.array_equal(got, proof)


@pytest.mark.parametrize("extra", [[], ["-DB200_OVERLAY_KEEP_REFERENCE_PROVER"]])
def test_overlay_header_compiles_against_the_reference(extra):
    """libsnark_b200_overlay.hpp force-included into an empty translation unit, with the reference's own headers: every
    explicit specialisation must match a primary template of the reference (a drifted signature is a hard error), with and
    without the prover entry-point specialisation. Needs the reference tree (absent on the GPU box: skipped there)."""
    ref = "/root/reference"
    if not os.path.isdir(os.path.join(ref, "libsnark")):
        pytest.skip("reference tree not present")
    src = os.path.join(ROOT, "tests", "_overlay_syntax.cpp")
    with open(src, "w") as f:
        f.write("int main() { return 0; }\n")
    try:
        cmd = ["g++", "-std=c++11", "-fsyntax-only", "-fopenmp", "-w", "-DNO_PROCPS", "-DBINARY_OUTPUT", "-DMONTGOMERY_OUTPUT",
               "-DUSE_ASM", "-DMULTICORE", "-DCURVE_ALT_BN128", *extra,
               "-I", os.path.join(ROOT, "oracle", "gmp_shim"), "-I", ref, "-I", os.path.join(ref, "depends", "libff"),
               "-I", os.path.join(ref, "depends", "libfqfft"), "-I", os.path.join(ROOT, "include"),
               "-I", os.path.join(ROOT, "libsnark_b200", "overlay"),
               "-include", os.path.join(ROOT, "libsnark_b200", "overlay", "libsnark_b200_overlay.hpp"), src]
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stderr[-3000:]
    finally:
        os.remove(src)
