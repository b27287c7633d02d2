"""Synthetic satisfiable R1CS of the shape of libsnark's generate_r1cs_example_with_field_input
(libsnark/relations/constraint_satisfaction_problems/r1cs/examples/r1cs_examples.tcc:25-100), built
without the reference so that bench.py can run on a box that has only this repo:

    variables x_0 = a, x_1 = b (random), then alternately x_{i+2} = x_i * x_{i+1} / x_{i+2} = x_i + x_{i+1};
    the last constraint sums all earlier variables. num_variables = num_constraints + 2.
    The first num_inputs variables are the primary input.

Returns the CSR that b200_r1cs_upload takes (rows = a | b | c combinations, column 0 = the constant ONE,
column k = variable k) and the witness in Montgomery form."""
import numpy as np

MODULI = {
    # scalar-field moduli (libff <curve>_init.cpp); R = 2^(64 * limbs)
    "alt_bn128": (21888242871839275222246405745257275088548364400416034343698204186575808495617, 4),
    "bls12_377": (8444461749428370424248824938781546531375899335154063827935233455917409239041, 4),
    "bw6_761": (258664426012969094010652733694893533536393512754914660539884262666720468348340822774968888139573360124440321458177, 6),
}


def _limbs(vals, n64):
    buf = b"".join(int(v).to_bytes(8 * n64, "little") for v in vals)
    return np.frombuffer(buf, dtype=np.uint64).reshape(-1, n64).copy()


def make(curve, num_constraints, num_inputs, seed=1):
    p, n64 = MODULI[curve]
    R = (1 << (64 * n64)) % p
    rng = np.random.default_rng(seed)
    nc, nv = num_constraints, num_constraints + 2
    a = int.from_bytes(rng.bytes(8 * n64), "little") % p
    b = int.from_bytes(rng.bytes(8 * n64), "little") % p
    x = [a, b]
    # CSR rows for A, B, C: lists of (col, coeff) with coeff in {1}
    rows_a, rows_b, rows_c = [], [], []
    for i in range(nc - 1):
        if i % 2:
            x.append(x[i] * x[i + 1] % p)
            rows_a.append([i + 1])
            rows_b.append([i + 2])
        else:
            x.append((x[i] + x[i + 1]) % p)
            rows_a.append([i + 1, i + 2])
            rows_b.append([0])
        rows_c.append([i + 3])
    # last constraint: (sum of all variables so far) * 1 = x_last
    x.append(sum(x) % p)
    rows_a.append(list(range(1, nv)))
    rows_b.append([0])
    rows_c.append([nv])
    assert len(x) == nv
    rows = rows_a + rows_b + rows_c
    counts = np.fromiter((len(r) for r in rows), dtype=np.int64, count=3 * nc)
    row_ptr = np.zeros(3 * nc + 1, dtype=np.uint32)
    row_ptr[1:] = np.cumsum(counts)
    col = np.fromiter((c for r in rows for c in r), dtype=np.uint32, count=int(counts.sum()))
    one = _limbs([R], n64)[0]
    val = np.tile(one, (col.shape[0], 1))
    w = _limbs([v * R % p for v in x], n64)
    return dict(curve=curve, nc=nc, ni=num_inputs, nv=nv, row_ptr=row_ptr, col=col, val=val,
                primary=w[:num_inputs].copy(), auxiliary=w[num_inputs:].copy(), one=one, modulus=p, R=R)
