"""GPU parity of the device field / group arithmetic (fp.cuh, curve.cuh) against libff's Fp_model,
Fp2_model and group law, element-wise on the reference's SHA512_rng inputs. Bit-exact.
Edge-case list from libff test_groups.cpp:24-57 (O+P, P+O, P+P, P+(-P))."""
import numpy as np
import pytest

from oracle import refo

pytestmark = pytest.mark.gpu

CURVES = ["alt_bn128", "bls12_377", "bw6_761"]


def _rand_field(ref, curve, field, n, seed):
    if field == refo.FQ2:
        a = ref.field_random(curve, refo.FQ, seed, 2 * n)
        return a.reshape(n, -1)
    return ref.field_random(curve, field, seed, n)


@pytest.mark.parametrize("curve", CURVES)
@pytest.mark.parametrize("field", [refo.FR, refo.FQ, refo.FQ2])
def test_field_ops(ref, b200, curve, field):
    if curve == "bw6_761" and field == refo.FQ2:
        pytest.skip("bw6_761 G2 is over Fq")
    n = 2000
    a = _rand_field(ref, curve, field, n, 10)
    b = _rand_field(ref, curve, field, n, 5000)
    # edge values: 0, 1 (Montgomery), p - 1
    L = a.shape[1]
    one = ref.field_consts(curve, refo.FR if field == refo.FR else refo.FQ)["one"]
    a[0] = 0
    a[1, :len(one)] = one
    a[1, len(one):] = 0
    a[2] = ref.field_op(curve, field, refo.OP_NEG, a[1:2])[0]
    b[0] = 0
    b[3] = a[3]
    for op in [refo.OP_ADD, refo.OP_SUB, refo.OP_MUL, refo.OP_SQR, refo.OP_NEG]:
        want = ref.field_op(curve, field, op, a, b)
        got = b200.test_field_op(curve, field, op, a, b)
        assert np.array_equal(got, want), (curve, field, op)
    k = 64
    want = ref.field_op(curve, field, refo.OP_INV, a[:k])
    got = b200.test_field_op(curve, field, refo.OP_INV, a[:k])
    assert np.array_equal(got, want)
    assert L == a.shape[1]


@pytest.mark.parametrize("curve", CURVES)
@pytest.mark.parametrize("field", [refo.FR, refo.FQ])
def test_montgomery_conversion(ref, b200, curve, field):
    a = ref.field_random(curve, field, 77, 500)
    std = ref.from_mont(curve, field, a)
    assert np.array_equal(b200.test_fp_convert(curve, field, False, a), std)
    assert np.array_equal(b200.test_fp_convert(curve, field, True, std), a)


@pytest.mark.parametrize("curve", CURVES)
@pytest.mark.parametrize("group", [refo.G1, refo.G2])
def test_group_law(ref, b200, curve, group):
    n = 256
    A = ref.make_bases(curve, group, n, 31)
    B = ref.make_bases(curve, group, n, 32)
    zero = ref.group_op(curve, group, refo.GOP_TO_AFFINE, np.zeros((1, A.shape[1]), np.uint64))[0]
    A[0] = zero                       # O + P
    B[1] = zero                       # P + O
    A[2] = zero
    B[2] = zero                       # O + O
    B[3] = A[3]                       # P + P -> dbl
    B[4] = ref.group_op(curve, group, refo.GOP_NEG, A[4:5])[0]   # P + (-P) -> O

    def norm(x):
        return ref.group_op(curve, group, refo.GOP_TO_AFFINE, x)

    add = norm(ref.group_op(curve, group, refo.GOP_ADD, A, B))
    assert np.array_equal(b200.test_group_op(curve, group, 0, A, B), add)
    assert np.array_equal(b200.test_group_op(curve, group, 1, A, B), add)
    dbl = ref.group_op(curve, group, refo.GOP_DBL, A)
    assert np.array_equal(b200.test_group_op(curve, group, 2, A, B), norm(dbl))
    twoA_B = norm(ref.group_op(curve, group, refo.GOP_ADD, dbl, B))
    assert np.array_equal(b200.test_group_op(curve, group, 3, A, B), twoA_B)
    assert np.array_equal(b200.test_group_op(curve, group, 4, A, B), twoA_B)
    A_2B = norm(ref.group_op(curve, group, refo.GOP_ADD, A, ref.group_op(curve, group, refo.GOP_DBL, B)))
    assert np.array_equal(b200.test_group_op(curve, group, 5, A, B), A_2B)
