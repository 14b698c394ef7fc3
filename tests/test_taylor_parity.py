"""Taylor-coefficient propagation: implementation under test vs the oracle (the
unmodified reference, oracle/_ref) on identical seeded inputs.

Mirrors the reference's white-box check `check_taylor_prop`
(tests/symbolic.cpp:89-137): for every order the bias, the pushed coefficient
and the Jacobian are compared.  Tolerance 1e-10 relative (BASELINE.json).
"""
import numpy as np
import pytest

from sanm_b200 import material as mat
from sanm_b200.capi import TaylorCoeffProp

TOL = 1e-10


def rel_err(a, b):
    scale = max(np.abs(b).max(), 1e-300)
    return np.abs(a - b).max() / scale


def run_prop(b, build, x_coeffs, order):
    """returns (y0, [bias_k], [y_k], jac_full)"""
    g = b.ComputingGraph()
    y = build(g)
    prop = TaylorCoeffProp(y, max_order=order)
    prop.push_xi(x_coeffs[0])
    y0 = prop.gather_yi()
    biases, ys = [], []
    jac = None
    for k in range(1, order + 1):
        biases.append(prop.compute_next_order_bias())
        if k == 1:
            jac = prop.jacobian_full()
        prop.push_xi(x_coeffs[k])
        ys.append(prop.gather_yi())
    return y0, biases, ys, jac


def compare(impl, ref, build, x_coeffs, order, tol=TOL):
    a = run_prop(impl, build, x_coeffs, order)
    r = run_prop(ref, build, x_coeffs, order)
    assert rel_err(a[0], r[0]) < tol, "y0"
    assert rel_err(a[3], r[3]) < tol, "jacobian"
    for k in range(order):
        scale = max(np.abs(r[2][k]).max(), 1e-300)
        assert np.abs(a[1][k] - r[1][k]).max() / scale < tol, f"bias {k + 1}"
        assert rel_err(a[2][k], r[2][k]) < tol, f"coeff {k + 1}"
    # the defining identity of the reference's check_taylor_prop:
    # J x_k + bias_k == y_k
    for k in range(order):
        jx = np.einsum("boi,bi->bo", a[3], x_coeffs[k + 1].reshape(len(a[3]), -1))
        lhs = jx.reshape(a[2][k].shape) + a[1][k]
        assert rel_err(lhs, a[2][k]) < 1e-9, f"identity at order {k + 1}"


def make_F_coeffs(rng, n, order, scale=0.5, amp=0.02):
    """F_0 = I + U(-0.2, 0.2) (det > 0), higher orders amp * scale^k * U(-1, 1).

    The reference's pow recurrence divides by x_0 (libsanm/analytic_unary.cpp:
    133-137), so rounding differences between two correct implementations grow
    like |x_1/x_0|^k; `amp` keeps that ratio O(1) for the elementwise F^2 of the
    incompressible neo-Hookean graph, as it is in real ANM runs where
    |x_1| ~ 1/sqrt(n).
    """
    xs = [np.eye(3)[None] + rng.uniform(-0.2, 0.2, (n, 3, 3))]
    for k in range(1, order + 1):
        xs.append(rng.uniform(-1, 1, (n, 3, 3)) * amp * scale ** k)
    return xs


MATERIAL = mat.MaterialProperty.from_young_poisson(1e6, 0.45)


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("energy", [mat.NEOHOOKEAN_I, mat.NEOHOOKEAN_C,
                                    mat.ARAP, mat.STVK_STRETCH])
def test_pk1_graphs(impl, ref, energy, fused, monkeypatch):
    """fused=True: the register-resident element kernels (csrc/fused_dev.h)
    take the three fea graphs; fused=False: the same graphs through the generic
    operator-list kernels (csrc/taylor_dev.h).  stvk_stretch is always generic."""
    monkeypatch.setenv("SANM_NO_FUSED", "0" if fused else "1")
    rng = np.random.default_rng(92702102)
    n, order = 37, 20
    xs = make_F_coeffs(rng, n, order)
    dminv = np.eye(3)[None] + rng.uniform(-0.1, 0.1, (n, 3, 3))
    bias = rng.uniform(-0.05, 0.05, (n, 3, 3))

    def build(g):
        Ds = g.placeholder() + g.constant(bias)
        F = Ds.batched_matmul(g.constant(dminv))
        return mat.pk1(energy, MATERIAL, F)

    compare(impl, ref, build, xs, order)


@pytest.mark.parametrize("energy", [mat.NEOHOOKEAN_I, mat.NEOHOOKEAN_C])
def test_cauchy_inverse_graphs(impl, ref, energy):
    rng = np.random.default_rng(5)
    n, order = 19, 10
    xs = make_F_coeffs(rng, n, order)
    ds = np.eye(3)[None] + rng.uniform(-0.1, 0.1, (n, 3, 3))

    def build(g):
        Dm = g.placeholder() + g.constant(np.zeros((n, 3, 3)))
        F = g.batched_mat_inv_mul(Dm, g.constant(ds), True)
        return mat.cauchy_stress(energy, MATERIAL, F)

    compare(impl, ref, build, xs, order)


def test_log_det_xtx(impl, ref):
    """Symbolic.LogDetTaylorProp, tests/symbolic.cpp:640-656"""
    rng = np.random.default_rng(1)
    n, order = 8, 13
    xs = [rng.uniform(-1, 1, (n, 3, 3)) + 2 * np.eye(3)[None]]
    xs += [rng.uniform(-1, 1, (n, 3, 3)) * 0.5 ** k for k in range(1, order + 1)]

    def build(g):
        x = g.placeholder()
        return x.batched_transpose().batched_matmul(x).batched_det().log()

    compare(impl, ref, build, xs, order)


def test_elemwise_and_reduce(impl, ref):
    """Symbolic.LinearCombination, tests/symbolic.cpp:301-322"""
    rng = np.random.default_rng(2)
    n, order = 9, 8
    xs = [rng.uniform(2, 5, (n, 4, 4))]
    xs += [rng.uniform(-1, 1, (n, 4, 4)) * 0.5 ** k for k in range(1, order + 1)]

    def build(g):
        x = g.placeholder()
        return g.linear_combine([(1.2, x.reduce_sum(-1)),
                                 (2.3, x.pow(2. / 3.)), (1.4, x.pow(1.5))], 2.5)

    compare(impl, ref, build, xs, order)


def test_matinvmul_nonsquare_matmul(impl, ref):
    """Symbolic.MatInvMul / TransMul shapes, tests/symbolic.cpp:179-233,408-426"""
    rng = np.random.default_rng(3)
    n, order = 6, 7
    xs = [rng.uniform(-1, 1, (n, 4, 4)) + 3 * np.eye(4)[None]]
    xs += [rng.uniform(-1, 1, (n, 4, 4)) * 0.5 ** k for k in range(1, order + 1)]
    a = rng.uniform(-1, 1, (n, 4, 4))
    c = rng.uniform(-1, 1, (n, 4, 6))

    def build(g):
        x = g.placeholder()
        y = g.batched_mat_inv_mul(x, g.constant(a), False)
        z = g.batched_mat_inv_mul(x * x, y, True)
        return z.batched_matmul(g.constant(c)).batched_transpose()

    compare(impl, ref, build, xs, order)


@pytest.mark.parametrize("m", [2, 4, 5, 7])
def test_batched_det_other_sizes(impl, ref, m):
    """Symbolic.BatchDeterminant for the sizes of tests/symbolic.cpp:324-358
    (`batched_det-mid` 4, `-large0` 5, `-large1` 7; y = det(x) * x): the
    reference expands the polynomial determinant (tensor_polymat.cpp:30-341),
    the device path integrates (det X)' = det X tr(X^-1 X') (taylor_dev.h
    op_det); both against each other at every order"""
    rng = np.random.default_rng(40 + m)
    n, order = 10, 10
    xs = [rng.uniform(-1, 1, (n, m, m)) + 2.5 * np.eye(m)[None]]
    xs += [rng.uniform(-1, 1, (n, m, m)) * 0.5 ** k for k in range(1, order + 1)]

    def build(g):
        x = g.placeholder()
        return x.batched_det() * x

    compare(impl, ref, build, xs, order)


@pytest.mark.parametrize("require_rotation", [False, True])
def test_svd_w_full_mode(impl, ref, require_rotation):
    """BatchSVDW with U and S READ by the graph: the full mode of the
    reference (libsanm/oprs/linalg.cpp:551-563, svd_w_taylor_fwd in
    tensor_svd.cpp:275-387) instead of the polar mode `fea` uses.
    y = (sum S) * U W^T - 2 W + U  exercises all three outputs."""
    rng = np.random.default_rng(77)
    n, order = 9, 8
    xs = make_F_coeffs(rng, n, order, amp=0.05)
    # distinct singular values (the U recurrence divides by S_i - S_j)
    xs[0] = xs[0] * np.array([1.0, 1.3, 1.7])[None, None, :]

    def build(g):
        x = g.placeholder()
        u, s, w = x.batched_svd_w(require_rotation)
        tr = s.reduce_sum(1)  # [N, 1]
        uwt = u.batched_matmul(w.batched_transpose())
        return g.linear_combine([(1.0, tr * uwt), (-2.0, w), (1.0, u)], 0.0)

    # (the U recurrence divides by S_i - S_j ~ 0.3: 5e-10 instead of 1e-10)
    compare(impl, ref, build, xs, order, tol=5e-10)
