"""SparseSolver: residual-defined parity (the reference's PARDISO is
a closed third-party binary; its tests pin A x == b and the Tikhonov normal
equations, tests/tensor.cpp:44-85)."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from sanm_b200.capi import SparseSolver


def build(b, A):
    A = sp.csr_matrix(A)
    A.sort_indices()
    n = A.shape[0]
    s = SparseSolver(b, n)
    rows = np.repeat(np.arange(n), np.diff(A.indptr))
    s.add_constraints(rows, A.indices, A.data)
    return s


def grid_matrix(nx, ny, nz, rng, nonsym=0.0, block=3):
    """block-structured FEA-like matrix on a 3-D grid (27-point coupling)"""
    idx = np.arange(nx * ny * nz).reshape(nx, ny, nz)
    rows, cols = [], []
    for dx in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dz in (-1, 0, 1):
                a = idx[max(0, -dx):nx - max(0, dx), max(0, -dy):ny - max(0, dy),
                        max(0, -dz):nz - max(0, dz)]
                b = idx[max(0, dx):nx - max(0, -dx), max(0, dy):ny - max(0, -dy),
                        max(0, dz):nz - max(0, -dz)]
                rows.append(a.ravel())
                cols.append(b.ravel())
    rows = np.concatenate(rows)
    cols = np.concatenate(cols)
    nv = nx * ny * nz
    G = sp.csr_matrix((np.ones(len(rows)), (rows, cols)), shape=(nv, nv))
    P = sp.kron(G, np.ones((block, block))).tocsr()
    P.data = rng.uniform(-1, 1, P.nnz)
    S = (P + P.T) * 0.5 + nonsym * sp.triu(P, 1)
    # diagonally dominant => well conditioned
    d = np.asarray(abs(S).sum(axis=1)).ravel()
    return (S + sp.diags(d * 0.6)).tocsr()


def test_reference_8x8(impl, ref):
    """Tensor.SparseSolver, tests/tensor.cpp:44-85 (random sparse 8x8 + l2)"""
    rng = np.random.default_rng(7)
    n = 8
    A = rng.uniform(-1, 1, (n, n)) * (rng.uniform(0, 1, (n, n)) < 0.5)
    A += np.diag(rng.uniform(1, 2, n) * 3)
    bvec = rng.uniform(-1, 1, n)
    for binding in (impl, ref):
        s = build(binding, A).prepare()
        x = s.solve(bvec)
        np.testing.assert_allclose(A @ x, bvec, atol=1e-12)
        np.testing.assert_allclose(s.apply(x), bvec, atol=1e-12)
        assert abs(s.coeff_l2() - np.linalg.norm(A)) < 1e-12
    xs = []
    for binding in (impl, ref):
        s = build(binding, A).prepare(0.3)
        x = s.solve(bvec)
        np.testing.assert_allclose((A.T @ A + 0.3 * np.eye(n)) @ x, A.T @ bvec,
                                   atol=1e-12)
        xs.append(x)
    np.testing.assert_allclose(xs[0], xs[1], rtol=1e-10)


def test_zero_diagonal_small(impl):
    """pivoting inside the root front: permuted identity-like system"""
    A = np.array([[0, 2, 0, 0], [1, 0, 0, 3], [0, 0, 0, 1], [0, 0, 4, 1.]])
    b = np.array([1., 2., 3., 4.])
    x = build(impl, A).prepare().solve(b)
    np.testing.assert_allclose(A @ x, b, atol=1e-13)


@pytest.mark.parametrize("shape,nonsym", [((5, 4, 3), 0.0), ((9, 8, 7), 0.3),
                                          ((16, 12, 10), 0.0),
                                          ((24, 16, 12), 0.2)])
@pytest.mark.parametrize("phased", [False, True])
def test_grid_blocks(impl, shape, nonsym, phased, monkeypatch):
    # default: one launch per level of the supernodal forest; SANM_LU_PHASED:
    # one persistent launch with grid barriers (read per solve() call)
    if phased:
        if shape == (24, 16, 12):
            pytest.skip("one size of the opt-in path is enough")
        monkeypatch.setenv("SANM_LU_PHASED", "1")
    else:
        monkeypatch.delenv("SANM_LU_PHASED", raising=False)
    rng = np.random.default_rng(11)
    A = grid_matrix(*shape, rng, nonsym=nonsym)
    n = A.shape[0]
    s = build(impl, A).prepare()
    st = s.stats()
    assert st["supernodes"] >= 1 and st["levels"] >= 1
    for _ in range(2):
        b = rng.uniform(-1, 1, n)
        x = s.solve(b)
        r = A @ x - b
        scale = abs(A).sum(axis=1).max() * np.abs(x).max() + np.abs(b).max()
        assert np.abs(r).max() <= 1e-12 * scale
        xr = spla.spsolve(A.tocsc(), b)
        assert np.abs(x - xr).max() <= 1e-10 * np.abs(xr).max()
        np.testing.assert_allclose(s.apply(x), A @ x, rtol=1e-13, atol=1e-13)


def test_row_order_and_small_drop(impl):
    """sparse_solver.cpp:284-305: rows must come in order, tiny entries vanish"""
    from sanm_b200.capi import SANMAssertionError
    s = SparseSolver(impl, 3)
    s.add_constraints([0, 0, 1], [0, 1, 1], [2., 1e-12, 3.])
    with pytest.raises(SANMAssertionError):
        s.add_constraints([0], [2], [1.])
    s.add_constraints([2], [2], [4.])
    s.prepare()
    np.testing.assert_allclose(s.solve([2., 3., 4.]), [1, 1, 1], atol=1e-14)
    assert abs(s.coeff_l2() - np.sqrt(4 + 9 + 16)) < 1e-14


@pytest.mark.parametrize("shape,nonsym", [((9, 8, 7), 0.3), ((16, 12, 10), 0.0),
                                          ((24, 16, 12), 0.2)])
def test_wide_front_path(impl, shape, nonsym, monkeypatch):
    """the superblock path of large fronts (128 pivots per step: shared-memory
    inverse of the pivot block, DMMA panels and rank-128 trailing update,
    csrc/sparse_lu_wide.h) and the solve variant that keeps a front's work
    vector in global memory, both forced onto small fronts through their
    thresholds; same residual / solution bars as the default path"""
    monkeypatch.setenv("SANM_LU_WIDE_MIN", "129")
    monkeypatch.setenv("SANM_LU_SOLVE_SMEM_NR", "200")
    rng = np.random.default_rng(13)
    A = grid_matrix(*shape, rng, nonsym=nonsym)
    n = A.shape[0]
    s = build(impl, A).prepare()
    for _ in range(2):
        b = rng.uniform(-1, 1, n)
        x = s.solve(b)
        r = A @ x - b
        scale = abs(A).sum(axis=1).max() * np.abs(x).max() + np.abs(b).max()
        assert np.abs(r).max() <= 1e-12 * scale
        xr = spla.spsolve(A.tocsc(), b)
        assert np.abs(x - xr).max() <= 1e-10 * np.abs(xr).max()


def test_solve_split_levels(impl, monkeypatch):
    """large mixed levels run their narrow fronts as a launch of their own
    (SANM_LU_SOLVE_SPLIT), in narrow blocks -- another, equally fixed summation
    tree for those fronts; forced onto a small system: same solution to
    rounding, bitwise repeatable, same residual bar"""
    rng = np.random.default_rng(31)
    A = grid_matrix(12, 10, 9, rng, nonsym=0.2)
    n = A.shape[0]
    b = rng.uniform(-1, 1, n)
    s = build(impl, A).prepare()
    monkeypatch.setenv("SANM_LU_SOLVE_SPLIT", "0")
    x0 = s.solve(b)
    monkeypatch.setenv("SANM_LU_SOLVE_SPLIT", "1")
    x1 = s.solve(b)
    x2 = s.solve(b)
    assert np.array_equal(x1, x2)
    assert np.abs(x1 - x0).max() <= 1e-12 * np.abs(x0).max()
    r = A @ x1 - b
    scale = abs(A).sum(axis=1).max() * np.abs(x1).max() + np.abs(b).max()
    assert np.abs(r).max() <= 1e-12 * scale


def test_adjacency_fast_path_same_analysis(impl, monkeypatch):
    """structurally symmetric patterns take their adjacency directly
    (sparse_symbolic.cpp sym_adj_fast); the analysis and the solution must be
    those of the general transposing build, and an unsymmetric pattern must
    fall back to it"""
    rng = np.random.default_rng(29)
    A = grid_matrix(7, 6, 5, rng, nonsym=0.3)
    n = A.shape[0]
    b = rng.uniform(-1, 1, n)
    res = []
    # the topology cache is keyed by the pattern and would hide the second analysis
    monkeypatch.setenv("SANM_NO_TOPO_CACHE", "1")
    for general in (False, True):
        if general:
            monkeypatch.setenv("SANM_ND_NO_FAST_ADJ", "1")
        else:
            monkeypatch.delenv("SANM_ND_NO_FAST_ADJ", raising=False)
        s = build(impl, A).prepare()
        res.append((s.stats(), s.solve(b)))
    assert res[0][0] == res[1][0]
    assert np.array_equal(res[0][1], res[1][1])
    # one-sided entries: the pattern is not symmetric any more
    B = A.tolil()
    B[0, n - 1] = 0.125
    B[3, n - 2] = -0.25
    B = B.tocsr()
    monkeypatch.delenv("SANM_ND_NO_FAST_ADJ", raising=False)
    x = build(impl, B).prepare().solve(b)
    xr = spla.spsolve(B.tocsc(), b)
    assert np.abs(x - xr).max() <= 1e-10 * np.abs(xr).max()


@pytest.mark.gpu
def test_large_system_gpu(cuda_impl, monkeypatch):
    """n = 140 k unknowns, nonsymmetric values: fronts of several thousand rows
    (wide path by default), top levels solved with global work vectors;
    residual bar of SURVEY 8c.  GPU only (5e11 flops)."""
    monkeypatch.setenv("SANM_LU_SOLVE_SMEM_NR", "3000")
    rng = np.random.default_rng(17)
    A = grid_matrix(36, 36, 36, rng, nonsym=0.2)
    n = A.shape[0]
    s = build(cuda_impl, A).prepare()
    st = s.stats()
    assert st["flops"] > 1e11
    for _ in range(2):
        b = rng.uniform(-1, 1, n)
        x = s.solve(b)
        r = A @ x - b
        scale = abs(A).sum(axis=1).max() * np.abs(x).max() + np.abs(b).max()
        assert np.abs(r).max() <= 1e-12 * scale


@pytest.mark.gpu
def test_wide_gemm_fast_tiles_bitwise_gpu(cuda_impl, monkeypatch):
    """interior tiles of the rank-128 trailing update take a leaner code path
    (sparse_lu_wide.h wide_gemm_tile_fast: accumulators start from -C, +A in the
    MMAs); it must reproduce the generic tile bit for bit"""
    rng = np.random.default_rng(23)
    A = grid_matrix(30, 30, 30, rng, nonsym=0.2)
    n = A.shape[0]
    b = rng.uniform(-1, 1, n)
    xs = []
    for cfg, slow in (("big", False), ("big", True), ("persist", False),
                      ("persist+lookahead", False)):
        monkeypatch.setenv("SANM_LU_WIDE_CFG", cfg.split("+")[0])
        if "lookahead" in cfg:
            monkeypatch.setenv("SANM_LU_LOOKAHEAD", "1")
        else:
            monkeypatch.delenv("SANM_LU_LOOKAHEAD", raising=False)
        if slow:
            monkeypatch.setenv("SANM_LU_WIDE_SLOW", "1")
        else:
            monkeypatch.delenv("SANM_LU_WIDE_SLOW", raising=False)
        xs.append(build(cuda_impl, A).prepare().solve(b))
    # the persistent kernel (next tile's C prefetched into shared memory), and
    # with the pivot-block inverse of the next step on a side stream
    assert np.array_equal(xs[0], xs[2])
    assert np.array_equal(xs[0], xs[3])
    assert np.array_equal(xs[0], xs[1])
    r = A @ xs[0] - b
    scale = abs(A).sum(axis=1).max() * np.abs(xs[0]).max() + np.abs(b).max()
    assert np.abs(r).max() <= 1e-12 * scale

