"""Independent numerical pins for the third-party pieces the oracle restates from their published algorithms
(the libraries are absent from this image): Eigen::JacobiSVD<Matrix3f>, pcl::TransformationFromCorrespondences
(Solver/SolverSE3.cpp:161-179), Eigen::LLT<Matrix3d>::solve inside errorFunction2 (:216-280) and Eigen::umeyama
(Solver/Ransac.cpp:231).  The independent implementation is numpy float64 (LAPACK), written from the reference's
call sites and the textbook definitions, not from the oracle's code.  These pin the VALUES to a stated tolerance;
the bit patterns of the libraries themselves stay unpinned."""
import ctypes as C

import numpy as np
import pytest


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def L(oracle):
    return oracle.lib()


def _svd_cases(rng):
    A = [rng.normal(0, 1, (6000, 3, 3))]
    A.append(rng.normal(0, 1, (1500, 3, 3)) * 10.0 ** rng.uniform(-6, 6, (1500, 1, 1)))  # scales 1e-6 .. 1e6
    # rank 2, rank 1, rank 0
    u, v, w = (rng.normal(0, 1, (600, 3)) for _ in range(3))
    A.append(u[:, :, None] * v[:, None, :] + w[:, :, None] * u[:, None, :])
    A.append(u[:, :, None] * v[:, None, :])
    A.append(np.zeros((3, 3, 3)))
    # repeated singular values: rotations, reflections, scaled identity, diagonal with ties
    q, _ = np.linalg.qr(rng.normal(0, 1, (600, 3, 3)))
    A.append(q)
    A.append(q * np.array([1.0, 1.0, -1.0]))  # det < 0
    A.append(np.eye(3)[None] * rng.uniform(0.1, 10, (50, 1, 1)))
    d = np.zeros((300, 3, 3))
    dv = rng.integers(1, 4, (300, 3)).astype(np.float64)
    d[:, 0, 0], d[:, 1, 1], d[:, 2, 2] = dv[:, 0], dv[:, 1], -dv[:, 2]
    A.append(d)
    # covariance-like matrices of noisy rigid motions (what TFC / umeyama feed the SVD), incl. near-planar sets
    for flat in (1.0, 1e-3):
        P = rng.normal(0, 1, (300, 40, 3)) * np.array([1.0, 1.0, flat])
        A.append(np.einsum("nki,nkj->nij", P @ q[:300], P) / 40)
    return np.concatenate(A).astype(np.float32)


def test_jacobi_svd3_against_lapack(L):
    rng = np.random.default_rng(2024)
    A = _svd_cases(rng)
    n = len(A)
    assert n >= 10000
    U, S, V = np.zeros((n, 3, 3), np.float32), np.zeros((n, 3), np.float32), np.zeros((n, 3, 3), np.float32)
    L.orc_jacobi_svd3_batch(n, _p(A), _p(U), _p(S), _p(V))
    A64, U64, S64, V64 = A.astype(np.float64), U.astype(np.float64), S.astype(np.float64), V.astype(np.float64)
    ref = np.linalg.svd(A64, compute_uv=False)
    scale = np.maximum(ref[:, :1], 1e-300)
    # singular values: non-negative, descending, equal to LAPACK's to 1e-5 of the largest
    assert (S >= 0).all() and (S[:, 0] >= S[:, 1]).all() and (S[:, 1] >= S[:, 2]).all()
    assert (np.abs(S64 - ref) / scale).max() < 1e-5
    # A = U diag(S) V^T, U and V orthogonal (the zero matrix returns identities)
    rec = np.einsum("nij,nj,nkj->nik", U64, S64, V64)
    assert (np.abs(rec - A64).reshape(n, -1).max(1) / scale[:, 0]).max() < 1e-5
    eye = np.eye(3)
    assert np.abs(np.einsum("nji,njk->nik", U64, U64) - eye).max() < 1e-5
    assert np.abs(np.einsum("nji,njk->nik", V64, V64) - eye).max() < 1e-5
    # reflection inputs keep det(U) det(V) = sign(det A): what the callers' sign fix relies on
    detA = np.linalg.det(A64)
    big = np.abs(detA) > 1e-3 * scale[:, 0] ** 3
    assert (np.sign(np.linalg.det(U64[big]) * np.linalg.det(V64[big])) == np.sign(detA[big])).all()


def _kabsch(p1, p2, w):
    """Weighted least-squares rigid motion p2 ~ R p1 + t (Kabsch / Umeyama without scale), float64."""
    w = w / w.sum()
    m1, m2 = (w[:, None] * p1).sum(0), (w[:, None] * p2).sum(0)
    H = (w[:, None, None] * (p2 - m2)[:, :, None] * (p1 - m1)[:, None, :]).sum(0)
    U, S, Vt = np.linalg.svd(H)
    D = np.diag([1.0, 1.0, np.sign(np.linalg.det(U) * np.linalg.det(Vt))])
    R = U @ D @ Vt
    return R, m2 - R @ m1, S


def _rigid_sets(rng, n_sets):
    """Point sets the RANSAC actually fits: depth 0.5 - 4.5 m, a rigid motion of <= 20 degrees / 0.5 m, sensor noise."""
    import synth

    out = []
    for k in range(n_sets):
        n = [4, 4, 5, 8, 30, 200, 800][k % 7]
        z = rng.uniform(0.5, 4.5, n)
        p1 = np.stack([rng.uniform(-1, 1, n) * z, rng.uniform(-0.8, 0.8, n) * z, z], 1)
        if k % 5 == 0:  # near-planar scene (a wall): the third singular value of the covariance is ~0
            p1[:, 2] = 2.0 + 0.001 * rng.normal(0, 1, n) + 0.1 * p1[:, 0]
        R = synth.rot_xyz(*(rng.uniform(-1, 1, 3) * np.deg2rad(20)))
        t = rng.uniform(-0.5, 0.5, 3)
        p2 = p1 @ R.T + t + rng.normal(0, 1, (n, 3)) * 0.0012 * p1[:, 2:3] ** 2
        out.append((p1.astype(np.float32), p2.astype(np.float32)))
    return out


def test_tfc_fit_against_weighted_kabsch(L, oracle):
    rng = np.random.default_rng(7)
    worst_R = worst_t = 0.0
    checked = 0
    for p1, p2 in _rigid_sets(rng, 700):
        n = len(p1)
        m = np.zeros(n, oracle.DM_DTYPE)
        m["queryIdx"] = m["trainIdx"] = np.arange(n)
        T = np.zeros(16, np.float32)
        L.orc_tfc_fit(_p(p1), _p(p2), _p(m), n, _p(T))
        T = T.reshape(4, 4).astype(np.float64)
        a, b = p1.astype(np.float64), p2.astype(np.float64)
        R, t, S = _kabsch(a, b, 1.0 / (a[:, 2] * b[:, 2]))  # TFC weights of SolverSE3.cpp:172-176
        if S[1] < 1e-3 * S[0]:
            continue  # collinear sample: the rotation about the line is undetermined for every solver
        checked += 1
        # conditioning of the rotation is set by the second singular value; FP32 data -> scale the bound with it
        tol = 1e-4 * max(1.0, 0.05 * S[0] / S[1])
        worst_R = max(worst_R, np.abs(T[:3, :3] - R).max() / tol)
        worst_t = max(worst_t, np.abs(T[:3, 3] - t).max() / (tol * max(1.0, np.abs(a).max())))
        assert np.abs(T[3] - [0, 0, 0, 1]).max() == 0
        assert abs(np.linalg.det(T[:3, :3]) - 1.0) < 1e-5
    assert checked > 600 and worst_R < 1.0 and worst_t < 1.0, (checked, worst_R, worst_t)


def test_umeyama_against_kabsch(L):
    rng = np.random.default_rng(8)
    checked = 0
    for p1, p2 in _rigid_sets(rng, 700):
        n = len(p1)
        T = np.zeros(16, np.float32)
        L.orc_umeyama3(_p(p1), _p(p2), n, _p(T))  # Eigen::umeyama(src, dst, false): dst ~ R src + t
        T = T.reshape(4, 4).astype(np.float64)
        a, b = p1.astype(np.float64), p2.astype(np.float64)
        R, t, S = _kabsch(a, b, np.ones(n))
        if S[1] < 1e-3 * S[0]:
            continue
        checked += 1
        tol = 1e-4 * max(1.0, 0.05 * S[0] / S[1])
        assert np.abs(T[:3, :3] - R).max() < tol, (n, S)
        assert np.abs(T[:3, 3] - t).max() < tol * max(1.0, np.abs(a).max()), (n, S)
    assert checked > 600


def test_mahalanobis_llt_against_numpy_solve(L, oracle):
    """errorFunction2: d^T (R^T S1 R + S2)^-1 d with S = diag(rcx z, rcy z, C) (SolverSE3.cpp:246-274)."""
    import synth

    rng = np.random.default_rng(9)
    n = 20000
    cov = (0.01 * 2.0 * 2.0) ** 2
    rcx = (3 * np.tan(58.0 / 180.0 * np.pi / 640)) ** 2
    rcy = (3 * np.tan(45.0 / 180.0 * np.pi / 480)) ** 2
    z = rng.uniform(0.4, 5.0, n)
    a = np.stack([rng.uniform(-1, 1, n) * z, rng.uniform(-0.8, 0.8, n) * z, z], 1).astype(np.float32)
    worst = 0.0
    for trial in range(4):
        R = synth.rot_xyz(*(rng.uniform(-1, 1, 3) * np.deg2rad(25)))
        T = np.eye(4)
        T[:3, :3], T[:3, 3] = R, rng.uniform(-0.4, 0.4, 3)
        T = T.astype(np.float32).astype(np.float64)  # the solver casts a float matrix to double (:186)
        b = (a.astype(np.float64) @ T[:3, :3].T + T[:3, 3] + rng.normal(0, 1, (n, 3)) * 10.0 ** rng.uniform(-4, -1, (n, 1))).astype(np.float32)
        out, full = np.zeros(n), np.zeros(n, np.int32)
        L.orc_mahal2_batch(n, _p(a), _p(b), _p(np.ascontiguousarray(T)), C.c_double(cov), _p(out), _p(full))
        a64, b64 = a.astype(np.float64), b.astype(np.float64)
        d = a64 @ T[:3, :3].T + T[:3, 3] - b64
        short = (d * d).sum(1) > 2.0 * (2 * max(rcx, cov))  # the error shortcut (:237-243)
        assert ((full == 0) == short).all()
        assert (out[short] == np.finfo(np.float64).max).all()
        idx = np.nonzero(~short)[0]
        assert len(idx) > 2000
        Rm = T[:3, :3]
        S1 = np.zeros((len(idx), 3, 3))
        S2 = np.zeros((len(idx), 3, 3))
        S1[:, 0, 0], S1[:, 1, 1], S1[:, 2, 2] = rcx * a64[idx, 2], rcy * a64[idx, 2], cov
        S2[:, 0, 0], S2[:, 1, 1], S2[:, 2, 2] = rcx * b64[idx, 2], rcy * b64[idx, 2], cov
        Sc = np.einsum("ji,njk,kl->nil", Rm, S1, Rm) + S2  # R^T S1 R + S2 (:262,271)
        ref = np.einsum("ni,ni->n", d[idx], np.linalg.solve(Sc, d[idx][:, :, None])[:, :, 0])
        worst = max(worst, (np.abs(out[idx] - ref) / np.maximum(ref, 1e-300)).max())
    assert worst < 1e-10, worst
