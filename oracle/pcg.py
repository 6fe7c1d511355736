"""ORACLE — TEST INFRASTRUCTURE ONLY.

numpy/scipy restatement of Eigen 3.4 `ConjugateGradient` (Eigen/src/IterativeLinearSolvers/ConjugateGradient.h,
internal::conjugate_gradient) as used by EigenIterativeSolver (src/linear_solvers/eigen_iterative_solver.h:27-65)
with DiagonalPreconditioner (linear_solver_factory.cpp:108-117) or BlockJacobiPreconditioner
(src/linear_solvers/preconditioners/block_jacobi.h:45-60).  The same recurrence is hand-written in the reference's
src/linear_solvers/pcg.h:11-80.  Eigen is absent from this container, so this follows the published algorithm.
"""
import numpy as np


def chebyshev_coefficients(lmax, ratio, degree):
    """(theta, [(c1, c2) for steps 1..degree-1]) of the Chebyshev iteration on [lmax / ratio, lmax]:
    d_0 = Dinv r / theta ; d_j = c1_j d_{j-1} + c2_j Dinv (r - A z_{j-1}) ; z_j = z_{j-1} + d_j"""
    lmin = lmax / ratio
    theta, delta = 0.5 * (lmax + lmin), 0.5 * (lmax - lmin)
    sigma = theta / delta
    rho = 1.0 / sigma
    coef = []
    for _ in range(degree - 1):
        rho_new = 1.0 / (2.0 * sigma - rho)
        coef.append((rho_new * rho, 2.0 * rho_new / delta))
        rho = rho_new
    return theta, coef


def chebyshev_lmax(A, dinv, npe, iters=12, safety=1.2):
    """largest eigenvalue of Dinv A as mfb_pcg_dev estimates it for precond 3: power iteration from an integer-hash
    vector, times a safety factor, capped by the rigorous bound lambda_max <= vertices per element"""
    i = np.arange(A.shape[0], dtype=np.uint64)
    v = ((i * np.uint64(2654435761)) & np.uint64(0xFFFFFFFF)).astype(np.float64) / 4294967296.0 - 0.5
    v /= np.sqrt(float(v @ v))
    lam = 1.0
    for _ in range(iters):
        w = dinv(A @ v)
        lam = np.sqrt(float(w @ w))
        v = w / lam
    return min(safety * lam, float(npe))


def make_chebyshev(A, dim, b, degree=6, ratio=120.0, npe=None):
    """precond 3 of mfb_pcg_dev: a fixed polynomial of degree-1 in (block-Jacobi)^-1 A applied to (block-Jacobi)^-1 r.
    Not in the reference (its non-Jacobi preconditioners are factorisations / AMG, src/factories/linear_solver_factory.cpp:
    119-173); symmetric positive definite for lmax >= lambda_max, hence a valid CG preconditioner."""
    dinv = make_precond(A, 2, dim)
    lmax = chebyshev_lmax(A, dinv, dim + 1 if npe is None else npe)
    theta, coef = chebyshev_coefficients(lmax, ratio, degree)

    def apply(r):
        d = dinv(r) / theta
        z = d.copy()
        res = r.copy()
        for c1, c2 in coef:
            res = res - A @ d
            d = c1 * d + c2 * dinv(res)
            z = z + d
        return z
    return apply


def make_precond(A, kind, dim):
    """kind: 0 identity, 1 diagonal (zero diagonal -> 1), 2 dim x dim block-Jacobi inverse"""
    n = A.shape[0]
    if kind == 0:
        return lambda r: r.copy()
    if kind == 1:
        d = A.diagonal().copy()
        inv = np.where(d != 0.0, 1.0 / np.where(d != 0.0, d, 1.0), 1.0)
        return lambda r: inv * r
    blocks = np.stack([np.linalg.inv(A[dim * i:dim * i + dim, dim * i:dim * i + dim].toarray()) for i in range(n // dim)])
    return lambda r: np.einsum("ijk,ik->ij", blocks, r.reshape(-1, dim)).reshape(-1)


def conjugate_gradient(A, b, x0, precond, tol, max_iters):
    """returns (x, iterations, tol_error) with Eigen's conventions"""
    x = x0.copy()
    residual = b - A @ x
    rhs_norm2 = float(b @ b)
    if rhs_norm2 == 0.0:
        return np.zeros_like(x), 0, 0.0
    threshold = max(tol * tol * rhs_norm2, np.finfo(np.float64).tiny)
    res2 = float(residual @ residual)
    if res2 < threshold:
        return x, 0, np.sqrt(res2 / rhs_norm2)
    p = precond(residual)
    abs_new = float(residual @ p)
    i = 0
    while i < max_iters:
        tmp = A @ p
        alpha = abs_new / float(p @ tmp)
        x += alpha * p
        residual -= alpha * tmp
        res2 = float(residual @ residual)
        if res2 < threshold:
            break
        z = precond(residual)
        abs_old = abs_new
        abs_new = float(residual @ z)
        p = z + (abs_new / abs_old) * p
        i += 1
    return x, i, np.sqrt(res2 / rhs_norm2)
