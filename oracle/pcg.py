"""ORACLE — TEST INFRASTRUCTURE ONLY.

numpy/scipy restatement of Eigen 3.4 `ConjugateGradient` (Eigen/src/IterativeLinearSolvers/ConjugateGradient.h,
internal::conjugate_gradient) as used by EigenIterativeSolver (src/linear_solvers/eigen_iterative_solver.h:27-65)
with DiagonalPreconditioner (linear_solver_factory.cpp:108-117) or BlockJacobiPreconditioner
(src/linear_solvers/preconditioners/block_jacobi.h:45-60).  The same recurrence is hand-written in the reference's
src/linear_solvers/pcg.h:11-80.  Eigen is absent from this container, so this follows the published algorithm.
"""
import numpy as np


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
