"""ORACLE — TEST INFRASTRUCTURE ONLY (imported by tests/ only; never by the product).

numpy/scipy restatement of the reference's two saddle-point formulations on top of the oracle's MixedStretch
(oracle/mfem_oracle.cpp) and the Displacement / BDF1 state of oracle/newton.py:

  SystemMatrixIndefinite::pre_solve / post_solve ... src/linear_solvers/linear_system.h:64-118
  BlockMatrix product (scaleAndAddTo) .............. src/linear_solvers/block_matrix.h:74-125  (same call order)
  DualCondensedSystem::pre_solve / post_solve ...... src/linear_solvers/linear_system.h:120-239 (explicit sparse lhs)
  MINRES ........................................... Eigen 3.4.0 unsupported/Eigen/src/IterativeSolvers/MINRES.h (NOT under
                                                     /root/reference: Paige-Saunders recurrence in its Lanczos form)

PARITY UNPINNED: the reference ships no fixtures for these systems; the explicit matrices built here
are cross-checked against each other (KKT elimination == condensed system) in tests/test_oracle_saddle.py.
"""
import numpy as np
import scipy.sparse as sp


class OracleSaddle:
    def __init__(self, var, newton):
        """var: orc.OracleMixedStretch after update(newton.x, h); newton: oracle.newton.OracleNewton."""
        self.var, self.nt = var, newton
        self.n = newton.free_dofs.size
        self.m = var.get("s").size

    # ---- Displacement::gradient (displacement.cpp:59-70)
    def grad_x(self):
        nt = self.nt
        diff = nt.x - nt.x_tilde() - nt.h ** 2 * nt.fext
        return nt.PM @ diff

    # ---- SystemMatrixIndefinite
    def kkt_rhs(self):
        return np.concatenate([-self.grad_x(), -self.var.get("grad"), -self.var.get("grad_la")])

    def kkt_product(self, v):
        """block_matrix.h:84-124, in the reference's order of calls"""
        var, n, m = self.var, self.n, self.m
        dx, ds, la = v[:n], v[n:n + m], v[n + m:]
        out_x = self.nt.PMP @ dx                                           # x_->product_hessian
        out_x = var.product("jacobian_x", la, out_x, transposed=True)      # -= PJW dSdF la
        out_s = var.product("hessian", ds, np.zeros(m))
        out_s = var.product("jacobian_mixed", la, out_s)
        out_l = var.product("jacobian_x", dx, np.zeros(m), transposed=False)
        out_l = var.product("jacobian_mixed", ds, out_l)
        return np.concatenate([out_x, out_s, out_l])

    def kkt_matrix(self):
        """explicit KKT matrix (small meshes): columns by applying the product to unit vectors"""
        nk = self.n + 2 * self.m
        K = np.zeros((nk, nk))
        e = np.zeros(nk)
        for j in range(nk):
            e[j] = 1.0
            K[:, j] = self.kkt_product(e)
            e[j] = 0.0
        return K

    # ---- DualCondensedSystem
    def jacobian_x(self):
        """B = -PJW blockdiag(dSdF) (mixed_stretch.cpp:385-390), dense n x m, column by column through the product"""
        B = np.zeros((self.n, self.m))
        e = np.zeros(self.m)
        for j in range(self.m):
            e[j] = 1.0
            B[:, j] = self.var.product("jacobian_x", e, np.zeros(self.n), transposed=True)
            e[j] = 0.0
        return B

    def _blockdiag(self, which):
        """explicit block-diagonal operator (hessian_inv(), jacobian_mixed(): mixed_stretch.cpp:373-404) from its product:
        the product does not mix elements, so applying it to the k-th unit vector of every element at once gives column k
        of every block"""
        N = 6 if self.nt.d == 3 else 3
        D = np.zeros((self.m, self.m))
        el = np.arange(self.m // N)
        for k in range(N):
            e = np.zeros(self.m)
            e[k::N] = 1.0
            out = self.var.product(which, e, np.zeros(self.m)).reshape(-1, N)
            for r in range(N):
                D[N * el + r, N * el + k] = out[:, r]
        return D

    def dual_system(self):
        """lhs = G^T G + C*C*H, rhs = B^T Minv (-gx) - C H gradient_mixed + gradient_dual (linear_system.h:127-169)"""
        lumped = np.asarray(self.nt.PMP @ np.ones(self.n)).reshape(-1)
        Minv = 1.0 / lumped
        B = self.jacobian_x()
        Cm = self._blockdiag("jacobian_mixed")
        H = self._blockdiag("hessian_inv")
        G = (Minv ** 0.5)[:, None] * B
        lhs = G.T @ G + Cm @ Cm @ H
        gx = self.grad_x()
        rhs = B.T @ (Minv * -gx)
        rhs = rhs - Cm @ (H @ self.var.get("grad"))
        rhs = rhs + self.var.get("grad_la")
        return lhs, rhs, dict(B=B, C=Cm, H=H, Minv=Minv, gx=gx)

    def dual_post_solve(self, la, parts):
        """linear_system.h:178-209: returns (dx, ds)"""
        var = self.var
        delta = var.product("jacobian_x", la, np.zeros(self.n), transposed=True)
        dx = parts["Minv"] * (-parts["gx"] - delta)
        tmp = var.product("jacobian_mixed", la, np.zeros(self.m))
        tmp = -var.get("grad") - tmp
        ds = var.product("hessian_inv", tmp, np.zeros(self.m))
        return dx, ds


def minres(A, b, x0, tol, max_iters):
    """Eigen::internal::minres with the identity preconditioner; A: callable or matrix. Returns (x, iters, rel. residual)."""
    mv = A if callable(A) else (lambda v: A @ v)
    x = np.array(x0, dtype=float)
    rhs2 = float(b @ b)
    if rhs2 == 0.0:
        return np.zeros_like(x), 0, 0.0
    thr2 = tol * tol * rhs2
    v = np.zeros_like(x)
    v_new = b - mv(x)
    r2 = float(v_new @ v_new)
    w_new = v_new.copy()
    beta_new = np.sqrt(float(v_new @ w_new))
    beta_one = beta_new
    c = c_old = 1.0
    s = s_old = 0.0
    p_old = np.zeros_like(x)
    p = p_old.copy()
    eta = 1.0
    n = 0
    if r2 < thr2:
        return x, 0, np.sqrt(r2 / rhs2)
    while n < max_iters:
        beta = beta_new
        v_old = v
        v_new = v_new / beta_new
        w_new = w_new / beta_new
        v = v_new
        w = w_new
        v_new = mv(w) - beta * v_old
        alpha = float(v_new @ w)
        v_new = v_new - alpha * v
        w_new = v_new.copy()
        beta_new = np.sqrt(float(v_new @ w_new))
        r2h = s * alpha + c * c_old * beta
        r3 = s_old * beta
        r1_hat = c * alpha - c_old * s * beta
        r1 = np.sqrt(r1_hat ** 2 + beta_new ** 2)
        c_old, s_old = c, s
        c, s = r1_hat / r1, beta_new / r1
        p_oold, p_old = p_old, p
        p = (w - r2h * p_old - r3 * p_oold) / r1
        x = x + beta_one * c * eta * p
        r2 *= s * s
        if r2 < thr2:
            break
        eta = -s * eta
        n += 1
    return x, n, np.sqrt(r2 / rhs2)
