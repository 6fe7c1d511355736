"""ORACLE — TEST INFRASTRUCTURE ONLY (imported by tests/ only; never by the product).

numpy/scipy restatement of one time step of the reference's Newton optimizer for a single mesh with the
mixed-stretch variable and no contact, on top of the oracle's MixedStretch (oracle/mfem_oracle.cpp):

  NewtonOptimizer<DIM>::step / update_system / substep ... src/optimizers/newton_optimizer.cpp:14-139
  linesearch_backtracking(state, alpha, f, c=0, p=0.5) .... src/linesearch.h:80-129
  Displacement<DIM>::energy / gradient / reset / post_solve src/variables/displacement.cpp:21-96
  BDF<1> .................................................. src/time_integrators/BDF.cpp:7-27
  SystemMatrixPD::pre_solve / post_solve .................. src/linear_solvers/linear_system.h:18-44
  consistent mass matrix (assembled over elements) ........ deps/Bartels/src/linear_tet_mass_matrix.cpp:8-43,
                                                            linear_tri2d_mass_matrix.cpp:8-36,
                                                            shells: src/mesh/tri_mesh.cpp:43-69
The linear system is solved directly (scipy SuperLU) unless cg=(precond, tol, max_iters) is given, in which case the
Eigen ConjugateGradient restatement of oracle/pcg.py runs from the reference's initial guess
(eigen_iterative_solver.h:44-57; the additive-CCD scale is 1 without contact).  PARITY UNPINNED: the reference ships
no Newton-history fixtures; this follows the control flow line by line.
"""
import time

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from oracle import pcg as opcg


def mass_matrix(V, T, vols, density):
    """d*nV square consistent mass matrix: element blocks mass*(c0 on the diagonal, c1 off it) (x) I_d.
    density: scalar or one value per element = the density of the element's OBJECT (each object's matrix is scaled by its own
    material, tet_mesh.cpp:36-40 / tri2d_mesh.cpp:36-41 / tri_mesh.cpp:67-68, and Meshes stacks them block-diagonally,
    meshes.cpp:100-108)."""
    nV, d = V.shape
    npe = T.shape[1]
    # tets 1/10, 1/20 ; triangles (2D and shells: TriMesh::mass_matrix, src/mesh/tri_mesh.cpp:43-69) 1/6, 1/12
    c0, c1 = (0.1, 0.05) if npe == 4 else (1.0 / 6.0, 1.0 / 12.0)
    mass = density * np.asarray(vols)
    rows, cols, vals = [], [], []
    for j in range(npe):
        for k in range(npe):
            w = mass * (c0 if j == k else c1)
            for l in range(d):
                rows.append(d * T[:, j] + l); cols.append(d * T[:, k] + l); vals.append(w)
    M = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(d * nV, d * nV))
    return M.tocsr()


class OracleNewton:
    def __init__(self, var, V, T, free_map, density=1000.0, h=0.0333, force=None, x0=None, v0=None):
        self.var = var
        self.d = V.shape[1]
        self.nV = V.shape[0]
        self.h = float(h)
        fm = np.asarray(free_map)
        free = np.nonzero(fm >= 0)[0]
        order = np.argsort(fm[free])
        self.free_v = free[order]
        self.free_dofs = (self.d * self.free_v[:, None] + np.arange(self.d)[None, :]).reshape(-1)
        vols = var.get("vols")
        self.M = mass_matrix(V, np.asarray(T), vols, density)
        self.PM = self.M[self.free_dofs, :]
        self.PMP = self.PM[:, self.free_dofs].tocsr()
        f = np.array([0.0, -9.8, 0.0])[: self.d] if force is None else np.asarray(force, dtype=float)[: self.d]
        self.fext = np.tile(f, self.nV)
        self.x = (np.asarray(V, dtype=float).reshape(-1) if x0 is None else np.asarray(x0, dtype=float)).copy()
        self.x_prev = self.x.copy()
        self.v_prev = np.zeros_like(self.x) if v0 is None else np.asarray(v0, dtype=float).copy()
        # wall-clock seconds under the optimizer's Timer keys (newton_optimizer.cpp:122,140,162,164,86,88); used by
        # bench.py's CPU baseline of the ms-per-Newton-iteration leg
        self.timers = {"update": 0.0, "rhs": 0.0, "linsolve": 0.0, "solve": 0.0, "LS": 0.0}

    def x_tilde(self):
        return self.x_prev + self.h * self.v_prev

    def _disp_energy(self, x):
        diff = x - self.x_tilde() - self.h ** 2 * self.fext
        return 0.5 * float(diff @ (self.M @ diff))

    def step(self, outer_steps=5, ls_iters=20, newton_tol=1e-10, cg=None, stretch=None):
        """stretch: an oracle.stretch.OracleStretch on the same mesh => log["fem_gradient_norm"] holds what
        apps/convergence.cpp:21-51 records at the top of every iteration (Optimizer::callback, newton_optimizer.cpp:37):
        |x.rhs() + Stretch.rhs()|_2 of the non-mixed system at the current mixed iterate"""
        var, h = self.var, self.h
        h2 = h * h
        log = {k: [] for k in ("energy", "energy_res", "decrement", "alpha", "cg_iters", "ls_iters", "fem_gradient_norm")}
        i, E_prev = 0, 0.0
        tm = self.timers
        while True:
            if stretch is not None:
                stretch.update(self.x, h)
                d0 = self.x - self.x_tilde() - h2 * self.fext
                log["fem_gradient_norm"].append(float(np.linalg.norm(-(self.PM @ d0) + stretch.rhs())))
            t0 = time.perf_counter()
            var.update(self.x, h)
            t1 = time.perf_counter(); tm["update"] += t1 - t0
            outer, inner, vals = var.csr()
            n = outer.size - 1
            A = sp.csr_matrix((vals, inner, outer), shape=(n, n))
            lhs = (self.PMP + A).tocsr()
            diff = self.x - self.x_tilde() - h2 * self.fext
            rhs = -(self.PM @ diff) + var.rhs()
            t2 = time.perf_counter(); tm["rhs"] += t2 - t1
            if cg is None:
                dx = spla.spsolve(lhs.tocsc(), rhs)
                cg_it = 0
            else:
                guess = -diff[self.free_dofs]
                dx, cg_it, _ = opcg.conjugate_gradient(lhs, rhs, guess, opcg.make_precond(lhs, cg[0], self.d), cg[1], cg[2])
            t3 = time.perf_counter(); tm["linsolve"] += t3 - t2
            var.solve(dx)
            t4 = time.perf_counter(); tm["solve"] += t4 - t3
            ds = var.get("ds")
            decrement = max(np.abs(dx).max(), np.abs(ds).max())
            s0 = var.get("s")

            def energy(a):
                x0 = self.x.copy()
                x0[self.free_dofs] += a * dx
                si = s0 + a * ds
                return self._disp_energy(x0) + h2 * var.energy(x0, si) + var.constraint_value(x0, si)

            E = energy(0.0)
            res = abs((E - E_prev) / (E + 1e-6))
            E_prev = E
            alpha, it = 1.0, 0
            while it < ls_iters:
                if energy(alpha) < E:
                    break
                alpha *= 0.5
                it += 1
            tm["LS"] += time.perf_counter() - t4
            if it < ls_iters:
                self.x[self.free_dofs] += alpha * dx
                var.set("s", s0 + alpha * ds)
            for k, v in (("energy", E), ("energy_res", res), ("decrement", decrement), ("alpha", alpha), ("cg_iters", cg_it), ("ls_iters", it)):
                log[k].append(v)
            i += 1
            if not (i < outer_steps and decrement > newton_tol and res > 1e-12):
                break
        # post_solve: BDF1 update, la_ = 0
        v = (self.x - self.x_prev) / h
        self.x_prev = self.x.copy()
        self.v_prev = v
        var.set("la", np.zeros_like(var.get("la")))
        out = {k: np.array(v) for k, v in log.items()}
        out["n_iters"] = i
        return out
