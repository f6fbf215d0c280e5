"""CPU feasibility study (oracle matrices, SciPy): geometric multigrid V-cycle as PCG preconditioner for the
Q2xQ1 RM-shell plate, against block-Jacobi PCG.  Development aid, not part of the product path."""
import sys, os, time
import numpy as np, scipy.sparse as sp, scipy.sparse.linalg as spla
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.rm_shell_ref import RMShellOracle, plate_mesh, basis_q1, basis_q2


def level(n, L=10.0, h=0.1):
    nodes, quads = plate_mesh(n, n, L, L)
    o = RMShellOracle(nodes, quads, False, False, 6)
    nn = nodes.shape[0]
    x, y = nodes[:, 0], nodes[:, 1]
    o.set_inputs(thickness=h * (1 + 0.5 * np.sin(2 * np.pi * x / 10) * np.cos(2 * np.pi * y / 10)), E=np.full(nn, 1e8),
                 nu=np.full(nn, 0.3), density=np.full(nn, 10.0), F_solid=np.tile([0, 0, 5.0], nn))
    o.set_strong_bc(lambda X: X[0] < 3e-16)
    A = o.apply_bc_matrix(o.assemble_dRdu()).tocsr()
    b = o.assemble_load(); b[o.bc_dofs] = 0
    return o, A, b


def prolongation(of, oc, n_c, L=10.0):
    """fine block node -> coarse cell (structured nested plates) -> coarse basis weights"""
    tf, tc = of.topo, oc.topo
    xf = tf.block_coords()
    hc = L / n_c
    rows, cols, vals = [], [], []
    for I in range(tf.nb):
        x, y = xf[I, 0], xf[I, 1]
        ci = min(int(x / hc + 1e-9), n_c - 1); cj = min(int(y / hc + 1e-9), n_c - 1)
        cell = ci * n_c + cj                       # plate_mesh numbers cells x-major
        X0 = x / hc - ci; X1 = y / hc - cj
        # tensor vertex order of the oracle: local X0 along x?  verify with vertex coordinates
        cvx = tc.x[tc.cv[cell]]
        # map (X0, X1) so that vertex 0 -> (0,0), 1 -> (1,0), 2 -> (0,1)
        e0 = cvx[1] - cvx[0]; e1 = cvx[2] - cvx[0]
        d = np.array([x, y, 0.0]) - cvx[0]
        X0 = d @ e0 / (e0 @ e0); X1 = d @ e1 / (e1 @ e1)
        if I < tf.nq2:
            w, _ = basis_q2(X0, X1); blk = tc.cell_blocks[cell, :9]
        else:
            w, _ = basis_q1(X0, X1); blk = tc.cell_blocks[cell, 9:]
        for k in range(len(w)):
            if abs(w[k]) > 1e-14:
                rows.append(I); cols.append(blk[k]); vals.append(w[k])
    Pb = sp.csr_matrix((vals, (rows, cols)), shape=(tf.nb, tc.nb))
    return sp.kron(Pb, sp.eye(3)).tocsr()


def block_jacobi(A, o):
    """6x6 vertex (u_v, theta_v) + 3x3 edge/centre block inverse, as in solver.cu"""
    t = o.topo
    n = A.shape[0]
    groups = []
    for v in range(t.nn):
        groups.append(np.r_[3 * v + np.arange(3), 3 * (t.nq2 + v) + np.arange(3)])
    for k in range(t.nn, t.nq2):
        groups.append(3 * k + np.arange(3))
    rows, cols, vals = [], [], []
    Ad = A.tocsr()
    for g in groups:
        B = Ad[g][:, g].toarray()
        Bi = np.linalg.inv(B)
        rr, cc = np.meshgrid(g, g, indexing="ij")
        rows.append(rr.ravel()); cols.append(cc.ravel()); vals.append(Bi.ravel())
    return sp.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(n, n))


def pcg(A, b, M, rtol=1e-10, maxit=200000):
    x = np.zeros_like(b); r = b.copy(); z = M(r); p = z.copy(); rz = r @ z; nb = np.linalg.norm(b)
    for it in range(1, maxit + 1):
        Ap = A @ p; al = rz / (p @ Ap); x += al * p; r -= al * Ap
        if np.linalg.norm(r) <= rtol * nb:
            return x, it
        z = M(r); rz2 = r @ z; p = z + (rz2 / rz) * p; rz = rz2
    return x, maxit


def cheb_smoother(A, Dinv, lmax, deg, lo_frac=0.1):
    """Chebyshev polynomial in Dinv A on [lo_frac lmax, 1.1 lmax] (symmetric, linear)"""
    a, bnd = lo_frac * lmax, 1.1 * lmax
    theta, delta = 0.5 * (bnd + a), 0.5 * (bnd - a)
    def apply(x, b):
        r = b - A @ x
        sigma = theta / delta; rho = 1.0 / sigma
        d = Dinv @ r / theta
        for k in range(deg):
            x = x + d
            if k == deg - 1: break
            r = r - A @ d
            rho_new = 1.0 / (2 * sigma - rho)
            d = rho_new * rho * d + 2 * rho_new / delta * (Dinv @ r)
            rho = rho_new
        return x
    return apply


def lam_max(A, Dinv, its=30):
    v = np.random.default_rng(0).standard_normal(A.shape[0])
    for _ in range(its):
        v = Dinv @ (A @ v); lam = np.linalg.norm(v); v /= lam
    return lam


if __name__ == "__main__":
    nf = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    nlev = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    deg = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    hth = float(sys.argv[4]) if len(sys.argv) > 4 else 0.1
    ns = [nf // 2 ** l for l in range(nlev)]
    lv = [level(n, h=hth) for n in ns]
    Ps = [prolongation(lv[l][0], lv[l + 1][0], ns[l + 1]) for l in range(nlev - 1)]
    # zero BC rows of P (fine BC dofs are fixed)
    for l in range(nlev - 1):
        keep = np.ones(lv[l][1].shape[0]); keep[lv[l][0].bc_dofs] = 0
        keepc = np.ones(lv[l + 1][1].shape[0]); keepc[lv[l + 1][0].bc_dofs] = 0
        Ps[l] = sp.diags(keep) @ Ps[l] @ sp.diags(keepc)
    Dinv = [block_jacobi(A, o) for o, A, b in lv]
    lm = [lam_max(lv[l][1], Dinv[l]) for l in range(nlev)]
    print("levels", ns, "dofs", [A.shape[0] for _, A, _ in lv], "lam_max(D^-1 A)", [round(x, 3) for x in lm])
    sm = [cheb_smoother(lv[l][1], Dinv[l], lm[l], deg) for l in range(nlev)]
    lu = spla.splu(lv[-1][1].tocsc())
    galerkin = "--galerkin" in sys.argv
    As = [lv[0][1]]
    for l in range(nlev - 1):
        As.append((Ps[l].T @ As[l] @ Ps[l]).tocsr() if galerkin else lv[l + 1][1])
    if galerkin:
        for l in range(1, nlev):
            bc = lv[l][0].bc_dofs
            As[l] = (As[l] + sp.csr_matrix((np.ones(bc.size), (bc, bc)), shape=As[l].shape)).tocsr()
        Dinv = [block_jacobi(As[l], lv[l][0]) for l in range(nlev)]
        lm = [lam_max(As[l], Dinv[l]) for l in range(nlev)]
        sm = [cheb_smoother(As[l], Dinv[l], lm[l], deg) for l in range(nlev)]
        lu = spla.splu(As[-1].tocsc())

    coarse_deg = 0
    for a in sys.argv:
        if a.startswith("--coarse-cheb="):
            coarse_deg = int(a.split("=")[1])
    lo = 0.1
    for a in sys.argv:
        if a.startswith("--coarse-lo="):
            lo = float(a.split("=")[1])
    if coarse_deg:
        smc = cheb_smoother(As[-1], Dinv[-1], lm[-1], coarse_deg, lo)

    def vcycle(l, b):
        if l == nlev - 1:
            if coarse_deg > 0:
                return smc(np.zeros_like(b), b)
            return lu.solve(b)
        x = sm[l](np.zeros_like(b), b)
        r = b - As[l] @ x
        x = x + Ps[l] @ vcycle(l + 1, Ps[l].T @ r)
        # symmetric post-smoothing: Chebyshev is a polynomial in D^-1 A -> same operator works
        x = sm[l](x, b)
        return x

    A, b = lv[0][1], lv[0][2]
    t0 = time.time(); x1, it1 = pcg(A, b, lambda r: vcycle(0, r)); t1 = time.time() - t0
    print("MG-PCG its", it1, "time %.1fs" % t1, "relres", np.linalg.norm(b - A @ x1) / np.linalg.norm(b))
    if "--jacobi" in sys.argv:
        x2, it2 = pcg(A, b, lambda r: Dinv[0] @ r)
        print("Jacobi-PCG its", it2, "diff", np.linalg.norm(x1 - x2) / np.linalg.norm(x2))
