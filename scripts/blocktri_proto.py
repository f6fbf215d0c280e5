"""NumPy prototype of csrc/coarse.cu: BFS level sets of the block-node graph -> block-tridiagonal LDL^T with explicit
inverses of the Schur complements (S_i^-1, G_i = O_i S_{i-1}^-1), recursive 2x2 SPD inversion, two sweeps per solve.
Checks the algebra against a sparse direct solve and prints the front sizes (memory / launch counts of the device version).

    python scripts/blocktri_proto.py plate 24 24 | wingbox
"""
import os
import sys

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def bfs_levels(indptr, indices, start):
    n = indptr.size - 1
    lev = np.full(n, -1, dtype=np.int64)
    lev[start] = 0
    front = np.array([start])
    k = 0
    while front.size:
        nb = np.unique(np.concatenate([indices[indptr[i]:indptr[i + 1]] for i in front]))
        nb = nb[lev[nb] < 0]
        k += 1
        lev[nb] = k
        front = nb
    return lev


def pseudo_peripheral(indptr, indices, seed=0):
    deg = np.diff(indptr)
    s, ecc = seed, -1
    for _ in range(8):
        lev = bfs_levels(indptr, indices, s)
        e = lev.max()
        if e <= ecc:
            break
        ecc = e
        last = np.nonzero(lev == e)[0]
        s = last[np.argmin(deg[last])]
    return s


def spd_inv(S, base=64):
    m = S.shape[0]
    if m <= base:
        return np.linalg.inv(S)
    m1 = ((m // 2 + 31) // 32) * 32
    A, B, C = S[:m1, :m1], S[m1:, :m1], S[m1:, m1:]
    Ai = spd_inv(A, base)
    W = B @ Ai
    Sc = C - W @ B.T
    Sci = spd_inv(Sc, base)
    O10 = -Sci @ W
    O00 = Ai - W.T @ O10
    return np.block([[O00, O10.T], [O10, Sci]])


def chol_inv(S, nb=64):
    """The device algorithm of csrc/coarse.cu: right-looking blocked Cholesky with explicit inverses T_kk of the nb x nb
    diagonal factors, recursive triangular inversion, S^-1 = Linv^T Linv."""
    import scipy.linalg as sl
    S = S.copy()
    m = S.shape[0]
    L = np.zeros_like(S)
    T = {}
    for k in range(0, m, nb):
        e = min(k + nb, m)
        Lkk = np.linalg.cholesky(S[k:e, k:e])
        Tkk = sl.solve_triangular(Lkk, np.eye(e - k), lower=True)
        L[k:e, k:e] = Lkk
        T[k] = Tkk
        if e < m:
            L[e:, k:e] = S[e:, k:e] @ Tkk.T
            S[e:, e:] -= L[e:, k:e] @ L[e:, k:e].T
    Li = np.zeros_like(S)

    def tri(r0, mm):
        if mm <= nb:
            Li[r0:r0 + mm, r0:r0 + mm] = T[r0]
            return
        m1 = ((mm // nb + 1) // 2) * nb
        tri(r0, m1); tri(r0 + m1, mm - m1)
        W = L[r0 + m1:r0 + mm, r0:r0 + m1] @ Li[r0:r0 + m1, r0:r0 + m1]
        Li[r0 + m1:r0 + mm, r0:r0 + m1] = -Li[r0 + m1:r0 + mm, r0 + m1:r0 + mm] @ W
    tri(0, m)
    return Li.T @ Li


def main():
    from oracle.rm_shell_ref import RMShellOracle, plate_mesh
    if sys.argv[1] == "plate":
        nx, ny = int(sys.argv[2]), int(sys.argv[3])
        nodes, quads = plate_mesh(nx, ny, 10.0, 10.0 * ny / nx)
        clamp = lambda x: np.less(x[0], 3e-16)
    else:
        g = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "wingbox_mesh.npz"))
        nodes, quads = g["nodes"], g["quads"]
        clamp = lambda x: np.less_equal(x[1], 1e-3)
    o = RMShellOracle(nodes, quads, False, False)
    nn = nodes.shape[0]
    o.set_inputs(thickness=np.full(nn, 0.05), E=np.full(nn, 1e8), nu=np.full(nn, 0.3), density=np.full(nn, 10.0),
                 F_solid=np.tile([0.0, 0.0, 5.0], nn))
    o.set_strong_bc(clamp)
    K = o.assemble_dRdu().tocsr()
    m = np.ones(K.shape[0]); m[o.bc_dofs] = 0
    D = sp.diags(m)
    A = (D @ K @ D + sp.diags(1 - m)).tocsr()
    n = A.shape[0]
    nb = n // 3
    # block graph: the PATTERN of K (clamped nodes stay connected; their rows of A are identity rows inside their front)
    Bg = sp.csr_matrix((np.ones(K.nnz), K.indices // 3, K.indptr), shape=(n, nb))
    Bg = sp.csr_matrix((np.ones(n), (np.arange(n) // 3, np.arange(n))), shape=(nb, n)) @ Bg
    Bg.sum_duplicates(); Bg.sort_indices()
    s = pseudo_peripheral(Bg.indptr, Bg.indices)
    lev = bfs_levels(Bg.indptr, Bg.indices, s)
    assert lev.min() >= 0, "disconnected graph"
    order = np.argsort(lev, kind="stable")
    nl = lev.max() + 1
    counts = np.bincount(lev, minlength=nl) * 3
    lp = np.concatenate([[0], np.cumsum(counts)])
    print("n = %d dofs, %d level sets, front max %d mean %.0f dofs; storage 3 sum m^2 = %.2f GB, dense n^2 = %.2f GB"
          % (n, nl, counts.max(), counts.mean(), 3 * (counts.astype(float) ** 2).sum() * 8 / 1e9, n * n * 8.0 / 1e9))
    if n > 60000:
        return
    perm = (3 * order[:, None] + np.arange(3)[None, :]).ravel()
    Ap = A[perm][:, perm].tocsr()
    Sinv, G = [], [None]
    for i in range(nl):
        Di = Ap[lp[i]:lp[i + 1], lp[i]:lp[i + 1]].toarray()
        if i:
            Oi = Ap[lp[i]:lp[i + 1], lp[i - 1]:lp[i]].toarray()
            Gi = Oi @ Sinv[-1]
            G.append(Gi)
            Di = Di - Gi @ Oi.T
        if os.environ.get("CHOL", "1") == "1":
            Sinv.append(chol_inv(Di))
        elif os.environ.get("EQUIL", "1") == "1":          # symmetric diagonal equilibration around the inversion
            d = 1.0 / np.sqrt(np.diag(Di))
            Sinv.append(d[:, None] * spd_inv(d[:, None] * Di * d[None, :], int(os.environ.get("BASE", "64"))) * d[None, :])
        else:
            Sinv.append(spd_inv(Di, int(os.environ.get("BASE", "64"))))
    rng = np.random.default_rng(0)
    b = rng.standard_normal(n)
    bp = b[perm]
    z = [bp[lp[0]:lp[1]]]
    for i in range(1, nl):
        z.append(bp[lp[i]:lp[i + 1]] - G[i] @ z[-1])
    x = [None] * nl
    x[-1] = Sinv[-1] @ z[-1]
    for i in range(nl - 2, -1, -1):
        x[i] = Sinv[i] @ z[i] - G[i + 1].T @ x[i + 1]
    xs = np.empty(n); xs[perm] = np.concatenate(x)
    ref = spla.splu(A.tocsc()).solve(b)
    r = b - A @ xs
    print("block-tridiagonal solve vs splu: rel.err %.2e, residual %.2e" % (np.abs(xs - ref).max() / np.abs(ref).max(),
                                                                          np.linalg.norm(r) / np.linalg.norm(b)))
    # one step of iterative refinement
    bp = r[perm]
    z = [bp[lp[0]:lp[1]]]
    for i in range(1, nl):
        z.append(bp[lp[i]:lp[i + 1]] - G[i] @ z[-1])
    x[-1] = Sinv[-1] @ z[-1]
    for i in range(nl - 2, -1, -1):
        x[i] = Sinv[i] @ z[i] - G[i + 1].T @ x[i + 1]
    dx = np.empty(n); dx[perm] = np.concatenate(x)
    xs += dx
    print("after one refinement: rel.err %.2e, residual %.2e" % (np.abs(xs - ref).max() / np.abs(ref).max(),
                                                                 np.linalg.norm(b - A @ xs) / np.linalg.norm(b)))


if __name__ == "__main__":
    main()
