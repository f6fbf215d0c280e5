"""GPU parity: the CUDA path through the C ABI against the CPU oracle (same seeded inputs).

Tolerances (BASELINE.json north_star): assembled matrix and vector relative error <= 1e-12;
displacement, compliance, adjoint thickness gradient <= 1e-8.
"""
import numpy as np
import pytest

from tests.problems import PROBLEMS, make_problem, oracle_for, device_for

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.fixture(scope="module", params=PROBLEMS)
def pair(request, built_lib):
    p = make_problem(request.param)
    return p, oracle_for(p), device_for(p)


def test_sizes_and_pattern(pair):
    p, o, d = pair
    t = o.topo
    assert (d.nn, d.ne, d.nc, d.ndof) == (t.nn, t.ne, t.nc, t.ndof)
    assert np.array_equal(d.cell_blocks(), t.cell_blocks)
    assert np.allclose(d.block_coords(), t.block_coords(), rtol=0, atol=1e-14)
    K = o.assemble_dRdu().tobsr((3, 3))
    K.sort_indices()
    rowptr, colidx, _ = d.get_bsr(values=False)
    assert np.array_equal(rowptr, K.indptr) and np.array_equal(colidx, K.indices)


def test_jacobian_matrix(pair):
    p, o, d = pair
    K = o.assemble_dRdu().tocsr()
    Kd = d.get_csr()
    assert rel(Kd.toarray(), K.toarray()) <= 1e-12
    A = o.apply_bc_matrix(K)
    Ad = d.get_csr(with_bc=True)
    assert rel(Ad.toarray(), A.toarray()) <= 1e-12


def test_assembly_bitwise_deterministic(pair):
    p, o, d = pair
    d.assemble_jacobian()
    v1 = d.get_bsr()[2].copy()
    d.set_input("thickness", p["thickness"])      # invalidates, forces a re-assembly
    d.assemble_jacobian()
    v2 = d.get_bsr()[2]
    assert np.array_equal(v1, v2)


def test_load_and_residual(pair):
    p, o, d = pair
    F = o.assemble_load()
    assert rel(d.assemble_load(), F) <= 1e-12
    rng = np.random.default_rng(3)
    w = rng.standard_normal(o.topo.ndof)
    R = o.assemble_residual(w)
    assert rel(d.assemble_residual(w), R) <= 1e-12
    assert rel(d.jacobian_apply(w), o.assemble_dRdu() @ w) <= 1e-12


def test_state_adjoint_and_gradient(pair):
    p, o, d = pair
    w = o.solve_state()
    it, rr = d.solve_state(rtol=1e-13, max_it=20000)
    wd = d.get_state()
    assert rr <= 1e-12
    assert rel(wd, w) <= 1e-8
    o.w = w
    # outputs on the oracle state (isolates the functional kernels from solver error)
    d.set_state(w)
    for name, ref in (("compliance", o.compliance(w)), ("mass", o.mass())):
        assert abs(d.output(name) - ref) <= 1e-11 * abs(ref), name
    # 0.5 w^T K w cancels ~7 digits (|K||w| >> |F|) in the oracle and on the device alike
    assert abs(d.output("elastic_energy") - o.elastic_energy(w)) <= 1e-7 * abs(o.elastic_energy(w))
    assert rel(d.output_grad("compliance", "disp_solid"), o.dcompliance_du(w)) <= 1e-12
    # the H1 regulariser gradient vanishes for uniform thickness: compare on the scale alpha1*|h|
    assert np.abs(d.output_grad("compliance", "thickness") - o.dcompliance_dh()).max() <= \
        1e-11 * max(np.abs(o.dcompliance_dh()).max(), 1e-2 * np.abs(o.h).max())
    assert rel(d.output_grad("mass", "thickness"), o.dmass_dh()) <= 1e-12
    assert rel(d.output_grad("mass", "density"), o.dmass_drho()) <= 1e-12
    assert rel(d.output_grad("elastic_energy", "disp_solid"), o.delastic_energy_du(w)) <= 1e-9
    assert rel(d.output_grad("elastic_energy", "thickness"), o.delastic_energy_dparam(w, "dh")) <= 1e-9
    # partials of the residual
    rng = np.random.default_rng(5)
    lam = rng.standard_normal(o.topo.ndof)
    for arg, mode in (("thickness", "dh"), ("E", "dE"), ("nu", "dnu")):
        M = o.assemble_dRdparam(w, mode)
        assert rel(d.dRdf_T_apply(arg, lam), M.T @ lam) <= 1e-11, arg
        dp = rng.standard_normal(M.shape[1])
        assert rel(d.dRdf_apply(arg, dp), M @ dp) <= 1e-11, arg
    Mf = o.assemble_dRdf()
    assert rel(d.dRdf_T_apply("F_solid", lam), Mf.T @ lam) <= 1e-12
    df = rng.standard_normal(Mf.shape[1])
    assert rel(d.dRdf_apply("F_solid", df), Mf @ df) <= 1e-12
    # adjoint solve + total thickness gradient of the compliance
    rhs = o.dcompliance_du(w)
    lam_o = o.solve_adjoint(rhs)
    lam_d = d.solve_adjoint(rhs, rtol=1e-13, max_it=20000)
    assert rel(lam_d, lam_o) <= 1e-8
    g_o = o.total_dcompliance_dh()
    g_d = d.output_grad("compliance", "thickness") - d.dRdf_T_apply("thickness", lam_d)
    assert rel(g_d, g_o) <= 1e-8


@pytest.mark.parametrize("name", ["plate_4x20", "warped"])
def test_penalty_bc(name, built_lib):
    p = make_problem(name)
    o, d = oracle_for(p, penalty=True), device_for(p, penalty=True)
    K = o.assemble_dRdu().tocsr()
    assert rel(d.get_csr().toarray(), K.toarray()) <= 1e-12
    w = o.solve_state()
    it, rr = d.solve_state(rtol=1e-13, max_it=40000)
    free = np.ones(o.topo.ndof, bool)
    assert rel(d.get_state(), w) <= 1e-7          # beta = 1e15: the direct solve itself carries ~1e-10
    d.set_state(w)
    assert abs(d.output("elastic_energy") - o.elastic_energy(w)) <= 1e-6 * abs(o.elastic_energy(w))


def test_quadrature_degree_and_errors(built_lib):
    p = make_problem("plate_dg0")
    o, d = oracle_for(p, quadrature_degree=4), device_for(p, quadrature_degree=4)
    assert d.sizes.n_qp == 9
    assert rel(d.get_csr().toarray(), o.assemble_dRdu().toarray()) <= 1e-12
    with pytest.raises(RuntimeError):
        d.set_input("thickness", np.ones(3))
    from femo_alpha_b200.shell_device import ShellDevice
    with pytest.raises(RuntimeError):
        ShellDevice(p["nodes"], p["quads"][:, :3])


@pytest.mark.parametrize("name", ["plate_varh", "plate_dg0", "warped"])
def test_stress_functionals(name, built_lib):
    p = make_problem(name)
    o, d = oracle_for(p), device_for(p)
    w = o.solve_state()
    d.set_state(w)
    m, rho = 1e-4, 6.0                      # moderate exponent: keeps the complex-step oracle well scaled
    d.set_stress_params(m, rho)
    ref = o.pnorm_stress(w, m=m, rho=rho)
    assert abs(d.pnorm_stress() - ref) <= 1e-10 * abs(ref)
    for arg, key in (("disp_solid", "w"), ("thickness", "thickness"), ("E", "E"), ("nu", "nu")):
        g_ref = o.pnorm_stress_grad(w, key, m=m, rho=rho)
        assert rel(d.pnorm_stress_grad(arg), g_ref) <= 1e-9, arg
    for surf in ("Top", "Mid", "Bot"):
        assert rel(d.stress_field(surf).reshape(-1, 4), o.stress_field(w, surf)) <= 1e-10
    # sub-domain p-norm with its own area
    cells = np.arange(0, o.topo.nc, 3)
    tags = -np.ones(o.topo.nc, dtype=np.int32); tags[cells] = 2
    d.set_cell_tags(tags)
    ref_t = o.pnorm_stress(w, m=m, rho=rho, cells=cells)
    assert abs(d.pnorm_stress(tag=2) - ref_t) <= 1e-10 * abs(ref_t)
    # the reference's parameters (m=1e-6, rho=100) evaluate finitely
    d.set_stress_params(1e-6, 100.0)
    assert np.isfinite(d.pnorm_stress())


def _sym_check(Kd):
    A = Kd.toarray()
    D = A - A.T
    blk = np.arange(A.shape[0]) // 3
    off = blk[:, None] != blk[None, :]
    assert not D[off].any()                     # block and transposed block are written from one value
    assert np.abs(D).max() <= 1e-14 * np.abs(A).max()   # 3x3 diagonal blocks: symmetric to round-off


@pytest.mark.parametrize("chunk", ["0", "1", "3", "5", "64"])
@pytest.mark.parametrize("name", ["plate_varh", "skew_planar", "plate_dg0"])
def test_chunked_symmetric_assembly_any_chunk_size(built_lib, monkeypatch, name, chunk):
    """Second-generation planar kernel (FEMO_TILE=0): destinations owned per chunk of cells, halo cells pulled;
    every chunk size (0 = plan disabled, contribution-gather kernel) must give the oracle's matrix."""
    monkeypatch.setenv("FEMO_TILE", "0")
    monkeypatch.setenv("FEMO_ASM_CHUNK", chunk)
    p = make_problem(name)
    o, d = oracle_for(p), device_for(p)
    K = o.assemble_dRdu().tocsr()
    Kd = d.get_csr()
    assert rel(Kd.toarray(), K.toarray()) <= 1e-12
    if chunk not in ("0", "64"):                # 64: the plan needs more shared memory than a CTA has -> gather kernel
        _sym_check(Kd)
    d.close()


@pytest.mark.parametrize("ws", ["2", "1", "0"])
@pytest.mark.parametrize("order", ["morton", "identity"])
@pytest.mark.parametrize("name", ["plate_4x20", "plate_varh", "skew_planar", "plate_dg0"])
def test_tiled_assembly(built_lib, monkeypatch, name, order, ws):
    """Default planar kernel (assemble_tiles.cu): register-tiled moments per chunk of 32 cells in Morton (or the
    caller's) order; the matrix is the oracle's, symmetric to the bit, and identical from run to run."""
    monkeypatch.setenv("FEMO_TILE_ORDER", order)
    monkeypatch.setenv("FEMO_TILE_WS", ws)          # 2: 8+8 warp roles (setmaxnreg), 1: 6+6 warp roles, 0: two plain CTAs per SM
    p = make_problem(name)
    o, d = oracle_for(p), device_for(p)
    K = o.assemble_dRdu().tocsr()
    Kd = d.get_csr()
    assert rel(Kd.toarray(), K.toarray()) <= 1e-12
    _sym_check(Kd)
    d.assemble_jacobian()
    assert np.array_equal(d.get_csr().data, Kd.data)
    d.close()


@pytest.mark.parametrize("shape", [(23, 17), (150, 40)])
@pytest.mark.parametrize("ws", ["2", "1"])
def test_tiled_assembly_many_chunks(built_lib, monkeypatch, shape, ws):
    """Plates spanning 25 / 375 chunks (the larger one gives every persistent CTA several chunks, so the
    double-buffered moment hand-over between compute and writer warps is exercised); quadrature degree 2, 4, 6."""
    from oracle.rm_shell_ref import plate_mesh, RMShellOracle
    from femo_alpha_b200.shell_device import ShellDevice
    monkeypatch.setenv("FEMO_TILE_WS", ws)
    rng = np.random.default_rng(7)
    nodes, quads = plate_mesh(shape[0], shape[1], 4.0, 3.0)
    nn = nodes.shape[0]
    nodes[:, :2] += 0.1 * min(4.0 / shape[0], 3.0 / shape[1]) * rng.standard_normal((nn, 2))
    inp = dict(thickness=0.05 * (1 + 0.5 * rng.random(nn)), E=1e6 * (1 + 0.2 * rng.random(nn)),
               nu=0.25 + 0.1 * rng.random(nn))
    for deg in ((2, 4, 6) if shape[0] < 100 else (6,)):
        o = RMShellOracle(nodes, quads, False, False, deg)
        o.set_inputs(**inp)
        d = ShellDevice(nodes, quads, False, False, deg)
        for k, v in inp.items():
            d.set_input(k, v)
        Kd = d.get_csr()
        K = o.assemble_dRdu(with_penalty=False).tocsr()
        assert abs(Kd - K).max() <= 1e-12 * abs(K).max()
        assert abs(Kd - Kd.T).max() <= 1e-14 * abs(K).max()
        d.close()


@pytest.mark.parametrize("name", ["plate_varh", "plate_dg0", "warped"])
def test_evaluate_residual_equals_separate_calls(built_lib, name):
    """femo_evaluate_residual (inputs up, assembly, residual back in row ranges, copies on a second stream) against
    set_input + assemble_residual and against the oracle; twice, with changed inputs, resident and explicit state."""
    p = make_problem(name)
    o, d = oracle_for(p), device_for(p)
    rng = np.random.default_rng(5)
    w = rng.standard_normal(d.ndof)
    for scale in (1.0, 1.2):
        inputs = {k: scale * np.asarray(p[k]) if k == "thickness" else np.asarray(p[k]) for k in ("thickness", "E", "nu", "F_solid")}
        if p["uhat"] is not None:
            inputs["uhat"] = np.asarray(p["uhat"])
        got = d.evaluate_residual(inputs, w)
        for k, v in inputs.items():
            d.set_input(k, v)
        ref = d.assemble_residual(w)
        assert np.array_equal(got, ref)
        o.set_inputs(thickness=inputs["thickness"])
        R = o.assemble_residual(w)
        assert np.abs(got - R).max() <= 1e-12 * np.abs(R).max()
    d.set_state(w)
    assert np.array_equal(d.evaluate_residual({"thickness": inputs["thickness"]}), ref)      # resident state, one input
    d.close()


@pytest.mark.parametrize("ws", ["2", "1"])
def test_halo_records_staged_in_shared_memory_change_nothing(built_lib, monkeypatch, ws):
    """The tiled kernel reads the records of a chunk's halo cells from global memory per halo item (default) or stages them
    in shared memory (FEMO_TILE_HREC=1, double-buffered cp.async): same numbers in the same order, so the
    matrices are bitwise equal -- on a distorted plate with ~90 chunks, tilted out of the xy plane."""
    from oracle.rm_shell_ref import plate_mesh
    from femo_alpha_b200.shell_device import ShellDevice
    monkeypatch.setenv("FEMO_TILE_WS", ws)
    rng = np.random.default_rng(9)
    nodes, quads = plate_mesh(48, 30, 4.0, 3.0)
    nn = nodes.shape[0]
    nodes[:, :2] += 0.1 * (4.0 / 48) * rng.standard_normal((nn, 2))
    nodes[:, 2] = 0.3 * nodes[:, 0]
    inp = dict(thickness=0.05 * (1 + 0.5 * rng.random(nn)), E=1e6 * (1 + 0.2 * rng.random(nn)), nu=0.25 + 0.1 * rng.random(nn))
    vals = []
    for hrec in ("1", "0"):
        monkeypatch.setenv("FEMO_TILE_HREC", hrec)
        d = ShellDevice(nodes, quads)
        assert d.sizes.all_planar == 1
        for k, v in inp.items():
            d.set_input(k, v)
        vals.append(d.get_bsr()[2].copy())
        d.close()
    assert np.array_equal(vals[0], vals[1])


def test_jacobian_at_benchmark_size_matches_oracle_strip(built_lib):
    """Parity at the size bench.py runs (config 3: 816 x 816 cells, 10 002 534 dofs, 81.9 M blocks; every persistent CTA of
    the tiled kernel walks ~281 chunks, int32 positions up to 8.2e7).  The C port of the oracle assembles two 816 x 6-cell
    strips of the SAME plate and inputs; rows of nodes strictly inside a strip couple to nodes of the strip only, so
    there  K_full . v  ==  K_strip . v|strip  and the diagonals agree, to 1e-12."""
    from bench import plate, inputs_for
    from femo_alpha_b200.shell_device import ShellDevice
    from oracle import c_port
    from oracle.rm_shell_ref import ShellTopology
    n, rows = 816, 6
    nodes, quads = plate(n, n, 0.0, 10.0, 10.0)
    inp = inputs_for(nodes)
    d = ShellDevice(nodes, quads)
    for k in ("thickness", "E", "nu"):
        d.set_input(k, inp[k])
    assert d.ndof == 10002534
    xb = d.block_coords()
    rng = np.random.default_rng(11)
    v = rng.standard_normal(d.ndof)
    Kv = d.jacobian_apply(v).reshape(-1, 3)
    diag = _bsr_diagonal(d)
    hy = 10.0 / n

    def key(x):                      # half-cell grid coordinates are integers; rotation nodes are told apart by `kind`
        return np.rint(2 * x[:, 0] / hy).astype(np.int64) * 4096 + np.rint(2 * x[:, 1] / hy).astype(np.int64)

    kfull = key(xb)
    kind_full = (np.arange(d.nb) >= d.nq2).astype(np.int64)
    order = np.lexsort((kfull, kind_full))
    skeys = kind_full[order] * (1 << 40) + kfull[order]
    worst = 0.0
    for j0 in (0, 405, n - rows):                                 # bottom edge, middle, top edge of the plate
        ys = j0 * hy
        sn, sq = plate(n, rows, 0.0, 10.0, rows * hy)
        sn = sn + np.array([0.0, ys, 0.0])
        topo = ShellTopology(sn, sq)
        x, y = sn[:, 0], sn[:, 1]
        h_s = 0.1 * (1 + 0.5 * np.sin(2 * np.pi * x / 10) * np.cos(2 * np.pi * y / 10))
        asm = c_port.CsrAssembler(topo)
        asm.assemble(h_s, np.full(sn.shape[0], 1e8), np.full(sn.shape[0], 0.3))
        Ks = asm.matrix()
        xs = topo.block_coords()
        kind_s = (np.arange(topo.nb) >= topo.nq2).astype(np.int64)
        pos = np.searchsorted(skeys, kind_s * (1 << 40) + key(xs))
        full_of = order[pos]                                      # strip block node -> block node of the full plate
        assert np.array_equal(kfull[full_of], key(xs)) and np.array_equal(kind_full[full_of], kind_s)
        vs = v.reshape(-1, 3)[full_of].ravel()
        Ksv = (Ks @ vs).reshape(-1, 3)
        inside = (xs[:, 1] > ys + 1e-9) & (xs[:, 1] < ys + rows * hy - 1e-9)
        if j0 == 0:
            inside |= xs[:, 1] < 1e-9                             # the plate's own edge is not a cut
        if j0 == n - rows:
            inside |= xs[:, 1] > 10.0 - 1e-9
        e1 = np.abs(Kv[full_of][inside] - Ksv[inside]).max() / np.abs(Ksv[inside]).max()
        ds = Ks.diagonal().reshape(-1, 3)
        e2 = np.abs(diag[full_of][inside] - ds[inside]).max() / np.abs(ds[inside]).max()
        print("strip at row %d: %d rows compared, K.v rel.err %.2e, diagonal rel.err %.2e" % (j0, 3 * inside.sum(), e1, e2))
        worst = max(worst, e1, e2)
    d.close()
    assert worst <= 1e-12


def _bsr_diagonal(d):
    """(nb, 3) diagonal of dR/du: the diagonal blocks are fetched by position (femo_get_blocks), not the 5.9 GB of values"""
    from femo_alpha_b200 import _lib as L
    rowptr, colidx, _ = d.get_bsr(values=False)
    rows = np.repeat(np.arange(d.nb, dtype=np.int64), np.diff(rowptr.astype(np.int64)))
    dpos = np.ascontiguousarray(np.nonzero(colidx == rows)[0], dtype=np.int64)
    assert dpos.size == d.nb
    blocks = np.empty((d.nb, 9))
    L.check(d.lib.femo_get_blocks(d._ctx, d.nb, L.ptr(dpos), L.ptr(blocks)))
    return np.ascontiguousarray(blocks[:, [0, 4, 8]])


@pytest.mark.parametrize("nb", ["1", "2", "3", "4", "legacy"])
def test_warped_cells_cell_centric_assembly(built_lib, monkeypatch, nb):
    """Non-planar cells (assemble.cu: k_warped_elem + k_warped_gather): element matrices per group of 32 cells with 1, 2 or
    3 node pairs per task, and the round-1 contribution-gather kernel (FEMO_WARPED=0), against the oracle on a doubly
    curved, distorted plate (437 cells: 14 groups, the last one partly filled) with mesh motion; symmetric and
    identical from run to run."""
    from oracle.rm_shell_ref import plate_mesh, RMShellOracle
    from femo_alpha_b200.shell_device import ShellDevice
    if nb == "legacy":
        monkeypatch.setenv("FEMO_WARPED", "0")
    else:
        monkeypatch.setenv("FEMO_WARPED_NB", nb)
    rng = np.random.default_rng(13)
    nodes, quads = plate_mesh(23, 19, 4.0, 3.0)
    nn = nodes.shape[0]
    nodes[:, :2] += 0.1 * (3.0 / 19) * rng.standard_normal((nn, 2))
    nodes[:, 2] = 0.3 * np.sin(nodes[:, 0]) * np.cos(1.3 * nodes[:, 1])
    inp = dict(thickness=0.05 * (1 + 0.5 * rng.random(nn)), E=1e6 * (1 + 0.2 * rng.random(nn)), nu=0.25 + 0.1 * rng.random(nn),
               uhat=0.01 * rng.standard_normal(3 * nn))
    for deg in (6, 4):
        o = RMShellOracle(nodes, quads, False, False, deg)
        o.set_inputs(**inp)
        d = ShellDevice(nodes, quads, False, False, deg)
        assert d.sizes.all_planar == 0
        for k, v in inp.items():
            d.set_input(k, v)
        Kd = d.get_csr()
        K = o.assemble_dRdu(with_penalty=False).tocsr()
        assert abs(Kd - K).max() <= 1e-12 * abs(K).max()
        if nb != "legacy":
            _sym_check(Kd)
        d.assemble_jacobian()
        assert np.array_equal(d.get_csr().data, Kd.data)
        d.close()
