"""femo_dRdf_T_apply_noalloc -- the allocation-free variant of femo_dRdf_T_apply (static device scratch, no cudaMalloc;
include/femo_b200.h) -- against the existing entry point, through the C ABI, on two calls in a row.  No counterpart in
the reference.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from tests.problems import make_problem, device_for

pytestmark = pytest.mark.gpu
MAX_CELLS = 65536           # FEMO_NOALLOC_MAX_CELLS


def _noalloc(d, arg, lam):
    """One call through the C ABI with device buffers owned by torch (the entry point itself allocates nothing)."""
    import torch
    from femo_alpha_b200 import _lib as L
    lam_t = torch.from_numpy(np.ascontiguousarray(lam, dtype=np.float64)).cuda()
    out_t = torch.full((d.n_mat,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    L.check(d.lib.femo_dRdf_T_apply_noalloc(d._ctx, L.INPUT_IDS[arg], C.c_void_p(lam_t.data_ptr()), C.c_void_p(out_t.data_ptr())))
    torch.cuda.synchronize()          # the call only queues work on the context's stream
    return out_t.cpu().numpy()


def _rel(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


@pytest.mark.parametrize("name", ["plate_varh", "plate_dg0", "warped"])
def test_no_alloc_entry_point_matches_regular_one(built_lib, name):
    p = make_problem(name)
    d = device_for(p)
    rng = np.random.default_rng(3)
    d.set_state(rng.standard_normal(d.ndof))
    # two calls in a row (different multipliers, different arguments): the shared scratch keeps nothing of the first
    for arg in ("thickness", "E", "thickness", "nu"):
        lam = rng.standard_normal(d.ndof)
        ref = d.dRdf_T_apply(arg, lam)                 # existing entry point (context-owned cudaMalloc'ed scratch)
        got = _noalloc(d, arg, lam)
        assert np.isfinite(got).all()
        assert _rel(got, ref) <= 1e-12, (name, arg, _rel(got, ref))
    # same input twice -> bitwise the same output
    lam = rng.standard_normal(d.ndof)
    assert np.array_equal(_noalloc(d, "thickness", lam), _noalloc(d, "thickness", lam))
    d.close()


def test_no_alloc_entry_point_rejects_what_it_cannot_serve(built_lib):
    import torch
    from femo_alpha_b200 import _lib as L
    from femo_alpha_b200.shell_device import ShellDevice
    from oracle.rm_shell_ref import plate_mesh
    d = device_for(make_problem("plate_varh"))
    buf = torch.zeros(max(d.ndof, d.n_mat), dtype=torch.float64, device="cuda")
    ptr = C.c_void_p(buf.data_ptr())
    assert d.lib.femo_dRdf_T_apply_noalloc(d._ctx, L.INPUT_IDS["F_solid"], ptr, ptr) != 0      # unsupported argument
    assert d.lib.femo_dRdf_T_apply_noalloc(d._ctx, L.INPUT_IDS["thickness"], None, ptr) != 0   # null pointer
    d.close()
    # one cell column more than the static scratch holds
    nodes, quads = plate_mesh(257, 256, 10.0, 10.0)
    assert quads.shape[0] > MAX_CELLS
    big = ShellDevice(nodes, quads)
    buf = torch.zeros(big.ndof, dtype=torch.float64, device="cuda")
    ptr = C.c_void_p(buf.data_ptr())
    assert big.lib.femo_dRdf_T_apply_noalloc(big._ctx, L.INPUT_IDS["thickness"], ptr, ptr) != 0
    msg = big.lib.femo_last_error().decode()
    assert "exceed the static scratch" in msg and str(MAX_CELLS) in msg
    # ... while the regular entry point serves it
    assert np.isfinite(big.dRdf_T_apply("thickness", np.ones(big.ndof))).all()
    big.close()
