"""The C/OpenMP CPU baseline (oracle/rm_shell_c.c) agrees with the NumPy oracle."""
import numpy as np
import pytest

from oracle import c_port
from tests.problems import make_problem, oracle_for


@pytest.mark.parametrize("name", ["plate_varh", "plate_dg0", "warped"])
def test_c_port_matches_numpy_oracle(name):
    o = oracle_for(make_problem(name))
    Ke = o.element_matrices()
    Kc = c_port.element_matrices(o.topo, o.h, o.E, o.nu, o.uhat, o.ewm)
    assert np.abs(Kc - Ke).max() <= 1e-12 * np.abs(Ke).max()
    asm = c_port.CsrAssembler(o.topo)
    nthreads = asm.assemble(o.h, o.E, o.nu, o.uhat, o.ewm)
    assert nthreads >= 1
    K = o.assemble_dRdu().toarray()
    assert np.abs(asm.matrix().toarray() - K).max() <= 1e-12 * np.abs(K).max()
