"""History <-> vector layout of the time-stepped path.

The CSDL state of the dynamic operations is ONE vector holding every time level: level k occupies entries
[k * n_dof, (k + 1) * n_dof), i.e. the (n_dof, n_levels) history array flattened column by column.  The names are those
the reference's operations import (femo_alpha/dynamic_rm_shell/utils.py:9-17); its PETSc / dense-matrix helpers have no
counterpart here because matrices and vectors live on the GPU."""
import numpy as np


def stack_array_into_vector(history):
    """(n_dof, n_levels) history -> flat vector, one time level after the other."""
    h = np.asarray(history)
    return np.ascontiguousarray(h.T).reshape(-1)


def reshape_vector_into_array(flat, n_levels):
    """Inverse of `stack_array_into_vector`; refuses a length that is not a whole number of time levels."""
    flat = np.asarray(flat)
    n_dof, rest = divmod(flat.shape[0], int(n_levels))
    if rest:
        raise ValueError("a history vector of %d entries does not split into %d time levels" % (flat.shape[0], n_levels))
    return flat.reshape(int(n_levels), n_dof).T
