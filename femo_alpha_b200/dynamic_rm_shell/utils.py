"""History <-> vector helpers -- drop-in for femo_alpha/dynamic_rm_shell/utils.py:9-23 (the PETSc helpers of that
file have no counterpart: vectors live on the GPU)."""
import numpy as np


def stack_array_into_vector(inp_arr):
    return np.ravel(inp_arr, order='F')


def reshape_vector_into_array(inp_vec, n_cols):
    n_rows = int(inp_vec.shape[0] / n_cols)
    if inp_vec.shape[0] - n_rows * n_cols != 0:
        raise ValueError("`inp_vec` of shape {} cannot be reshaped into {} rows and {} columns".format(
            inp_vec.shape, n_rows, n_cols))
    return np.reshape(inp_vec, (n_rows, n_cols), order='F')


def apply_hom_DirichletBCs_to_matrix(mat, bc_dof_idxs):
    mat[bc_dof_idxs, :] = 0.
    for bc_idx in bc_dof_idxs:
        mat[bc_idx, bc_idx] = 1.
    return mat
