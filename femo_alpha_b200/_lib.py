"""ctypes binding of libfemo_b200.so (C ABI declared in include/femo_b200.h).

There is no CPU fallback: if the shared library is missing, or no B200 is
visible when a context is created, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfemo_b200.so")
if os.environ.get("FEMO_LIB_PATH"):          # experiments: another build of the same library (scripts/asm_experiments.py)
    LIB_PATH = os.environ["FEMO_LIB_PATH"]

THICKNESS, F_SOLID, E, NU, DENSITY, UHAT = range(6)
COMPLIANCE, MASS, ELASTIC_ENERGY, PNORM_STRESS, VOLUME = range(5)
ARG_STATE = 100
INPUT_IDS = {"thickness": THICKNESS, "F_solid": F_SOLID, "E": E, "nu": NU, "density": DENSITY, "uhat": UHAT}
OUTPUT_IDS = {"compliance": COMPLIANCE, "mass": MASS, "elastic_energy": ELASTIC_ENERGY, "pnorm_stress": PNORM_STRESS,
              "volume": VOLUME}


class ShellOpts(C.Structure):
    _fields_ = [("element_wise_material", C.c_int32), ("elementwise_pressure", C.c_int32),
                ("quadrature_degree", C.c_int32), ("reserved", C.c_int32)]


class Sizes(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("n_nodes", "n_edges", "n_cells", "n_blocks", "n_dof", "nnz_blocks",
                                         "n_contrib", "n_material", "n_force")] + \
               [("n_qp", C.c_int32), ("all_planar", C.c_int32)]


_P = C.c_void_p
_D = C.POINTER(C.c_double)
_I = C.POINTER(C.c_int32)

# name -> argtypes (restype is int unless stated); must list every symbol of include/femo_b200.h
SIGNATURES = {
    "femo_ctx_create": [C.c_int, C.POINTER(_P)],
    "femo_ctx_destroy": [_P],
    "femo_set_stream": [_P, _P],
    "femo_mesh_set": [_P, C.c_int64, _P, C.c_int64, C.c_int, _P],
    "femo_shell_setup": [_P, C.POINTER(ShellOpts)],
    "femo_get_sizes": [_P, C.POINTER(Sizes)],
    "femo_get_block_coords": [_P, _P],
    "femo_get_cell_blocks": [_P, _P],
    "femo_set_strong_bc": [_P, C.c_int64, _P],
    "femo_set_penalty_bc": [_P, C.c_int64, _P],
    "femo_set_input": [_P, C.c_int, _P, C.c_int64, C.c_int],
    "femo_get_input": [_P, C.c_int, _P, C.c_int64, C.c_int],
    "femo_set_state": [_P, _P, C.c_int],
    "femo_get_state": [_P, _P, C.c_int],
    "femo_assemble_jacobian": [_P],
    "femo_assemble_residual": [_P, _P, _P, C.c_int],
    "femo_evaluate_residual": [_P, C.c_int, _P, _P, _P, _P, _P],
    "femo_assemble_load": [_P, _P, C.c_int],
    "femo_jacobian_apply": [_P, _P, _P, C.c_int, C.c_int],
    "femo_get_bsr": [_P, _P, _P, _P, C.c_int],
    "femo_get_blocks": [_P, C.c_int64, _P, _P],
    "femo_solve_state": [_P, C.c_double, C.c_int, C.POINTER(C.c_int), _D],
    "femo_solve_adjoint": [_P, _P, _P, C.c_double, C.c_int, C.POINTER(C.c_int), _D, C.c_int],
    "femo_solve_forward": [_P, _P, _P, C.c_double, C.c_int, C.POINTER(C.c_int), _D, C.c_int],
    "femo_dRdf_T_apply": [_P, C.c_int, _P, _P, C.c_int],
    "femo_dRdf_apply": [_P, C.c_int, _P, _P, C.c_int],
    "femo_dRdf_T_apply_noalloc": [_P, C.c_int, _P, _P],
    "femo_output": [_P, C.c_int, _D],
    "femo_output_grad": [_P, C.c_int, C.c_int, _P, C.c_int],
    "femo_set_stress_params": [_P, C.c_double, C.c_double],
    "femo_set_cell_tags": [_P, C.c_int64, _P],
    "femo_pnorm_stress": [_P, C.c_int, _D],
    "femo_pnorm_stress_grad": [_P, C.c_int, C.c_int, _P, C.c_int],
    "femo_stress_field": [_P, C.c_int, _P, C.c_int],
    "femo_nccl_unique_id": [_P],
    "femo_dist_setup": [_P, C.c_int, C.c_int, C.c_int, _P, _P, _P, _P, _P, _P, C.c_int64],
    "femo_dist_init": [_P, _P],
    "femo_dist_share_comm": [_P, _P],
    "femo_halo_exchange": [_P, _P, C.c_int],
    "femo_halo_accumulate": [_P, _P, C.c_int],
    "femo_dist_allreduce": [_P, _D, C.c_int],
    "femo_dist_counts": [_P, C.POINTER(C.c_int64), C.POINTER(C.c_int64)],
    "femo_dist_overlap_info": [_P, _D],
    "femo_mg_create": [C.c_int, C.POINTER(_P), C.POINTER(_P)],
    "femo_mg_destroy": [_P],
    "femo_mg_set_transfer": [_P, C.c_int, _P, _P, _P, _P, _P, _P],
    "femo_mg_setup": [_P, C.c_int, C.c_int],
    "femo_mg_get_lambda": [_P, C.c_int, _D],
    "femo_mg_set_refine": [_P, C.c_int, C.c_double],
    "femo_mg_solve": [_P, _P, _P, C.c_double, C.c_int, C.POINTER(C.c_int), _D, C.c_int, C.c_int],
    "femo_direct_setup": [_P],
    "femo_direct_solve": [_P, _P, _P, C.c_int, _D, C.c_int, C.c_int],
    "femo_direct_info": [_P, _D],
    "femo_dyn_create": [_P, C.c_double, C.c_int, C.c_int, C.c_double, C.POINTER(_P)],
    "femo_dyn_destroy": [_P],
    "femo_dyn_set_mg": [_P, _P],
    "femo_dyn_setup": [_P, C.c_int],
    "femo_dyn_solve": [_P, _P, C.c_int64, C.c_double, C.c_int, C.POINTER(C.c_int64), C.c_int],
    "femo_dyn_get_history": [_P, _P, C.c_int],
    "femo_dyn_set_history": [_P, _P, C.c_int],
    "femo_dyn_get_energies": [_P, _P],
    "femo_dyn_solve_adjoint": [_P, _P, _P, C.c_double, C.c_int, C.POINTER(C.c_int64), C.c_int],
    "femo_dyn_solve_forward": [_P, _P, _P, C.c_double, C.c_int, C.POINTER(C.c_int64), C.c_int],
    "femo_dyn_dRdw_apply": [_P, _P, _P, C.c_int, C.c_int],
    "femo_dyn_dRdh_T_apply": [_P, _P, _P, C.c_int],
    "femo_dyn_dRdh_apply": [_P, _P, _P, C.c_int],
    "femo_dyn_total_strain_energy": [_P, _D],
    "femo_dyn_total_strain_energy_grad": [_P, C.c_int, _P, C.c_int],
    "femo_dyn_mass_apply": [_P, _P, _P, C.c_int],
    "femo_dyn_info": [_P, _D, _D, C.POINTER(C.c_int)],
    "femo_hex_setup": [_P, C.c_int64, _P, C.c_int64, _P, C.c_int, C.c_double, C.c_double, C.c_double],
    "femo_hex_set_input": [_P, C.c_int, _P, C.c_int64, C.c_int],
    "femo_hex_get_input": [_P, C.c_int, _P, C.c_int64, C.c_int],
    "femo_hex_dRdf_T_apply": [_P, C.c_int, _P, _P, C.c_int],
    "femo_hex_dRdf_apply": [_P, C.c_int, _P, _P, C.c_int],
    "femo_hex_output": [_P, C.c_int, _D],
    "femo_hex_output_grad": [_P, C.c_int, C.c_int, _P, C.c_int],
    "femo_hex_get_mesh": [_P, _P, _P],
    "femo_launch_count": [_P, C.POINTER(C.c_int64)],
    "femo_last_timings": [_P, _D],
    "femo_pcg_fixed": [_P, C.c_int, _D],
    "femo_measure_peaks": [_P, _D, _D],
    "femo_last_dmma_tflops": [_P, _D],
}

_lib = None


def load():
    """Load the shared library (fails loudly when it has not been built)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "femo_alpha_b200: %s is missing -- build it with `python -m femo_alpha_b200.build` "
            "(nvcc, sm_100a). There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    lib.femo_last_error.restype = C.c_char_p
    lib.femo_last_error.argtypes = []
    lib.femo_version.restype = C.c_int
    lib.femo_version.argtypes = []
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = C.c_int
    _lib = lib
    return lib


def check(status):
    if status != 0:
        raise RuntimeError("libfemo_b200: " + load().femo_last_error().decode("utf-8", "replace"))


def ptr(a):
    """Raw pointer of a C-contiguous numpy array, a torch CUDA tensor, or None."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    raise TypeError("expected numpy array or torch tensor")


def on_device(a):
    return 1 if (a is not None and not isinstance(a, np.ndarray) and hasattr(a, "is_cuda") and a.is_cuda) else 0
