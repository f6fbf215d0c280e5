"""Element arithmetic of the warped-cell assembly kernel (csrc/assemble.cu: warped_uu from per-point M matrices, warped_ut and
warped_tt in W form) exercised on the CPU: the same __host__ __device__ functions, compiled by nvcc into a host program
(tests/warped_host.cu), against block_general -- the literal row form the GPU parity tests pin to the oracle
(tests/test_gpu_parity.py::test_warped_cells_cell_centric_assembly runs the kernels themselves).  All 91 node pairs of an
element, 1 to 4 columns per task, plus the task lists covering every pair exactly once."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def test_warped_tile_arithmetic_matches_row_form(built_lib):
    exe = os.path.join(HERE, "_build", "warped_host")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    src = os.path.join(HERE, "warped_host.cu")
    csrc = os.path.join(ROOT, "femo_alpha_b200", "csrc")
    libdir = os.path.join(ROOT, "femo_alpha_b200")
    deps = [src, os.path.join(csrc, "assemble.cu"), os.path.join(csrc, "femo_internal.cuh")]
    if not os.path.exists(exe) or any(os.path.getmtime(exe) < os.path.getmtime(d) for d in deps):
        r = subprocess.run(["nvcc", "-std=c++17", "-O1", "-gencode", "arch=compute_100a,code=sm_100a", "-I", csrc, src, "-L", libdir,
                            "-lfemo_b200", "-Xlinker", "-rpath=" + libdir, "-o", exe], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
    r = subprocess.run([exe], capture_output=True, text=True, env=dict(os.environ, LD_LIBRARY_PATH=libdir))
    print(r.stdout)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout + r.stderr
    for nb in (1, 2, 3, 4):
        assert ("NB=%d:" % nb) in r.stdout
