"""torchrun worker: femo_halo_accumulate (reverse halo exchange, csrc/dist.cu) on N GPUs.  x = 1 on every local block:
after the call an owned entry holds the number of ranks that keep a copy of its block, which the partition tables give
independently; x = a global field: owner + sum of ghost copies = copies * field."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from femo_alpha_b200.dist_shell import DistributedShell
    from tests.problems import make_problem
    p = make_problem("plate_4x20")
    ds = DistributedShell(p["nodes"], p["quads"], rank, world, device=local)
    ds.init_comm()
    lp = ds.lp
    # copies of every block over the ranks, from the partition tables (all-gathered keys)
    n_keys = 2 * lp.n_global_vertices + lp.n_global_edges + lp.n_global_cells          # same on every rank
    keys = torch.zeros(n_keys, dtype=torch.int64, device="cuda")
    keys[torch.from_numpy(lp.block_gkey).cuda()] += 1
    dist.all_reduce(keys)
    copies = keys.cpu().numpy()[lp.block_gkey].astype(np.float64)
    own = np.repeat(lp.own_block_mask, 3)
    x = ds.halo_accumulate(np.ones(ds.dev.ndof))
    e1 = np.abs(x[own] - np.repeat(copies, 3)[own]).max()
    field = np.cos(0.37 * lp.block_gkey.repeat(3) + np.tile(np.arange(3), lp.block_gkey.size))      # same value on every copy
    y = ds.halo_accumulate(field.copy())
    e2 = np.abs(y[own] - (np.repeat(copies, 3) * field)[own]).max()
    e3 = np.abs(y[~own] - field[~own]).max() if (~own).any() else 0.0                                # ghosts untouched
    print("rank %d/%d halo_accumulate: counts err %.1e | field err %.1e | ghosts changed by %.1e | max copies %d"
          % (rank, world, e1, e2, e3, int(copies.max())), flush=True)
    ok = e1 == 0.0 and e2 < 1e-14 and e3 == 0.0 and copies.max() >= 2
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ds.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
