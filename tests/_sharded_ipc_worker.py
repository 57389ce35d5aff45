"""Worker of tests/test_gpu_sharded_ipc.py (not collected): one rank of a block-row sharded operator in its own
process, peers connected through CUDA IPC handles swapped over torch.distributed (gloo).  Writes this rank's rows."""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
import torch
import torch.distributed as dist

from butterfly_b200 import capi, synth
from butterfly_b200.sharded import ShardRankOperator


def main():
    out_dir = Path(sys.argv[1])
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    torch.cuda.set_device(rank)
    raw = synth.random_hodbf(n=3000, dtype=capi.COMPLEX_F64, seed=3)

    def exchange(blob):
        got = [None] * world
        dist.all_gather_object(got, blob)
        return got

    op = ShardRankOperator(raw, world, rank, rank, exchange, fractions=[0.45, 0.55][:world] if world == 2 else None)
    rng = np.random.default_rng(17)  # the operand is replicated: same seed on every rank
    ys = []
    for rep in range(3):
        xh = rng.standard_normal((raw.num_cols, 1)) + 1j * rng.standard_normal((raw.num_cols, 1))
        x = torch.from_numpy(xh).cuda()
        y = torch.full((raw.num_rows, 1), float("nan"), dtype=torch.complex128, device="cuda")
        op.apply(x, y)
        torch.cuda.synchronize()
        r0, r1 = op.y_range
        ys.append(y[r0:r1].cpu().numpy())
    dist.barrier()  # nobody tears its workspace down while a peer may still push into it
    np.savez(out_dir / f"rank{rank}.npz", rows=np.array(op.y_range), exchanged=op.exchanged_rows, **{f"y{i}": v for i, v in enumerate(ys)})
    op.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
