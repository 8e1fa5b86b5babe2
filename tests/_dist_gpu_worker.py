"""world_size-2 worker on GPU(s): h2b_msm_sharded across PROCESSES -- every rank a single-GPU context of its own (device rank mod
the visible GPUs; on a one-GPU box both ranks share it), mailboxes opened through h2b_peer_export's CUDA IPC handles, which travel
over gloo.  Each rank must return the whole result, equal to the CPU port's, on three calls in a row (both mailbox parities),
then again after a second h2b_peer_attach."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import h2b200
from oracle import cref as C


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    n_total = int(sys.argv[1])
    ndev = h2b200.load().h2b_device_count()
    ctx = h2b200.Context([rank % ndev])
    ctx.set_option("peer_timeout_ms", 20000)
    eng = h2b200.B200Engine(ctx)
    lo, hi = h2b200.shard_range(n_total, rank, world)
    C.use_all_host_cores()                  # (torchrun exports OMP_NUM_THREADS=1)
    b_all = C.gen_bases(n_total, 31)
    desc = eng.get_base_descriptor(np.ascontiguousarray(b_all[lo:hi]))
    ok = 1
    for session in range(2):
        eng.peer_attach(rank, world, h2b200.dist_all_gather_bytes)
        dist.barrier()                      # every rank attached (mailboxes cleared) before any rank publishes
        for rep in range(3):
            s_all = C.random_fr(n_total, 900 + 10 * session + rep)
            got = eng.msm_sharded(np.ascontiguousarray(s_all[lo:hi]), desc, n_total)
            exp = C.to_affine(C.best_multiexp(s_all, b_all))
            if not np.array_equal(got[:8], exp):
                print(f"rank {rank}: session {session} call {rep} differs from the oracle", file=sys.stderr)
                ok = 0
        dist.barrier()                      # nobody re-attaches (clears a mailbox) while a peer is still in its last call
    desc.release()
    ctx.close()
    flag = torch.tensor([ok])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
