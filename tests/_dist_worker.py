"""world_size-2 worker (gloo, CPU): the multi-process MSM plumbing -- shard by point range, all-gather the per-window sums, fold
with h2b_msm_combine (the path for callers without peer access) -- with the per-shard window sums produced by the host-emulated
kernels; and the handle exchange that sets up h2b_msm_sharded (the peer-memory path bench.py uses)."""
import ctypes
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
    hc = ctypes.CDLL(os.path.join(ROOT, "yet-another-halo2-fork_b200", "libh2b200_hostcheck.so"))
    s_all, b_all = C.random_fr(n_total, 77), C.gen_bases(n_total, 3)
    lo, hi = h2b200.shard_range(n_total, rank, world)
    c = h2b200.window_bits(n_total)
    nw = (255 + c - 1) // c
    s, b = np.ascontiguousarray(s_all[lo:hi]), np.ascontiguousarray(b_all[lo:hi])
    win = np.zeros((nw, 16), dtype=np.uint64)
    hc.hc_msm_windows(s.ctypes.data_as(ctypes.c_void_p), b.ctypes.data_as(ctypes.c_void_p), hi - lo, c, win.ctypes.data_as(ctypes.c_void_p))
    mine = torch.from_numpy(win.view(np.int64))
    allw = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allw, mine)
    ok = 1
    if rank == 0:
        parts = np.stack([t.numpy().view(np.uint64) for t in allw])
        out = h2b200.B200Engine.msm_combine(parts, c)
        exp = C.to_affine(C.best_multiexp(s_all, b_all))
        ok = int(np.array_equal(out[:8], exp))
    # the start-up exchange of h2b_msm_sharded: every rank's 64-byte mailbox handle, in rank order (B200Engine.peer_attach)
    from h2b200.engine import dist_all_gather_bytes
    blobs = dist_all_gather_bytes(bytes([rank + 1]) * h2b200._lib.PEER_HANDLE_BYTES)
    if [b for b in blobs] != [bytes([r + 1]) * h2b200._lib.PEER_HANDLE_BYTES for r in range(world)]:
        ok = 0
    flag = torch.tensor([ok])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)        # any rank's failure fails all
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
