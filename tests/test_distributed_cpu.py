"""N > 1 host logic on CPU: two gloo ranks shard an MSM by point range, all-gather window sums, rank 0 combines."""
import os
import subprocess
import sys

import pytest

import h2b200

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_partition():
    for n in (1, 7, 1000, 1 << 24):
        for world in (1, 2, 3, 4, 8):
            rs = [h2b200.shard_range(n, r, world) for r in range(world)]
            assert rs[0][0] == 0 and rs[-1][1] == n
            assert all(rs[i][1] == rs[i + 1][0] for i in range(world - 1))


@pytest.mark.parametrize("n", [5, 1000])
def test_two_rank_gloo_msm(n):
    subprocess.check_call([sys.executable, "-c", "import sys; sys.path.insert(0, %r); import build; build.build_hostcheck()" %
                           os.path.join(ROOT, "yet-another-halo2-fork_b200")])
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "_dist_worker.py"), str(n)]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
