"""Small fixed workload for ncu captures: one MSM and one NTT (sizes from argv: msm_k ntt_k), H2B_PROF_REPS (default 2) repetitions."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import h2b200
from oracle import cref as C

mk = int(sys.argv[1]) if len(sys.argv) > 1 else 20
nk = int(sys.argv[2]) if len(sys.argv) > 2 else 22
reps = int(os.environ.get("H2B_PROF_REPS", "2"))
ctx = h2b200.Context()
eng = h2b200.B200Engine(ctx)
if mk:
    n = 1 << mk
    s = torch.from_numpy(C.random_fr(n, 1).view(np.int64)).cuda()
    desc = eng.get_base_descriptor(C.gen_bases(n, 1))
    for _ in range(reps):
        out = eng.msm_with_cached_base(s, desc)
    print("msm", mk, ctx.last_call_ms())
if nk:
    dom = h2b200.EvaluationDomain(1, nk, ctx)
    d = torch.from_numpy(C.random_fr(1 << nk, 2).view(np.int64)).cuda()
    for _ in range(reps):
        dom.coeff_to_lagrange(d)
    print("ntt", nk, ctx.last_call_ms())
