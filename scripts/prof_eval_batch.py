"""ncu target: the batched evaluation of a k = 20 proof's evaluation phase (41 pairs over 12 polynomials of 2^20 coefficients)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import h2b200

k = int(sys.argv[1]) if len(sys.argv) > 1 else 20
ctx = h2b200.Context()
ops = h2b200.PolyOps(ctx)
polys = [torch.from_numpy(h2b200.synth.random_scalars(1 << k, 7 + i).view(np.int64)).cuda() for i in range(12)]
pairs = [(polys[i % 12], 1000003 + i) for i in range(41)]
for _ in range(3):
    vals = ops.eval_polynomials(pairs)
print("ok", ctx.last_call_ms(), hex(vals[0])[:18])
