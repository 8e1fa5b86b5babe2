"""Small fixed workload for ncu captures of the polynomial-helper and quotient kernels: one call of each at 2^k rows."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import h2b200
from h2b200 import evaluation as E

k = int(sys.argv[1]) if len(sys.argv) > 1 else 22
n = 1 << k
ctx = h2b200.Context()
ops = h2b200.PolyOps(ctx)
d = torch.from_numpy(h2b200.synth.random_scalars(n, 7).view(np.int64)).cuda()
for _ in range(2):
    ops.batch_invert(d)
    ops.grand_product(d, 1, n)
    ops.eval_polynomial(d, 12345)
    ops.kate_division(d, 12345)
    g = E.GraphEvaluator()
    gate = ("prod", ("query", "fixed", 0, 0), ("sum", ("prod", ("query", "advice", 0, 0), ("query", "advice", 1, 1)),
                                               ("neg", ("query", "advice", 0, -1))))
    g.add_calculation(("horner", (10,), (g.add_expression(gate),), (9,)))
    vals = torch.zeros_like(d)
    g.evaluate(ctx, [d], [d, d], [], [], 3, 5, 7, 11, k - 3, k, previous=vals, out=vals)
    E.lookup_fold(ctx, vals, d, d, d, d, d, d, d, 3, 5, 7, k - 3, k)
    E.permutation_fold(ctx, vals, [d], [d, d, d, d], [d, d, d, d], 4, d, d, d, 3, 5, 7, h2b200.fields.FR_ZETA, 9, 11, 5, k - 3, k)
print("ok", ctx.last_call_ms())
