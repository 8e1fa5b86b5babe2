"""Small workload that runs every kernel family once (MSM layouts, NTT pass counts, polynomial helpers, quotient folds) at sizes
that finish in seconds: a quick all-kernels smoke run (compute-sanitizer is closed on this GPU pool, so it is run plain)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import h2b200
from h2b200 import evaluation as E

ctx = h2b200.Context([0])
eng = h2b200.B200Engine(ctx)
ops = h2b200.PolyOps(ctx)
# MSM: classic layout (n = 3000), precomputed windows + long bucket + batch (n = 2^16 + 5)
for n in (3000, (1 << 16) + 5):
    b = h2b200.synth.device_bases(ctx, n, 1)
    desc = eng.get_base_descriptor(b)
    s = h2b200.synth.random_scalars(n, 3)
    s[::2] = h2b200.fields.fr_mont(1)
    eng.msm_with_cached_base(s, desc)
    eng.msm_batch([s, s[::-1].copy()], desc)
    desc.release()
# NTT: one-, two- and three-pass sizes, coset transforms
for k in (5, 12, 17):
    dom = h2b200.EvaluationDomain(5, k, ctx)
    a = h2b200.synth.random_scalars(1 << k, 4)
    c = dom.lagrange_to_coeff(a.copy())
    e = dom.coeff_to_extended(c)
    dom.extended_to_coeff(e, divide_by_vanishing=True)
# polynomial helpers and quotient folds
n = (1 << 13) + 77
d = torch.from_numpy(h2b200.synth.random_scalars(n, 7).view(np.int64)).cuda()
ops.batch_invert(d.clone()); ops.grand_product(d, 1, n); ops.eval_polynomial(d, 5); ops.kate_division(d, 5)
k = 12
d = torch.from_numpy(h2b200.synth.random_scalars(1 << k, 8).view(np.int64)).cuda()
g = E.GraphEvaluator()
gate = ("prod", ("query", "fixed", 0, 0), ("sum", ("prod", ("query", "advice", 0, 0), ("query", "advice", 1, 1)), ("neg", ("query", "advice", 0, -1))))
g.add_calculation(("horner", (10,), (g.add_expression(gate),), (9,)))
vals = torch.zeros_like(d)
g.evaluate(ctx, [d], [d, d], [], [], 3, 5, 7, 11, k - 2, k, previous=vals, out=vals)
E.lookup_fold(ctx, vals, d, d, d, d, d, d, d, 3, 5, 7, k - 2, k)
E.shuffle_fold(ctx, vals, d, d, d, d, d, d, 7, k - 2, k)
E.permutation_fold(ctx, vals, [d, d], [d, d, d], [d, d, d], 2, d, d, d, 3, 5, 7, h2b200.fields.FR_ZETA, 9, 11, 5, k - 2, k)
torch.cuda.synchronize()
print("sanitize target ok")
