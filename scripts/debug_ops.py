import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import h2b200
from h2b200 import fields as F
from h2b200.plonk import circuit as C, polyexpr as PX
from h2b200.plonk.ops import DeviceOps
k = int(sys.argv[1]); j = 6
ctx = h2b200.Context([0]); engine = h2b200.B200Engine(ctx)
dom = h2b200.EvaluationDomain(j, k, ctx)
ops = DeviceOps(None, engine, dom)
n = 1 << k; r = F.R_MOD
p = ops.upload(h2b200.synth.random_scalars(n, 5))
q = ops.upload(h2b200.synth.random_scalars(n, 6))
idx = [0, 1, 2, 3, 5, n // 2 - 1, n // 2, n - 2, n - 1, 151551, 151552, 151553, 200000]
idx = [i for i in idx if i < n]
pi = ops.to_ints(p); qi = ops.to_ints(q)
# pointwise at n rows
prod = ops.to_ints(PX.pointwise(ops, C.mul(C.add(PX.col(0), C.const(7)), PX.col(1)), [p, q]))
bad = [i for i in range(n) if prod[i] != (pi[i] + 7) * qi[i] % r]
print("pointwise n rows bad:", len(bad), bad[:5], flush=True)
rot = ops.to_ints(PX.pointwise(ops, C.sub(PX.col(0, 1), PX.col(1, -1)), [p, q]))
bad = [i for i in range(n) if rot[i] != (pi[(i + 1) % n] - qi[(i - 1) % n]) % r]
print("rotations n rows bad:", len(bad), bad[:5], flush=True)
# horner
hz = ops.to_ints(PX.horner(ops, [p, q, p], 11))
bad = [i for i in range(n) if hz[i] != ((pi[i] * 11 + qi[i]) * 11 + pi[i]) % r]
print("horner bad:", len(bad), flush=True)
# grand product
gp = ops.to_ints(ops.grand_product(p, 3, n))
acc = 3; badc = 0
for i in range(n):
    if gp[i] != acc: badc += 1
    acc = acc * pi[i] % r
print("grand_product bad:", badc, flush=True)
inv = ops.to_ints(ops.batch_invert(ops.clone(p)))
bad = [i for i in range(0, n, 997) if inv[i] * pi[i] % r != 1]
print("batch_invert bad (sampled):", len(bad), flush=True)
# canonical round trip
can = ops.to_canonical(p, n - 6)
back = ops.to_ints(ops.from_canonical(can, n))
print("canonical round trip:", back[:n - 6] == pi[:n - 6], back[n - 6:] == [0] * 6, flush=True)
# coset extension vs direct evaluation
ext = ops.coeff_to_extended(p)
en = dom.extended_len()
for i in [0, 1, 7, en // 3, en - 1]:
    pt = F.FR_ZETA * pow(dom.extended_omega, i, r) % r
    print("ext", i, ops.read_ints(ext, i, 1)[0] == ops.eval_polynomial(p, pt), flush=True)
back = ops.extended_to_coeff(ext)
bi = ops.to_ints(back)
print("extended_to_coeff round trip:", bi[:n] == pi, all(v == 0 for v in bi[n:]), len(bi), flush=True)
lag = ops.lagrange_to_coeff(ops.clone(p))
for i in [0, 1, n - 1]:
    print("lagrange", i, ops.eval_polynomial(lag, pow(dom.omega, i, r)) == pi[i], flush=True)
# kate division
b = 12345
qd = ops.kate_division(p, b)
z = 98765
lhs = ops.eval_polynomial(p, z); rem = ops.eval_polynomial(p, b)
print("kate:", (ops.eval_polynomial(qd, z) * (z - b) + rem) % r == lhs, flush=True)
