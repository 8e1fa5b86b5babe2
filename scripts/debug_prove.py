import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import h2b200
from h2b200 import plonk, fields as F
from h2b200.plonk import example, rng as RNG, arguments as A, prover as PR
from h2b200.plonk.ops import DeviceOps
from oracle import plonk_verifier, pyref as R
k = int(sys.argv[1]); SECRET = 0x1234567
ctx = h2b200.Context([0]); engine = h2b200.B200Engine(ctx)
circ = example.MyCircuit(k, 42); cc = circ.compile(); cs = plonk.circuit.ConstraintSystemBack(cc.cs)
params = h2b200.ParamsKZG.setup(k, SECRET, engine)
ops = DeviceOps(params, engine, h2b200.EvaluationDomain(cs.degree(), k, ctx))
# SRS sanity: commit(p) == [p(s)]G for a random p
p = ops.upload(h2b200.synth.random_scalars(1 << k, 5))
exp = R.g1_mul(R.G1_GEN, ops.eval_polynomial(p, SECRET))
print("commit == [p(s)]G:", ops.commit(p) == exp, flush=True)
lag = ops.lagrange_to_coeff(ops.clone(p))
print("commit_lagrange == commit(iNTT):", ops.commit_lagrange(p) == ops.commit(lag), flush=True)
vk = plonk.keygen_vk(ops, cc); pk = plonk.keygen_pk(ops, vk, cc)
cap = {}
orig = A.vanishing_evaluate
def hook(ops_, c, x, xn, tr):
    e = orig(ops_, c, x, xn, tr); cap['h_poly'] = e.h_poly; cap['x'] = x; return e
A.vanishing_evaluate = hook
orig_h = PR.evaluate_h
def hook_h(*a, **kw):
    v = orig_h(*a, **kw); cap['h_ext'] = v.clone(); return v
PR.evaluate_h = hook_h
tr = plonk.Blake2bWrite()
plonk.create_proof(ops, pk, circ.instances(), circ.witness, RNG.OneNg(), tr, num_threads=16)
proof = tr.finalize()
dbg = {}
ok = plonk_verifier.verify_proof(vk, circ.instances(), proof, SECRET, 1 << k, dbg)
h_at_x = ops.eval_polynomial(cap['h_poly'], cap['x'])
print("verify:", ok, " x match:", cap['x'] == dbg['x'], " h(x) == expected:", h_at_x == dbg['expected_h_eval'], flush=True)
# which expression is non-zero on the domain?  evaluate each constraint's contribution: quotient must be a polynomial, i.e. numerator vanishes on H.
# check numerator on the extended coset is consistent: h_ext / Z_H back to coeffs has degree < n*(d-1): compare full inverse transform tail
dom = ops.domain
import ctypes
full = torch.empty((dom.extended_len(), 4), dtype=torch.int64, device='cuda')
