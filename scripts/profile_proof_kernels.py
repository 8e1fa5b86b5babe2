"""Per-kernel device time of one k-row proof (h2b_profile), next to the wall clock of create_proof and its stages."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import h2b200
from h2b200 import plonk
from h2b200.plonk import example, rng as RNG, arguments as A, prover as PR, shplonk as SH
from h2b200.plonk.ops import DeviceOps
k = int(sys.argv[1]) if len(sys.argv) > 1 else 20
ctx = h2b200.Context([int(x) for x in os.environ.get('H2B_DEVICES','0').split(',')]); engine = h2b200.B200Engine(ctx)
circ = example.MyCircuit(k, 42); cc = circ.compile(); cs = plonk.circuit.ConstraintSystemBack(cc.cs)
ops = DeviceOps(h2b200.ParamsKZG.setup(k, 0x1234567, engine), engine, h2b200.EvaluationDomain(cs.degree(), k, ctx))
vk = plonk.keygen_vk(ops, cc); pk = plonk.keygen_pk(ops, vk, cc)
stages = {}
def wrap(mod, name):
    f = getattr(mod, name)
    def g(*a, **kw):
        torch.cuda.synchronize(); t = time.perf_counter(); r = f(*a, **kw); torch.cuda.synchronize()
        stages[name] = stages.get(name, 0) + (time.perf_counter() - t) * 1e3; return r
    setattr(mod, name, g)
for n in ["lookup_commit_permuted", "permutation_commit", "lookup_commit_product", "shuffle_commit_product", "vanishing_commit",
          "vanishing_construct", "vanishing_evaluate", "permutation_pk_evaluate", "permutation_evaluate", "lookup_evaluate", "shuffle_evaluate"]:
    wrap(A, n)
wrap(PR, "evaluate_h"); wrap(SH, "create_proof")
orig_commit_phase = PR.ProverSingle.commit_phase
def cp(self, phase, w):
    torch.cuda.synchronize(); t = time.perf_counter(); r = orig_commit_phase(self, phase, w); torch.cuda.synchronize()
    stages[f"commit_phase{phase}"] = (time.perf_counter() - t) * 1e3; return r
PR.ProverSingle.commit_phase = cp
def run():
    tr = plonk.Blake2bWrite()
    plonk.create_proof(ops, pk, circ.instances(), circ.witness, RNG.OneNg(), tr, num_threads=16)
    torch.cuda.synchronize()
run(); stages.clear()
t = time.perf_counter(); run(); wall = (time.perf_counter() - t) * 1e3
print("wall ms", round(wall, 1)); print(json.dumps({k_: round(v, 2) for k_, v in stages.items()}, indent=0))
print("sum of stages", round(sum(stages.values()), 1))
ctx.profile_enable(True); stages.clear(); run(); prof = ctx.profile_read(); ctx.profile_enable(False)
tot = sum(v["ms"] for v in prof.values())
print("kernel ms total", round(tot, 1))
for name, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])[:25]:
    print(f"{name:28s} {v['ms']:8.2f} ms {v['launches']:5d}")
