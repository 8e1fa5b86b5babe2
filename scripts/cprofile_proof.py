"""Host-side profile (cProfile, self time) of one k-row create_proof: where the wall clock goes outside the library calls."""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import h2b200
from h2b200 import plonk
from h2b200.plonk import example, rng as RNG
from h2b200.plonk.ops import DeviceOps
k = int(sys.argv[1]) if len(sys.argv) > 1 else 20
ctx = h2b200.Context([0]); engine = h2b200.B200Engine(ctx)
circ = example.MyCircuit(k, 42); cc = circ.compile(); cs = plonk.circuit.ConstraintSystemBack(cc.cs)
ops = DeviceOps(h2b200.ParamsKZG.setup(k, 0x1234567, engine), engine, h2b200.EvaluationDomain(cs.degree(), k, ctx))
vk = plonk.keygen_vk(ops, cc); pk = plonk.keygen_pk(ops, vk, cc)
def run():
    tr = plonk.Blake2bWrite()
    plonk.create_proof(ops, pk, circ.instances(), circ.witness, RNG.OneNg(), tr, num_threads=16)
    torch.cuda.synchronize()
run(); run()
t = time.perf_counter(); run(); print("wall ms", round((time.perf_counter() - t) * 1e3, 1))
pr = cProfile.Profile(); pr.enable(); run(); pr.disable()
st = pstats.Stats(pr); st.sort_stats("tottime").print_stats(45)
