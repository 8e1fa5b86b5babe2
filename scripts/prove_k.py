"""One real proof of the frontend_backend_split circuit at 2^k rows on the GPU, with a per-stage wall-clock breakdown.
usage: python scripts/prove_k.py [k] [repeats]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import h2b200
from h2b200 import plonk
from h2b200.plonk import example, rng as RNG
from h2b200.plonk.ops import DeviceOps

k = int(sys.argv[1]) if len(sys.argv) > 1 else 20
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
SECRET = 0x1234567
devices = [int(x) for x in os.environ.get('H2B_DEVICES', '0').split(',')]
ctx = h2b200.Context(devices)
out_devices = devices
engine = h2b200.B200Engine(ctx)
t = time.time()
circ = example.MyCircuit(k, 42)
cc = circ.compile()
cs = plonk.circuit.ConstraintSystemBack(cc.cs)
print(f"circuit: {circ.units} units, {len(cc.preprocessing.copies)} copies, {time.time() - t:.2f} s", flush=True)
t = time.time()
params = h2b200.ParamsKZG.setup(k, SECRET, engine)
torch.cuda.synchronize()
print(f"ParamsKZG.setup: {time.time() - t:.2f} s", flush=True)
ops = DeviceOps(params, engine, h2b200.EvaluationDomain(cs.degree(), k, ctx))
t = time.time()
vk = plonk.keygen_vk(ops, cc)
pk = plonk.keygen_pk(ops, vk, cc)
torch.cuda.synchronize()
print(f"keygen: {time.time() - t:.2f} s", flush=True)
out = {"k": k, "devices": out_devices, "runs": []}
for rep in range(reps):
    tr = plonk.Blake2bWrite()
    l0 = ctx.launch_count()
    torch.cuda.synchronize()
    t = time.time()
    plonk.create_proof(ops, pk, circ.instances(), circ.witness, RNG.OneNg(), tr, num_threads=16)
    torch.cuda.synchronize()
    dt = time.time() - t
    proof = tr.finalize()
    out["runs"].append({"ms": dt * 1e3, "launches": ctx.launch_count() - l0, "proof_bytes": len(proof)})
    print(f"create_proof: {dt * 1e3:.1f} ms, {ctx.launch_count() - l0} launches, {len(proof)} bytes", flush=True)
if os.environ.get("H2B_VERIFY", "1") == "1":
    from oracle import plonk_verifier
    t = time.time()
    ok = plonk_verifier.verify_proof(vk, circ.instances(), proof, SECRET, 1 << k)
    print(f"verify (python restatement): {ok}, {time.time() - t:.2f} s")
    out["verified"] = bool(ok)
free, total = torch.cuda.mem_get_info()
out["torch_peak_gib"] = torch.cuda.max_memory_allocated() / 2**30
out["device0_used_gib"] = (total - free) / 2**30
print(json.dumps(out))
