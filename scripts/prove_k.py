"""One real proof of the frontend_backend_split circuit at 2^k rows on the GPU, with a per-stage wall-clock breakdown.
usage: python scripts/prove_k.py [k] [repeats]      H2B_DEVICES=0,1,2,3 for a multi-device context;
H2B_COSETS=1: evaluate the quotient coset by coset over those devices (plonk/cosets.py) with a key that holds no extended vector."""
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
use_cosets = os.environ.get("H2B_COSETS", "0") == "1"
pk = plonk.keygen_pk(ops, vk, cc, extended=not use_cosets)
if use_cosets:
    from h2b200.plonk.cosets import CosetProver
    ops.cosets = CosetProver(ops, devices)
torch.cuda.synchronize()
print(f"keygen: {time.time() - t:.2f} s", flush=True)
out = {"k": k, "devices": out_devices, "cosets": use_cosets, "runs": []}
stage_ms = {}
if os.environ.get("H2B_STAGES", "0") == "1":          # wall-clock of the quotient stage inside create_proof (adds synchronisations)
    from h2b200.plonk import prover as _prover

    def _timed(fn, name):
        def wrapper(*a, **kw):
            for dev in sorted(set(devices)):
                torch.cuda.synchronize(dev)
            t0 = time.time()
            r = fn(*a, **kw)
            for dev in sorted(set(devices)):
                torch.cuda.synchronize(dev)
            stage_ms.setdefault(name, []).append((time.time() - t0) * 1e3)
            return r
        return wrapper
    _prover.evaluate_h = _timed(_prover.evaluate_h, "evaluate_h")
    for _name in ("lookup_commit_permuted", "permutation_commit", "lookup_commit_product", "shuffle_commit_product", "vanishing_commit",
                  "vanishing_construct", "vanishing_construct_pieces", "vanishing_evaluate"):
        setattr(_prover.A, _name, _timed(getattr(_prover.A, _name), _name))
    _prover.shplonk.create_proof = _timed(_prover.shplonk.create_proof, "multiopen")
    _prover.Prover.commit_phase = _timed(_prover.Prover.commit_phase, "commit_phase")
    if use_cosets:
        ops.cosets.evaluate_h_pieces = _timed(ops.cosets.evaluate_h_pieces, "evaluate_h_pieces")
        ops.cosets._run_worker = _timed(ops.cosets._run_worker, "worker")
kprof = os.environ.get("H2B_KPROF", "0") == "1"     # per-kernel-family device time of the LAST proof (main context only)
for rep in range(reps):
    if kprof and rep == reps - 1:
        ctx.profile_enable(True)
    tr = plonk.Blake2bWrite()
    l0 = ctx.launch_count()
    torch.cuda.synchronize()
    t = time.time()
    if os.environ.get("H2B_CPROFILE", "0") == "1" and rep == reps - 1:      # host-side call tree of the last proof
        import cProfile
        import pstats
        prof = cProfile.Profile()
        prof.runcall(plonk.create_proof, ops, pk, circ.instances(), circ.witness, RNG.OneNg(), tr, num_threads=16)
        torch.cuda.synchronize()
        dt = time.time() - t
        pstats.Stats(prof).sort_stats("cumulative").print_stats(45)
    else:
        plonk.create_proof(ops, pk, circ.instances(), circ.witness, RNG.OneNg(), tr, num_threads=16)
        torch.cuda.synchronize()
        dt = time.time() - t
    proof = tr.finalize()
    import hashlib
    out["proof_sha256"] = hashlib.sha256(proof).hexdigest()
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
out["per_device_used_gib"] = []
for dev in sorted(set(devices)):
    f_, t_ = torch.cuda.mem_get_info(dev)
    out["per_device_used_gib"].append(round((t_ - f_) / 2**30, 2))
if kprof:
    out["kernel_ms"] = {k_: round(v["ms"], 2) for k_, v in sorted(ctx.profile_read().items(), key=lambda kv: -kv[1]["ms"])}
out["stage_ms"] = {k_: [round(v, 1) for v in vs[-4:]] for k_, vs in stage_ms.items()}
print(json.dumps(out))
if ops.cosets:
    ops.cosets.close()
    ops.cosets = None
