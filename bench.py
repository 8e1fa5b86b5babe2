#!/usr/bin/env python
"""bench.py -- BN254 MSM 2^24 (headline) and Fr NTT 2^24 on B200, with the CPU restatement timed beside it.

    python bench.py --gpus N --steps K --warmup W            this repo's CUDA path (one process per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...  the reference's CPU algorithm (oracle port, host cores)

A "step" is one 2^24-point MSM over resident bases (BASELINE.json configs[1]).  With N > 1 GPUs the MSM is sharded
by contiguous point range (strong scaling): every rank reduces its own range, the last kernel of the reduction stores the
rank's reduced terms (20 points) into every peer's mailbox over NVLink, waits for theirs and folds them (h2b_msm_sharded:
no NCCL call and no host staging on the data path).  The inputs are the SAME 2^24 scalars / bases at every N (ChaCha20
seeds of SURVEY.md 8d, sliced by rank), so `result_sha256` must not change with N.  One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MSM_K = int(os.environ.get("H2B_BENCH_MSM_K", "24"))
NTT_K = int(os.environ.get("H2B_BENCH_NTT_K", "24"))
CPU_BUDGET_S = float(os.environ.get("H2B_BENCH_CPU_BUDGET_S", "240"))     # wall-clock cap of the reference arm's timed loop
NTT_BATCH_COLS = int(os.environ.get("H2B_BENCH_NTT_COLS", "16"))
NTT_BATCH_K = int(os.environ.get("H2B_BENCH_NTT_BATCH_K", "22"))          # columns of 2^22 coefficients -> 2^24 evaluations (e = 4)
# SURVEY.md §8(d) algorithmic work: XYZZ mixed additions for the best window size, 10 field multiplications each,
# 128 IMAD.WIDE (32x32->64 multiply-add) per Montgomery multiplication (64 product + 64 reduction).
MSM_ADDS = {16: 1.475e6, 18: 5.014e6, 20: 1.769e7, 22: 6.488e7, 24: 2.317e8, 26: 8.556e8}
WIDE_PER_FIELD_MUL = 128
MULS_PER_MADD = 10


def workload_config(world: int) -> dict:
    return {"workload": f"configs[1]: best_multiexp BN254 G1 MSM, 2^{MSM_K} points (bases [t0+i]G, t0 from ChaCha20 seed [2;32]; scalars = "
                        f"Fr::random draws of ChaCha20 seed [1;32]; SURVEY.md 8d), sharded by point range over {world} GPU(s)",
            "window_bits": int(os.environ.get("H2B_MSM_C", "0")) or "auto",
            "l2": "inputs (1.5 GiB) exceed the 126 MB L2; no flush needed"}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


TRAFFIC_FILE = "r2_full_size_dram_traffic.json"


def ncu_traffic(kernel_prefix: str, size_k: int):
    """DRAM bytes per launch (read + write) from the committed `ncu` capture of the same kernel at the same size
    (profiles/r2_full_size_dram_traffic.json, taken at 2^24 on this round's code: scripts/capture_traffic.sh); None when the
    size differs or the capture is absent.  Returns (bytes, source)."""
    path = os.path.join(ROOT, "profiles", TRAFFIC_FILE)
    if size_k != 24 or not os.path.exists(path):
        return None, None
    doc = json.load(open(path))
    rows = [r for r in doc["kernels"] if r["kernel"].startswith(kernel_prefix)]
    if not rows:
        return None, None
    return float(np.mean([r["dram_read_bytes"] + r["dram_write_bytes"] for r in rows])), f"profiles/{TRAFFIC_FILE} (commit {doc.get('commit', '?')})"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------
# CPU restatement (oracle port), all host threads, at the REAL size of the metric (no extrapolation)
# ---------------------------------------------------------------------------------------------------------
_CPU_INPUTS = {}


def cpu_inputs(k: int):
    """the same seeded inputs as the GPU arm, generated by the oracle's own ChaCha20 (cached per size)"""
    from oracle import cref as C
    if k not in _CPU_INPUTS:
        C.use_all_host_cores()                   # torchrun exports OMP_NUM_THREADS=1 to its ranks
        n = 1 << k
        _CPU_INPUTS[k] = (C.random_fr_chacha20(n, C.SEED_SCALARS), C.gen_bases(n, C.bench_t0()))
    return _CPU_INPUTS[k]


def cpu_msm(k: int):
    """One best_multiexp restatement (oracle/h2ref.c) on 2^k points; returns (ms, threads, affine result)."""
    from oracle import cref as C
    s, b = cpu_inputs(k)
    t = time.perf_counter()
    r = C.best_multiexp(s, b)
    ms = (time.perf_counter() - t) * 1e3
    return ms, C.num_threads(), C.to_affine(r)


def cpu_ntt(k: int):
    from oracle import cref as C
    from oracle import pyref as R
    C.use_all_host_cores()
    a = C.random_fr_chacha20(1 << k, C.SEED_SCALARS)
    w = C.mont_from_int(pow(R.ROOT_OF_UNITY, 1 << (28 - k), R.R_MOD), R.R_MOD)
    lib, p = C.lib(), a.ctypes.data_as(__import__("ctypes").c_void_p)
    t = time.perf_counter()
    lib.h2ref_best_fft(p, w.ctypes.data_as(__import__("ctypes").c_void_p), k)      # in place: no copy inside the timed region
    return (time.perf_counter() - t) * 1e3


def point_sha256(affine_or_jac) -> str:
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(affine_or_jac, dtype=np.uint64)[:8].tobytes()).hexdigest()


# ---------------------------------------------------------------------------------------------------------
# k = 20 proof-shaped kernel sequence (BASELINE.json configs[3]): the MSM / NTT calls ONE create_proof makes for
# the frontend_backend_split circuit at k = 20 (9 commit_lagrange + 8 commit MSMs of 2^20, 10 iNTTs of
# 2^20, 10 coset NTTs 2^20 -> 2^23, 1 inverse coset NTT of 2^23), issued through the reference-facing interface with
# HOST buffers.  It is the accelerated part of the prover, not a proof: witness synthesis, evaluate_h and the transcript
# stay on the CPU in the reference integration and are not timed here.
# ---------------------------------------------------------------------------------------------------------
SEQ_K, SEQ_J = 20, 6
SEQ_COUNTS = {"commit_lagrange_uniform": 4, "commit_lagrange_skewed": 5, "commit": 8, "lagrange_to_coeff": 10,
              "coeff_to_extended": 10, "extended_to_coeff": 1}


def skew(scalars: np.ndarray, one: np.ndarray, seed: int) -> np.ndarray:
    """selector-like column: 45 % zeros, 45 % ones, 10 % uniform (SURVEY.md 8d non-uniform set (i))."""
    rng = np.random.default_rng(seed)
    u = rng.random(scalars.shape[0])
    out = scalars.copy()
    out[u < 0.45] = 0
    out[(u >= 0.45) & (u < 0.9)] = one
    return out


def proof_sequence_gpu(ctx, eng):
    import h2b200
    n = 1 << SEQ_K
    dom = h2b200.EvaluationDomain(SEQ_J, SEQ_K, ctx)
    g = eng.get_base_descriptor(h2b200.synth.device_bases(ctx, n, 1))
    gl = eng.get_base_descriptor(h2b200.synth.device_bases(ctx, n, 1 + n))
    uni = h2b200.synth.random_scalars(n, 5)
    skw = skew(uni, h2b200.fields.fr_mont(1), 6)
    ext = np.empty((dom.extended_len(), 4), dtype=np.uint64)

    def once():
        for _ in range(SEQ_COUNTS["commit_lagrange_uniform"]):
            eng.msm_with_cached_base(uni, gl)
        for _ in range(SEQ_COUNTS["commit_lagrange_skewed"]):
            eng.msm_with_cached_base(skw, gl)
        for _ in range(SEQ_COUNTS["commit"]):
            eng.msm_with_cached_base(uni, g)
        for _ in range(SEQ_COUNTS["lagrange_to_coeff"]):
            dom.lagrange_to_coeff(uni.copy())
        for _ in range(SEQ_COUNTS["coeff_to_extended"]):
            dom.coeff_to_extended(uni, out=ext)
        for _ in range(SEQ_COUNTS["extended_to_coeff"]):
            dom.extended_to_coeff(ext)

    once()
    l0 = ctx.launch_count()
    t = time.perf_counter()
    once()
    ms = (time.perf_counter() - t) * 1e3
    return ms, ctx.launch_count() - l0


# The same prover, as the device-resident integration runs it: columns are uploaded once, every
# commit / transform / quotient fold / opening helper consumes and produces device buffers, and only 96-byte points and 32-byte
# evaluations come back.  Counts follow SURVEY.md 8(d) for the k = 20 frontend_backend_split circuit (A=4, F=8, I=1, L=1, S=1,
# P=1, d=6 -> extended_k = 23): the MSM / NTT calls of SEQ_COUNTS plus 3 batch inversions + 3 running products (permutation,
# lookup, shuffle Z polynomials), the custom-gate / permutation / lookup / shuffle folds over 2^23 rows, and the SHPLONK side
# (32 polynomial evaluations, 2 quotient divisions).  Still not a proof: witness synthesis and the transcript are host work.
RES_COUNTS = {"upload_columns": 6, "batch_invert": 3, "grand_product": 3, "eval_polynomial": 32, "kate_division": 2,
              "graph_rows": 1 << 23, "graph_programs": 3, "permutation_fold": 1, "lookup_fold": 1, "shuffle_fold": 1}


def proof_sequence_resident(ctx, eng):
    import h2b200
    import torch
    from h2b200 import evaluation as E
    n = 1 << SEQ_K
    dom = h2b200.EvaluationDomain(SEQ_J, SEQ_K, ctx)
    ops = h2b200.PolyOps(ctx)
    ext_k = dom.extended_k
    g = eng.get_base_descriptor(h2b200.synth.device_bases(ctx, n, 1))
    gl = eng.get_base_descriptor(h2b200.synth.device_bases(ctx, n, 1 + n))
    uni_h = torch.from_numpy(h2b200.synth.random_scalars(n, 5).view(np.int64)).pin_memory()
    skw_h = torch.from_numpy(skew(uni_h.numpy().view(np.uint64), h2b200.fields.fr_mont(1), 6).view(np.int64)).pin_memory()
    # a representative custom-gate program: three degree-3 gates over 4 advice / 2 fixed columns with rotations, folded with y
    adv = lambda c, r=0: ("query", "advice", c, r)
    fix = lambda c, r=0: ("query", "fixed", c, r)
    gates = [("prod", fix(0), ("sum", ("prod", adv(0), adv(1, 1)), ("neg", adv(2)))),
             ("prod", fix(1), ("sum", ("prod", ("prod", adv(3), adv(3)), adv(0, -1)), ("neg", adv(1)))),
             ("prod", fix(0, 1), ("sum", adv(2, 1), ("neg", ("prod", adv(2), ("const", 3)))))]
    custom = E.GraphEvaluator()
    custom.add_calculation(("horner", (10,), tuple(custom.add_expression(e) for e in gates), (9,)))
    lookup = E.GraphEvaluator()                       # evaluation.rs:233-262 shape with degree-2 inputs
    lc_in = lookup.add_calculation(("horner", (0, 0), (lookup.add_expression(("prod", fix(0), adv(0))),), (8,)))
    lc_tab = lookup.add_calculation(("horner", (0, 0), (lookup.add_expression(("prod", fix(1), adv(1))),), (8,)))
    lookup.add_calculation(("mul", lookup.add_calculation(("add", lc_in, (6,))), lookup.add_calculation(("add", lc_tab, (7,)))))
    shuf = E.GraphEvaluator()
    shuf.add_calculation(("add", shuf.add_calculation(("horner", (0, 0), (shuf.add_expression(("prod", fix(0), adv(2))),), (8,))), (7,)))
    beta, gamma, theta, y = 11, 13, 17, 19
    delta = pow(7, 1 << 28, h2b200.fields.R_MOD)

    def once():
        cols = [uni_h.cuda(non_blocking=True) for _ in range(4)] + [skw_h.cuda(non_blocking=True) for _ in range(2)]   # witness upload
        # commits that the prover issues back to back with no challenge in between go through h2b_msm_batch:
        # the 4 advice columns (plonk/prover.rs:436-440), the 2 permuted lookup columns (lookup/prover.rs:130-135); the three
        # Z-polynomial commits each follow a fresh challenge and stay single calls.
        eng.msm_batch(cols[0:4], gl)
        eng.msm_batch(cols[4:6], gl)
        for i in range(3):
            eng.msm_with_cached_base(cols[4 + i % 2], gl)
        for i in range(RES_COUNTS["batch_invert"]):
            z = ops.grand_product(ops.batch_invert(cols[i].clone()), 1, n)
        coeffs = [dom.lagrange_to_coeff(cols[i % 6].clone()) for i in range(SEQ_COUNTS["lagrange_to_coeff"])]
        eng.msm_batch(coeffs[0:5], g)                      # the h pieces (vanishing/prover.rs:136-140)
        for i in range(3):                                 # random poly, SHPLONK h1, h2: one at a time
            eng.msm_with_cached_base(coeffs[5 + i], g)
        ext = [dom.coeff_to_extended(c) for c in coeffs[: SEQ_COUNTS["coeff_to_extended"]]]
        values = torch.zeros((1 << ext_k, 4), dtype=torch.int64, device="cuda")
        custom.evaluate(ctx, ext[4:6], ext[0:4], [], [], beta, gamma, theta, y, SEQ_K, ext_k, previous=values, out=values)
        E.permutation_fold(ctx, values, [ext[6]], ext[0:4], ext[4:8], 4, ext[8], ext[9], ext[9], beta, gamma, y, h2b200.fields.FR_ZETA,
                           delta, dom.extended_omega, 5, SEQ_K, ext_k)
        table_values = lookup.evaluate(ctx, ext[4:6], ext[0:4], [], [], beta, gamma, theta, y, SEQ_K, ext_k)
        E.lookup_fold(ctx, values, table_values, ext[6], ext[7], ext[5], ext[8], ext[9], ext[9], beta, gamma, y, SEQ_K, ext_k)
        sv = shuf.evaluate(ctx, ext[4:6], ext[0:4], [], [], beta, gamma, theta, y, SEQ_K, ext_k)
        E.shuffle_fold(ctx, values, sv, table_values, ext[6], ext[8], ext[9], ext[9], y, SEQ_K, ext_k)
        h = dom.extended_to_coeff(values, divide_by_vanishing=True)
        ops.eval_polynomials([(coeffs[i % len(coeffs)], 123457 + i) for i in range(RES_COUNTS["eval_polynomial"])])   # one call, as the prover does
        for i in range(RES_COUNTS["kate_division"]):
            ops.kate_division(coeffs[i], 987 + i)
        return h

    once()
    torch.cuda.synchronize()
    l0 = ctx.launch_count()
    t = time.perf_counter()
    once()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t) * 1e3
    return ms, ctx.launch_count() - l0


def k20_proof(ctx, eng, k: int = SEQ_K, reps: int = 3, circuit: str = "MyCircuit", threads: int = 16):
    """A real proof: the reference's frontend_backend_split circuit (gate with rotation, lookup, shuffle, copy constraints,
    second-phase column, instance) at 2^k rows, through the prover mirror (plonk.create_proof = keygen'd pk + ProverSingle +
    SHPLONK + Blake2b transcript) with every vector on the device.  tests/test_gpu_plonk.py checks the same path byte for byte
    against the CPU restatement and with the verify_proof restatement; here it is only timed.  Timed region = create_proof incl.
    witness columns handed over as host arrays (H2D inside) and the proof bytes coming back."""
    import hashlib
    import torch
    import h2b200
    from h2b200 import plonk
    from h2b200.plonk import example, rng as RNG
    from h2b200.plonk.ops import DeviceOps
    t0 = time.perf_counter()
    circ = example.MyCircuit(k, 42) if circuit == "MyCircuit" else example.StandardPlonk(k, 7, 5)
    cc = circ.compile()
    cs = plonk.circuit.ConstraintSystemBack(cc.cs)
    params = h2b200.ParamsKZG.setup(k, 0x1234567, eng)
    ops = DeviceOps(params, eng, h2b200.EvaluationDomain(cs.degree(), k, ctx))
    pk = plonk.keygen_pk(ops, plonk.keygen_vk(ops, cc), cc)
    circ.witness(0, {})                                       # first-phase columns are an input of the prover: laid out once
    setup_s = time.perf_counter() - t0
    times, launches, proof = [], 0, b""
    for rep in range(reps + 1):
        tr = plonk.Blake2bWrite()
        l0 = ctx.launch_count()
        torch.cuda.synchronize()
        t = time.perf_counter()
        plonk.create_proof(ops, pk, circ.instances(), circ.witness, RNG.OneNg(), tr, num_threads=threads)
        torch.cuda.synchronize()
        if rep:
            times.append((time.perf_counter() - t) * 1e3)
        launches = ctx.launch_count() - l0
        proof = tr.finalize()
    d = ops.domain
    # the same proof written by the independent python-int oracle (tests/golden/make_golden_large.py; same circuit, SRS secret, RNG and
    # thread count): the bytes must be identical.  None = no stored proof for this k.
    gold_name = f"frontend_backend_split_k{k}_oneng_t{threads}" if circuit == "MyCircuit" else f"standard_plonk_k{k}_mindeg5_oneng_t{threads}"
    gold_path = os.path.join(ROOT, "tests", "golden", "large", gold_name + ".json")
    matches = (json.load(open(gold_path))["sha256"] == hashlib.sha256(proof).hexdigest()) if os.path.exists(gold_path) else None
    if matches is False:
        raise SystemExit(f"bench: the k = {k} proof differs from the oracle's stored proof ({gold_path})")
    return {"matches_oracle_sha": matches,
            "what": (f"create_proof of the frontend_backend_split circuit at k = {k} (KZG / SHPLONK / Blake2b; degree {cs.degree()}, "
                     f"extended_k = {d.extended_k}, {circ.units} units of 51 rows), host witness columns in, proof bytes out") if circuit == "MyCircuit" else
                    (f"create_proof of the simple example circuit of halo2_proofs/benches/plonk.rs (StandardPlonk, minimum_degree 5) at k = {k} "
                     f"(KZG / SHPLONK / Blake2b; extended_k = {d.extended_k}): BASELINE.json configs[0]; launch / latency bound at this size"),
            "value": min(times), "unit": "ms", "runs_ms": times, "gpu_launches": launches, "proof_bytes": len(proof),
            "proof_sha256": hashlib.sha256(proof).hexdigest(), "rng": f"OneNg, num_threads = {threads}",
            "h2d_bytes_per_proof": 4 * 32 * (1 << k), "d2h_bytes_per_proof": len(proof),
            "setup_keygen_s": setup_s, "verified": "tests/test_gpu_plonk.py (verify_proof restatement, and proof bytes == CPU restatement)"}


def msm_raw_bench(ctx, eng):
    """What an unmodified caller gets: ParamsKZG::commit_lagrange calls plain `engine.msm(&scalars, &bases)` (kzg/commitment.rs:305-317 ->
    zal.rs:58), i.e. h2b_msm_raw: the bases cross PCIe for this call only and no window multiples exist (one bucket set per window)."""
    out = {"what": "h2b_msm_raw (MsmAccel::msm): pageable host scalars AND pageable host bases in, nothing registered or cached; "
                   "best of 3 after one warm-up call", "unit": "ms", "sizes": {}}
    import torch
    import h2b200
    for k in (20, 24):
        n = 1 << k
        b = h2b200.synth.device_bases(ctx, n, h2b200.synth.bench_t0()).cpu().numpy().view(np.uint64)
        s = h2b200.synth.chacha_scalars_device(ctx, n).cpu().numpy().view(np.uint64)
        torch.cuda.empty_cache()
        ts, res = [], None
        for rep in range(4):
            t = time.perf_counter()
            res = eng.msm(s, b)
            if rep:
                ts.append((time.perf_counter() - t) * 1e3)
        out["sizes"][f"2^{k}"] = {"value": min(ts), "runs_ms": ts, "h2d_bytes": 96 * n, "result_sha256": point_sha256(res)}
        gold = json.load(open(os.path.join(ROOT, "tests", "golden", "msm_results.json"))).get(str(k))
        if gold:          # the seeded problem of the headline bench: same committed result, through the uncached path
            out["sizes"][f"2^{k}"]["result_matches_golden_sha"] = point_sha256(res) == gold["affine_sha256"]
            if not out["sizes"][f"2^{k}"]["result_matches_golden_sha"]:
                raise SystemExit(f"bench: h2b_msm_raw at 2^{k} differs from the committed oracle result")
    return out


def multi_gpu_proof(world: int, timeout_s: int = 420):
    """BASELINE.json configs[4] under the driver's eyes: after the ranks of the sharded bench have released their GPUs, rank 0 runs ONE
    process with one context over all N devices (scripts/prove_k.py, H2B_COSETS=1: MSMs sharded by point range inside the library, the
    quotient evaluated coset by coset over the devices, plonk/evaluation.rs:344-361) -- k = 24 on 8 GPUs, 22 on 4, 21 on 2 -- one
    warm-up proof and two timed ones, verified by the verify_proof restatement.  Any failure is reported, never fatal to the line."""
    k = {2: 21, 4: 22, 8: 24}.get(world)
    if k is None:
        return {"skipped": f"no multi-GPU proof size is defined for {world} GPUs"}
    env = dict(os.environ, H2B_DEVICES=",".join(str(i) for i in range(world)), H2B_COSETS="1", H2B_VERIFY="1")
    for var in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "OMP_NUM_THREADS"):
        env.pop(var, None)
    t0 = time.perf_counter()
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "prove_k.py"), str(k), "3"], env=env, capture_output=True,
                           text=True, timeout=timeout_s)
    except subprocess.TimeoutExpired:
        return {"k": k, "skipped": f"scripts/prove_k.py did not finish within {timeout_s} s"}
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    if r.returncode != 0 or not lines:
        return {"k": k, "failed": (r.stderr or r.stdout)[-600:]}
    out = json.loads(lines[-1])
    runs = [x["ms"] for x in out["runs"]]
    return {"what": f"create_proof of the frontend_backend_split circuit at k = {k}, one process, one context over {world} GPUs, quotient by cosets "
                    "(configs[4] shape); host witness columns in, proof bytes out; first run (warm-up: tables, key cosets) excluded",
            "k": k, "devices": out["devices"], "value": min(runs[1:]), "unit": "ms", "runs_ms": runs, "gpu_launches_device0_context": out["runs"][-1]["launches"],
            "proof_sha256": out["proof_sha256"], "verified": out.get("verified"), "per_device_used_gib": out.get("per_device_used_gib"),
            "script_wall_s": time.perf_counter() - t0}


def proof_sequence_cpu():
    """One call of each kind through the oracle port, scaled by SEQ_COUNTS (bounded sample)."""
    from oracle import cref as C
    from oracle import pyref as R
    C.use_all_host_cores()
    n = 1 << SEQ_K
    d = R.EvaluationDomain(SEQ_J, SEQ_K)
    m = lambda x: C.mont_from_int(x, R.R_MOD)
    uni, bases = C.random_fr(n, 5), C.gen_bases(n, 1)
    skw = skew(uni, m(1), 6)
    parts = {}

    def timed(name, fn):
        t = time.perf_counter()
        r = fn()
        parts[name] = (time.perf_counter() - t) * 1e3
        return r

    timed("commit_lagrange_uniform", lambda: C.best_multiexp(uni, bases))
    timed("commit_lagrange_skewed", lambda: C.best_multiexp(skw, bases))
    parts["commit"] = parts["commit_lagrange_uniform"]
    coeff = timed("lagrange_to_coeff", lambda: C.lagrange_to_coeff(uni, SEQ_K, m(d.omega_inv), m(d.ifft_divisor)))
    ext = timed("coeff_to_extended", lambda: C.coeff_to_extended(coeff, SEQ_K, d.extended_k, m(d.extended_omega), m(d.g_coset), m(d.g_coset_inv)))
    timed("extended_to_coeff", lambda: C.extended_to_coeff(ext, d.extended_k, m(d.extended_omega_inv), m(d.extended_ifft_divisor),
                                                            m(d.g_coset), m(d.g_coset_inv), n * (SEQ_J - 1)))
    total = sum(parts[k] * SEQ_COUNTS[k] for k in SEQ_COUNTS)
    return total, parts, C.num_threads()


def run_reference(args, rank: int, world: int):
    """The reference's CPU algorithm for the path (C port of halo2curves 0.6 best_multiexp: the reference is Rust and cannot be
    built here), all host threads, on the SAME 2^MSM_K inputs as the GPU arm.  Every step is one full-size MSM: nothing is
    extrapolated.  If `--steps` full-size runs would exceed CPU_BUDGET_S of wall clock the loop stops early and the line's
    `steps` says how many were timed (`steps_requested` keeps the flag)."""
    if rank != 0:
        return
    from oracle import cref as C
    C.use_all_host_cores()
    t_gen = time.perf_counter()
    cpu_inputs(MSM_K)
    gen_s = time.perf_counter() - t_gen
    if args.warmup > 0:                            # one small warm-up pass (thread pool, page tables); a CPU loop needs no more
        s, b = cpu_inputs(MSM_K)
        C.best_multiexp(np.ascontiguousarray(s[: 1 << 16]), np.ascontiguousarray(b[: 1 << 16]))
    times, res = [], None
    t_loop = time.perf_counter()
    for i in range(args.steps):
        ms, threads, res = cpu_msm(MSM_K)
        times.append(ms)
        elapsed = time.perf_counter() - t_loop
        if i + 1 < args.steps and elapsed + elapsed / (i + 1) > CPU_BUDGET_S:
            break
    ms_per_step = float(np.mean(times))
    ntt_ms = cpu_ntt(NTT_K)
    sample = (f"C restatement of halo2curves 0.6 best_multiexp (rayon chunk-per-thread Booth Pippenger), {len(times)} full-size "
              f"2^{MSM_K} MSM(s) on {C.num_threads()} threads, mean; inputs generated in {gen_s:.1f} s outside the timed loop")
    line = {
        "impl": "reference", "metric": f"bn254_msm_2^{MSM_K}_ms", "value": ms_per_step, "unit": "ms", "n_gpus": args.gpus,
        "steps": len(times), "steps_requested": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "measured_at": MSM_K, "runs_ms": times,
        "higher_is_better": False, "scaling": "strong",
        "vs_baseline": None, "dtype": "u32x8 (254-bit Montgomery integers)", "data": "synthetic",
        "config": workload_config(args.gpus), "parity": "unpinned",
        "result_sha256": point_sha256(res),
        "cpu_baseline": {"value": ms_per_step, "unit": "ms", "cores": C.num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": ms_per_step, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "ntt": {"metric": f"fr_ntt_2^{NTT_K}_ms", "value": ntt_ms, "unit": "ms", "measured_at": NTT_K,
                "sample": f"best_fft restatement (bit-reverse + radix-2, recursive split over threads) on 2^{NTT_K}, one run"},
        "gpu_launches": 0,
    }
    emit(line)


# ---------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------
def run_gpu(args, rank: int, world: int, local_rank: int):
    import torch
    import torch.distributed as dist
    import h2b200

    torch.cuda.set_device(local_rank)
    if world > 1:
        # NCCL_DEBUG=VERSION (set by some images) makes NCCL print a banner on stdout; the contract is ONE JSON line.
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = h2b200.Context([local_rank])
    eng = h2b200.B200Engine(ctx)
    # NCCL carries only the set-up and timing plumbing (mailbox handles, barriers, max over ranks): the MSM's own exchange is
    # the library's peer-store kernel
    eng.peer_attach(rank, world, h2b200.dist_all_gather_bytes if world > 1 else None)
    if world > 1:
        dist.barrier()          # h2b_peer_attach clears this rank's mailbox: every rank is attached before any rank publishes

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- inputs: the SAME 2^MSM_K problem at every N, this rank's contiguous point range of it --------------------
    n_total = 1 << MSM_K
    lo, hi = n_total * rank // world, n_total * (rank + 1) // world
    n_loc = hi - lo
    t0 = h2b200.synth.bench_t0()
    d_bases = h2b200.synth.device_bases(ctx, n_loc, t0 + lo)
    torch.cuda.synchronize()
    t_reg = time.perf_counter()
    desc = eng.get_base_descriptor(d_bases)          # registration = copy-free adoption + window-multiple precomputation
    torch.cuda.synchronize()
    registration_ms = (time.perf_counter() - t_reg) * 1e3
    pre_c, registration_bytes = desc.info()
    all_scalars = h2b200.synth.chacha_scalars_device(ctx, n_total)
    d_scalars = all_scalars[lo:hi].clone()
    del all_scalars
    torch.cuda.empty_cache()
    pageable = d_scalars.cpu().numpy().view(np.uint64)             # an ordinary host array: what a Rust Vec<Fr> is
    h_pinned = torch.from_numpy(pageable.view(np.int64)).pin_memory()

    def msm_step(scalars):
        """one whole MSM; every rank returns the normalised result"""
        if world == 1:
            return eng.msm_with_cached_base(scalars, desc)
        return eng.msm_sharded(scalars, desc, n_total)

    sampler = ClockSampler(local_rank)
    sampler.start()                      # nvidia-smi needs ~0.3 s to deliver its first row: start it under the warm-up load
    for _ in range(args.warmup):
        msm_step(d_scalars)
    ctx.profile_enable(True)
    ctx.profile_read()
    launches0 = ctx.launch_count()
    barrier()
    n_rows_before = len(sampler.rows)
    t0w = time.perf_counter()
    dev_ms = []
    for _ in range(args.steps):
        res_dev = msm_step(d_scalars)
        dev_ms.append(ctx.last_call_ms())          # CUDA events on the launching stream, first to last enqueued operation
    barrier()
    wall_ms = (time.perf_counter() - t0w) * 1e3
    n_rows_timed = len(sampler.rows) - n_rows_before
    launches = ctx.launch_count() - launches0
    prof = ctx.profile_read()
    ctx.profile_enable(False)
    ms_per_step = max_over_ranks(wall_ms / args.steps)
    max_dev_ms = max_over_ranks(float(np.mean(dev_ms)))

    # ---- e2e: host scalars in, 96-byte point out, through the same public call.  `e2e` = PAGEABLE buffers (what the reference
    # hands over), `e2e_pinned` = page-locked ones ----------------------------------------------------------------------
    def timed_e2e(src):
        for _ in range(min(args.warmup, 2)):
            msm_step(src)
        barrier()
        t = time.perf_counter()
        for _ in range(args.steps):
            r = msm_step(src)
        barrier()
        return max_over_ranks((time.perf_counter() - t) * 1e3 / args.steps), r

    e2e_ms, res_e2e = timed_e2e(pageable)
    e2e_pinned_ms, res_pin = timed_e2e(h_pinned.numpy().view(np.uint64))
    if not (np.array_equal(res_dev, res_e2e) and np.array_equal(res_dev, res_pin)):
        raise SystemExit("bench: device-resident and host-buffer MSM results differ")

    # ---- NTT 2^24: device-resident, then host buffers in/out ----------------------------------------------
    dom = h2b200.EvaluationDomain(1, NTT_K, ctx)
    d_poly = h2b200.synth.chacha_scalars_device(ctx, 1 << NTT_K)
    hp = d_poly.cpu().numpy().view(np.uint64)                       # pageable
    hp_pinned = torch.from_numpy(hp.view(np.int64)).pin_memory()
    for _ in range(args.warmup):
        dom.coeff_to_lagrange(d_poly)
    ctx.profile_enable(True)
    ctx.profile_read()
    barrier()
    t0w = time.perf_counter()
    for _ in range(args.steps):
        dom.coeff_to_lagrange(d_poly)
    barrier()
    ntt_ms = max_over_ranks((time.perf_counter() - t0w) * 1e3 / args.steps)
    ntt_prof = ctx.profile_read()
    ctx.profile_enable(False)

    def timed_ntt_host(buf):
        dom.coeff_to_lagrange(buf)
        barrier()
        t = time.perf_counter()
        for _ in range(args.steps):
            dom.coeff_to_lagrange(buf)
        barrier()
        return max_over_ranks((time.perf_counter() - t) * 1e3 / args.steps)

    ntt_e2e_ms = timed_ntt_host(hp)
    ntt_e2e_pinned_ms = timed_ntt_host(hp_pinned.numpy().view(np.uint64))
    del d_poly, hp_pinned

    # ---- batched coset NTTs dealt over the ranks (north_star: "the independent per-column NTTs of the extended-domain quotient
    # computation are batched across GPUs"): NTT_BATCH_COLS columns of coeff_to_extended 2^22 -> 2^24 (evaluate_h,
    # plonk/evaluation.rs:344-361); rank r takes the columns c = r (mod N); no collective ---------------------------------
    bdom = h2b200.EvaluationDomain(5, NTT_BATCH_K, ctx)             # quotient degree 4 -> e = 4
    my_cols = list(range(rank, NTT_BATCH_COLS, world))
    nb_in, nb_out = 1 << NTT_BATCH_K, bdom.extended_len()
    col_src = h2b200.synth.chacha_scalars_device(ctx, nb_in, bytes([7]) * 32)
    d_cols = [col_src.roll(c, 0).contiguous() for c in my_cols]     # distinct columns, cheaply
    d_outs = [torch.empty((nb_out, 4), dtype=torch.int64, device="cuda") for _ in my_cols]
    bsteps = max(1, min(args.steps, 5))

    def batch_dev():
        bdom.coeff_to_extended_many(d_cols, d_outs)

    batch_dev()
    barrier()
    t0w = time.perf_counter()
    for _ in range(bsteps):
        batch_dev()
    barrier()
    batch_ms = max_over_ranks((time.perf_counter() - t0w) * 1e3 / bsteps)
    batch_digest = None
    if my_cols:
        import hashlib
        batch_digest = hashlib.sha256(d_outs[0][:4096].cpu().numpy().tobytes()).hexdigest()
    h_cols = [c.cpu().numpy().view(np.uint64) for c in d_cols]      # pageable host columns
    h_outs = [np.empty((nb_out, 4), dtype=np.uint64) for _ in my_cols]
    for o in h_outs:
        o.fill(0)                                                   # touch the pages outside the timed region
    del d_cols, d_outs, col_src
    torch.cuda.empty_cache()

    def batch_host():
        bdom.coeff_to_extended_batch(h_cols, h_outs)

    batch_host()
    barrier()
    t0w = time.perf_counter()
    for _ in range(bsteps):
        batch_host()
    barrier()
    batch_e2e_ms = max_over_ranks((time.perf_counter() - t0w) * 1e3 / bsteps)
    del h_cols, h_outs

    clocks = sampler.stop()             # sampled over every timed region of this run (MSM, e2e MSM, NTT, e2e NTT, batch)
    clocks["samples_in_msm_timed_region"] = n_rows_timed
    wide_peak, fmul_peak = ctx.imad_peak()
    notice = ctx.last_notice()
    if rank != 0:
        ctx.close()
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (bucket accumulation), from CUDA events on the launching stream ---------
    hbm_peak, peak_src = peaks()
    acc = prof.get("msm_accumulate", {"ms": 0.0, "launches": 1})
    # The scalars are processed in up to 8 chunks (one accumulate launch each, binning of the next chunk overlapped on a second
    # stream), so the unit is the STEP: all accumulate launches of one MSM against the whole MSM's algorithmic additions.
    acc_ms = acc["ms"] / max(1, args.steps)
    acc_launches_per_step = acc["launches"] / max(1, args.steps)
    kernel_ms = {k: v["ms"] / args.steps for k, v in prof.items()}
    adds = MSM_ADDS.get(MSM_K, 0.0) / world                 # this rank's share of the SURVEY 8(d) addition count
    alg_bytes = 96.0 * n_loc                                # 32 B scalar + 64 B base per point
    imad_achieved = adds * MULS_PER_MADD * WIDE_PER_FIELD_MUL / (acc_ms * 1e-3) if acc_ms else 0.0
    # The binding roofline of both kernels is the integer-multiply pipe ("roofline.achieved(MSM) = adds(k) x
    # IMAD / t / IMAD_peak"), so that is what `achieved/peak/frac` report; the HBM figures the contract names sit in `hbm`.
    hbm_achieved = alg_bytes / (acc_ms * 1e-3) / 1e9 if acc_ms else 0.0
    acc_traffic, acc_traffic_src = ncu_traffic("msm_accumulate_kernel", MSM_K)
    # `achieved` / `frac` count the IMAD.WIDE instructions the kernel EXECUTES (a fraction of the pipe's time, <= 1 by construction);
    # SURVEY's table assumes 10 full multiplications per mixed addition, more than the kernel spends, so the table-count figure can
    # exceed the peak and is reported beside it as `frac_vs_survey_table`.
    executed_rate = (n_loc * 13 * 1160.0 / (acc_ms * 1e-3)) if (acc_ms and pre_c == 20) else None
    sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
    nominal_peak = sms * 32.0 * clocks["sm_mhz"] * 1e6 if clocks.get("sm_mhz") else None      # 4 SMSPs x 32 lanes / 4 cycles per SM
    roofline = {
        "kernel": "msm_accumulate_kernel", "launches_per_step": acc_launches_per_step,
        "bound": "imad", "bound_note": "integer-multiply pipe: neither hbm nor tensor binds this path (SURVEY.md 8d, north_star); the hbm figures are under `hbm`",
        "achieved": executed_rate if executed_rate else imad_achieved, "peak": wide_peak, "unit": "IMAD.WIDE/s",
        "frac": ((executed_rate if executed_rate else imad_achieved) / wide_peak) if wide_peak else None,
        "frac_counts": "executed instructions" if executed_rate else "SURVEY.md 8d table (no executed count for this window size)",
        "achieved_vs_survey_table": imad_achieved, "frac_vs_survey_table": imad_achieved / wide_peak if wide_peak else None,
        "peak_nominal": nominal_peak, "peak_nominal_source": f"{sms} SMs x 4 sub-partitions x 32 lanes / 4 cycles x the SM clock sampled under load",
        "peak_measured_over_nominal": (wide_peak / nominal_peak) if (wide_peak and nominal_peak) else None,
        "peak_source": "h2b_imad_peak: IMAD.WIDE.U32.X carry chains measured live in this run (half-rate on sm_100a; "
                       "profiles/r1_imad_micro.txt, r1_fp64_micro.txt: FP64 and ALU ops do not co-issue beside it)",
        "field_muls_per_s_peak": fmul_peak,
        "algorithmic": f"{adds:.4g} XYZZ mixed adds x {MULS_PER_MADD} field muls x {WIDE_PER_FIELD_MUL} IMAD.WIDE per step (SURVEY.md 8d table), over the summed duration of the step's accumulate launches",
        "executed": "n x 13 windows (c = 20, precomputed window multiples) mixed adds x 1160 IMAD.WIDE "
                    "(6 mul + 2 dedicated squares + one fused product difference): fewer than the table's count, hence frac < frac_vs_survey_table",
        "traffic": (acc_traffic * acc_launches_per_step) if (acc_traffic and world == 1) else None,
        "traffic_source": (acc_traffic_src if world == 1 else None) or "no ncu capture of this round's code at this size / rank count: null rather than a stale figure",
        "traffic_note": "ncu dram__bytes_read+write per accumulate launch x launches per step; 13 gathered 64-byte multiples per "
                        "point at two DRAM atoms each dominate (algorithmic input = 96 B per point)",
        "hbm": {"bound": "hbm", "achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_achieved / hbm_peak,
                "peak_source": peak_src, "algorithmic": "96 B per point (32 B scalar + 64 B base)"},
        "kernel_ms_per_step": kernel_ms,
    }
    npass = sum(v["launches"] for v in ntt_prof.values()) // max(1, args.steps)
    ntt_kernel_ms = sum(v["ms"] for v in ntt_prof.values()) / args.steps
    ntt_bytes = 128.0 * (1 << NTT_K) if NTT_K > 21 else 64.0 * (1 << NTT_K)
    ntt_muls = (1 << NTT_K) / 2 * NTT_K
    ntt_imad = ntt_muls * WIDE_PER_FIELD_MUL / (ntt_kernel_ms * 1e-3)
    ntt_hbm = ntt_bytes / (ntt_kernel_ms * 1e-3) / 1e9
    ntt_traffic, ntt_traffic_src = ncu_traffic("ntt_pass_kernel", NTT_K)
    ntt = {
        "metric": f"fr_ntt_2^{NTT_K}_ms", "value": ntt_ms, "unit": "ms", "passes": npass,
        "replicas_per_s": world * 1e3 / ntt_ms,
        "replicas_note": "a single transform stays on one GPU; with N ranks these are N independent replicas (see ntt_batch for the sharded workload)",
        "e2e": {"value": ntt_e2e_ms, "unit": "ms", "h2d_bytes_per_step": 32 << NTT_K, "d2h_bytes_per_step": 32 << NTT_K, "host_buffers": "pageable"},
        "e2e_pinned": {"value": ntt_e2e_pinned_ms, "unit": "ms"},
        "roofline": {"kernel": "ntt_pass_kernel (all passes of one transform)",
                     "bound": "imad", "bound_note": "SURVEY.md 8d: max(hbm_fraction, imad_fraction); 254-bit butterflies are multiplier bound",
                     "achieved": ntt_imad, "peak": wide_peak, "unit": "IMAD.WIDE/s", "frac": ntt_imad / wide_peak if wide_peak else None,
                     "algorithmic": f"(n/2) log2 n = {ntt_muls:.4g} field muls x {WIDE_PER_FIELD_MUL} IMAD.WIDE",
                     "traffic": (ntt_traffic * npass) if ntt_traffic else None, "traffic_source": ntt_traffic_src,
                     "traffic_note": f"sum over the {npass} passes of one transform (ncu dram__bytes_read+write); {npass} passes move "
                                     f"{npass * 64} B per element + the boundary-twiddle table, against SURVEY's 2-pass figure of 128 B: a deliberate trade of "
                                     "DRAM traffic (8-9 % of HBM peak) for multiplications on a multiplier-bound kernel",
                     "hbm": {"bound": "hbm", "achieved": ntt_hbm, "peak": hbm_peak, "unit": "GB/s", "frac": ntt_hbm / hbm_peak,
                             "algorithmic": "128 B per element (two HBM round trips, SURVEY.md 8d)"}},
    }
    col_in_b, col_out_b = 32 * nb_in, 32 * nb_out
    ntt_batch = {
        "metric": f"coeff_to_extended 2^{NTT_BATCH_K} -> 2^{bdom.extended_k}, {NTT_BATCH_COLS} independent columns dealt over the ranks (evaluate_h's coset NTTs)",
        "value": NTT_BATCH_COLS * 1e3 / batch_ms, "unit": "columns/s", "ms_per_batch": batch_ms, "steps": bsteps, "scaling": "strong",
        "columns_per_rank": len(my_cols), "first_column_sha256_4096_rows": batch_digest,
        "e2e": {"value": NTT_BATCH_COLS * 1e3 / batch_e2e_ms, "unit": "columns/s", "ms_per_batch": batch_e2e_ms,
                "h2d_bytes_per_step": col_in_b * NTT_BATCH_COLS, "d2h_bytes_per_step": col_out_b * NTT_BATCH_COLS, "host_buffers": "pageable",
                "host_gb_per_s": (col_in_b + col_out_b) * NTT_BATCH_COLS / (batch_e2e_ms * 1e-3) / 1e9,
                "note": "bounded by the host side: every rank's columns cross the same host memory / PCIe root complexes; see DESIGN.md 5"},
    }
    # the result of the seeded 2^24 problem is a committed constant (tests/golden/msm_results.json, from the CPU port): checked at
    # EVERY rank count, not only where the CPU port runs beside the GPU (N = 1)
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "msm_results.json"))).get(str(MSM_K))
    golden_ok = (point_sha256(res_dev) == golden["affine_sha256"]) if golden else None
    if golden_ok is False:
        raise SystemExit(f"bench: the 2^{MSM_K} MSM result {point_sha256(res_dev)} differs from the committed oracle result {golden['affine_sha256']}")
    line = {
        "metric": f"bn254_msm_2^{MSM_K}_ms", "value": ms_per_step, "unit": "ms", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "u32x8 (254-bit Montgomery integers)", "data": "synthetic",
        "config": workload_config(world), "parity": "unpinned",
        "result_sha256": point_sha256(res_dev), "result_matches_golden_sha": golden_ok,
        "e2e": {"value": e2e_ms, "unit": "ms", "h2d_bytes_per_step": 32 * n_total, "d2h_bytes_per_step": 96, "host_buffers": "pageable"},
        "e2e_pinned": {"value": e2e_pinned_ms, "unit": "ms", "host_buffers": "page-locked"},
        "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "ntt": ntt, "ntt_batch": ntt_batch,
        "msm_points_per_s": n_total / (ms_per_step * 1e-3),
        "registration": {"ms": registration_ms, "device_bytes": registration_bytes, "precompute": pre_c != 0, "window_bits": pre_c,
                         "points": n_loc, "notice": notice,
                         "what": "one-off cost of get_base_descriptor on this rank's bases (device-resident): the multiples 2^(c*j) * P of every window"},
        "exchange": ("none (one GPU)" if world == 1 else
                     "h2b_msm_sharded: reduced terms stored into every peer's mailbox over NVLink by the last kernel of the reduction, "
                     "folded on the device; every rank returns the whole result (NCCL only for set-up, barriers and max-over-ranks timing)"),
        "timing": {"value": "host clock around the synchronous calls, bracketed by barrier + cuda synchronize, max over ranks "
                            "(includes the host-side Horner over the reduced terms and the normalisation)",
                   "device_events_ms_per_step": max_dev_ms},
    }
    if world == 1 and not args.no_seq:
        seq_ms, seq_launches = proof_sequence_gpu(ctx, eng)
        line["k20_kernel_sequence"] = {
            "what": "MSM + NTT calls of one k=20 create_proof (SURVEY 8d counts), host buffers in/out, through MsmAccel / EvaluationDomain mirrors; not a proof",
            "counts": SEQ_COUNTS, "value": seq_ms, "unit": "ms", "gpu_launches": seq_launches}
        res_ms, res_launches = proof_sequence_resident(ctx, eng)
        line["k20_kernel_sequence"]["device_resident"] = {
            "what": "same MSM / NTT counts with device-resident columns, plus the Z-polynomial inversions / running products, the "
                    "quotient folds over 2^23 rows and the opening helpers (SURVEY 8f N1-N3) on the device; witness upload included",
            "extra_counts": RES_COUNTS, "value": res_ms, "unit": "ms", "gpu_launches": res_launches}
        line["k20_proof"] = k20_proof(ctx, eng)
        line["k12_proof"] = k20_proof(ctx, eng, k=12, reps=5, circuit="StandardPlonk", threads=2)
        line["msm_raw"] = msm_raw_bench(ctx, eng)
        if not args.no_cpu:
            cpu_total, cpu_parts, thr = proof_sequence_cpu()
            line["k20_kernel_sequence"]["cpu_baseline"] = {"value": cpu_total, "unit": "ms", "cores": thr, "kind": "port",
                                                           "sample": "one call of each kind timed, scaled by the counts", "per_call_ms": cpu_parts}
            line["k20_proof"]["cpu_baseline"] = {"value": cpu_total, "unit": "ms", "cores": thr, "kind": "port",
                                                 "sample": "the MSM + NTT calls of the same proof only (a lower bound on the reference prover): "
                                                           "one call of each kind timed, scaled by the counts"}
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
        if not args.no_seq:
            torch.cuda.empty_cache()
            line["proof_multi_gpu"] = multi_gpu_proof(world)
    if world == 1 and not args.no_cpu:
        # the CPU port beside it, on the SAME inputs at the SAME size (one full 2^24 MSM, ~10 s; one full 2^24 NTT, ~3 s), and the
        # one check of the GPU result against the oracle at the headline size
        ms, threads, cpu_res = cpu_msm(MSM_K)
        line["cpu_baseline"] = {"value": ms, "unit": "ms", "cores": threads, "kind": "port", "measured_at": MSM_K,
                                "sample": f"C restatement of halo2curves 0.6 best_multiexp (rayon chunk-per-thread Booth Pippenger): one full 2^{MSM_K} MSM on the same inputs, no extrapolation"}
        line["result_matches_cpu_oracle"] = bool(np.array_equal(cpu_res, res_dev[:8]))
        nms = cpu_ntt(NTT_K)
        ntt["cpu_baseline"] = {"value": nms, "unit": "ms", "cores": threads, "kind": "port", "measured_at": NTT_K,
                               "sample": f"best_fft restatement on 2^{NTT_K}: one full-size run"}
        if not line["result_matches_cpu_oracle"]:
            raise SystemExit("bench: the 2^%d MSM result differs from the CPU oracle" % MSM_K)
    emit(line)


_JSON_FD = None


def capture_stdout():
    """The contract is ONE JSON line on stdout.  Native libraries (NCCL's version banner, CUDA warnings) write to fd 1
    behind Python's back, so fd 1 is pointed at stderr for the whole run and the JSON line goes to the saved descriptor."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    capture_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-seq", action="store_true", help="skip the k=20 proof-shaped kernel sequence")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.warmup < 3:
        args.warmup = 3
    run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
