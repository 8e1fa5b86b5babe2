"""Size sweep of the two hot kernels (BASELINE.json configs[1] and configs[2]): MSM and NTT 2^16..2^26, device resident and through
host buffers, with the per-kernel CUDA-event breakdown; MSM inputs are the ChaCha20-seeded sets of SURVEY.md 8(d), uniform and the
two non-uniform ones.  Prints one JSON object.

    python scripts/sweep.py [--msm 16,18,20,22,24] [--ntt 16,18,20,22,24,26] [--reps 5]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import h2b200


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3 / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--msm", default="16,18,20,22,24")
    ap.add_argument("--ntt", default="16,18,20,22,24,26")
    ap.add_argument("--poly", default="20,23")
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    ctx = h2b200.Context([0])
    eng = h2b200.B200Engine(ctx)
    out = {"msm": {}, "ntt": {}, "poly": {}}
    # point additions of the best classic Pippenger at each size (x 2720 IMAD issue slots = 1280 IMAD.WIDE)
    table_adds = {16: 1.475e6, 18: 5.014e6, 20: 1.769e7, 22: 6.488e7, 24: 2.317e8, 26: 8.556e8}
    import ctypes
    wide = ctypes.c_double()
    muls = ctypes.c_double()
    ctx.check(ctx.lib.h2b_imad_peak(ctx.handle, ctypes.byref(wide), ctypes.byref(muls)))
    out["imad_wide_peak_per_s"] = wide.value
    poly = h2b200.PolyOps(ctx)
    for k in [int(x) for x in args.msm.split(",") if x]:
        n = 1 << k
        d_bases = h2b200.synth.device_bases(ctx, n, h2b200.synth.bench_t0())
        t0 = time.perf_counter()
        desc = eng.get_base_descriptor(d_bases)
        torch.cuda.synchronize()
        reg_ms = (time.perf_counter() - t0) * 1e3
        d = h2b200.synth.chacha_scalars_device(ctx, n, h2b200.synth.SEED_SCALARS)
        h = d.cpu().pin_memory()
        eng.msm_with_cached_base(d, desc)
        ctx.profile_enable(True)
        ctx.profile_read()
        dev_ms = timed(lambda: eng.msm_with_cached_base(d, desc), args.reps)
        prof = ctx.profile_read()
        ctx.profile_enable(False)
        hn = h.numpy().view(np.uint64)
        host_ms = timed(lambda: eng.msm_with_cached_base(hn, desc), args.reps)
        pg = np.array(hn)                 # a pageable copy: what a Rust Vec<Fr> is
        pageable_ms = timed(lambda: eng.msm_with_cached_base(pg, desc), args.reps)
        rec = {"device_ms": dev_ms, "host_ms": host_ms, "host_pageable_ms": pageable_ms, "points_per_s": n / (dev_ms * 1e-3),
               "registration_ms": reg_ms, "kernels_ms": {name: v["ms"] / (args.reps + 1) for name, v in prof.items()}}
        if k in table_adds:
            rec["imad_fraction_of_measured_peak"] = table_adds[k] * 1280 / (dev_ms * 1e-3) / wide.value
        # the two non-uniform scalar sets of SURVEY.md 8(d): (i) 90 % zeros / ones + 10 % uniform, (ii) every scalar < 2^16
        gen = torch.Generator(device="cuda")
        gen.manual_seed(5)
        u = torch.rand(n, device="cuda", generator=gen)
        one = torch.from_numpy(h2b200.fields.fr_mont(1).view(np.int64).reshape(1, 4)).cuda()
        sk = d.clone()
        sk[u < 0.45] = 0
        sk[(u >= 0.45) & (u < 0.9)] = one
        rec["device_ms_90pct_zeros_ones"] = timed(lambda: eng.msm_with_cached_base(sk, desc), args.reps)
        raw = torch.zeros((n, 4), dtype=torch.int64, device="cuda")
        raw[:, 0] = torch.randint(0, 1 << 16, (n,), device="cuda", generator=gen)
        r_mont = torch.from_numpy(h2b200.fields.fr_mont(pow(2, 256, h2b200.fields.R_MOD)).view(np.int64).reshape(1, 4)).cuda()
        small = poly.mul(raw, r_mont.expand(n, 4).contiguous())       # (v, 0, 0, 0) * mont(R) = v * R = Montgomery form of v
        rec["device_ms_scalars_below_2p16"] = timed(lambda: eng.msm_with_cached_base(small, desc), args.reps)
        out["msm"][k] = rec
        desc.release()
        del d_bases, d, h, sk, raw, small, u
        torch.cuda.empty_cache()
    for k in [int(x) for x in args.ntt.split(",") if x]:
        n = 1 << k
        dom = h2b200.EvaluationDomain(1, k, ctx)
        h = torch.from_numpy(h2b200.synth.random_scalars(n, 99).view(np.int64)).pin_memory()
        d = h.cuda()
        dev_ms = timed(lambda: dom.coeff_to_lagrange(d), args.reps)
        l2c_ms = timed(lambda: dom.lagrange_to_coeff(d), args.reps)
        hn = h.numpy().view(np.uint64)
        host_ms = timed(lambda: dom.coeff_to_lagrange(hn), max(1, args.reps // 2))
        pg = np.array(hn)
        pageable_ms = timed(lambda: dom.coeff_to_lagrange(pg), max(1, args.reps // 2))
        rec = {"forward_device_ms": dev_ms, "lagrange_to_coeff_device_ms": l2c_ms, "forward_host_ms": host_ms,
               "forward_host_pageable_ms": pageable_ms,
               "field_muls_per_s": n / 2 * k / (dev_ms * 1e-3), "alg_GBps": (128 if k > 21 else 64) * n / (dev_ms * 1e-3) / 1e9}
        if k >= 18 and k <= 26:
            # coset extension n/4 -> n (quotient degree 4) and back, device resident
            edom = h2b200.EvaluationDomain(5, k - 2, ctx)
            small = d[: n // 4].contiguous()
            ext = edom.coeff_to_extended(small)
            rec["coeff_to_extended_device_ms"] = timed(lambda: edom.coeff_to_extended(small), args.reps)
            rec["extended_to_coeff_device_ms"] = timed(lambda: edom.extended_to_coeff(ext), args.reps)
            del ext, small
        out["ntt"][k] = rec
        del d, h
        torch.cuda.empty_cache()
    ops = h2b200.PolyOps(ctx)
    for k in [int(x) for x in args.poly.split(",") if x]:
        n = 1 << k
        d = torch.from_numpy(h2b200.synth.random_scalars(n, 7).view(np.int64)).cuda()
        rec = {}
        rec["batch_invert_ms"] = timed(lambda: ops.batch_invert(d), args.reps)
        rec["grand_product_ms"] = timed(lambda: ops.grand_product(d, 1, n), args.reps)
        rec["eval_polynomial_ms"] = timed(lambda: ops.eval_polynomial(d, 12345), args.reps)
        rec["kate_division_ms"] = timed(lambda: ops.kate_division(d, 12345), args.reps)
        # one degree-3 gate over the rows of this domain (4 loads, 3 multiplications + the y fold per row)
        from h2b200 import evaluation as E
        g = E.GraphEvaluator()
        gate = ("prod", ("query", "fixed", 0, 0), ("sum", ("prod", ("query", "advice", 0, 0), ("query", "advice", 1, 1)),
                                                   ("neg", ("query", "advice", 0, -1))))
        g.add_calculation(("horner", (10,), (g.add_expression(gate),), (9,)))
        vals = torch.zeros_like(d)
        kk = max(k - 3, 1)
        rec["graph_gate_ms"] = timed(lambda: g.evaluate(ctx, [d], [d, d], [], [], 3, 5, 7, 11, kk, k, previous=vals, out=vals), args.reps)
        rec["lookup_fold_ms"] = timed(lambda: E.lookup_fold(ctx, vals, d, d, d, d, d, d, d, 3, 5, 7, kk, k), args.reps)
        rec["permutation_fold_4cols_ms"] = timed(lambda: E.permutation_fold(ctx, vals, [d], [d, d, d, d], [d, d, d, d], 4, d, d, d, 3, 5, 7,
                                                                            h2b200.fields.FR_ZETA, 9, 11, 5, kk, k), args.reps)
        out["poly"][k] = rec
        del d, vals
        torch.cuda.empty_cache()
    print(json.dumps(out))
    ctx.close()


if __name__ == "__main__":
    main()
