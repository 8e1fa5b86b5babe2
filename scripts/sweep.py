"""Size sweep of the two hot kernels (BASELINE.json configs[1] and configs[2]): MSM 2^16..2^24 and NTT 2^16..2^26, device
resident and through host buffers, with the per-kernel CUDA-event breakdown.  Prints one JSON object.

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
    for k in [int(x) for x in args.msm.split(",") if x]:
        n = 1 << k
        d_bases = h2b200.synth.device_bases(ctx, n, 1)
        desc = eng.get_base_descriptor(d_bases)
        h = torch.from_numpy(h2b200.synth.random_scalars(n, 1234).view(np.int64)).pin_memory()
        d = h.cuda()
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
        out["msm"][k] = {"device_ms": dev_ms, "host_ms": host_ms, "host_pageable_ms": pageable_ms, "points_per_s": n / (dev_ms * 1e-3),
                         "kernels_ms": {name: v["ms"] / (args.reps + 1) for name, v in prof.items()}}
        desc.release()
        del d_bases, d, h
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
