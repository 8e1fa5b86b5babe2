"""BASELINE.json configs[4] shape on one box: ONE process, one h2b context over all visible GPUs.  The 2^24-point MSM is sharded
by point range inside the library (bases registered once, shards run concurrently, host folds the partial sums) and the
independent per-column coset NTTs (k = 24 -> extended 2^25) are dealt round-robin over the devices (h2b_coeff_to_extended_batch).
Checks each result against the single-device context, prints one JSON object.

    python scripts/k24_multi.py [--k 24] [--cols 8]
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=24)
    ap.add_argument("--cols", type=int, default=8)
    args = ap.parse_args()
    ndev = h2b200.load().h2b_device_count()
    n = 1 << args.k
    out = {"devices": ndev, "k": args.k}
    one = h2b200.Context([0])
    eng1 = h2b200.B200Engine(one)
    bases = h2b200.synth.device_bases(one, n, 1).cpu().numpy().view(np.uint64)      # host copy: registration shards it
    scal = h2b200.synth.random_scalars(n, 42)
    d1 = eng1.get_base_descriptor(bases)
    ref = eng1.msm_with_cached_base(scal, d1)
    t = time.perf_counter(); eng1.msm_with_cached_base(scal, d1); out["msm_1gpu_host_ms"] = (time.perf_counter() - t) * 1e3
    dom1 = h2b200.EvaluationDomain(3, args.k, one)
    cols = [h2b200.synth.random_scalars(n, 100 + i) for i in range(args.cols)]
    ref_ext0 = dom1.coeff_to_extended(cols[0])
    t = time.perf_counter()
    for c in cols:
        dom1.coeff_to_extended(c)
    out["c2e_1gpu_host_ms"] = (time.perf_counter() - t) * 1e3
    d1.release(); one.close()
    if ndev > 1:
        allc = h2b200.Context(list(range(ndev)))
        eng = h2b200.B200Engine(allc)
        t = time.perf_counter(); d = eng.get_base_descriptor(bases); out["register_ms"] = (time.perf_counter() - t) * 1e3
        got = eng.msm_with_cached_base(scal, d)
        assert np.array_equal(got, ref), "sharded MSM differs from the single-device result"
        t = time.perf_counter(); eng.msm_with_cached_base(scal, d); out["msm_sharded_host_ms"] = (time.perf_counter() - t) * 1e3
        dom = h2b200.EvaluationDomain(3, args.k, allc)
        ext = dom.coeff_to_extended_batch(cols)
        assert np.array_equal(ext[0], ref_ext0), "batched coset NTT differs from the single-device result"
        t = time.perf_counter(); dom.coeff_to_extended_batch(cols); out["c2e_batch_host_ms"] = (time.perf_counter() - t) * 1e3
        out["extended_k"] = dom.extended_k
        allc.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
