"""A/B of one context knob on the device-resident and pageable-host MSM: python scripts/knob_ab.py <knob> <a> <b> [sizes=20,21,24] [reps=10]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import h2b200

knob, va, vb = sys.argv[1], sys.argv[2], sys.argv[3]
sizes = [int(x) for x in (sys.argv[4] if len(sys.argv) > 4 else "20,21,24").split(",")]
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 10
ctx = h2b200.Context([0])
eng = h2b200.B200Engine(ctx)
for k in sizes:
    n = 1 << k
    desc = eng.get_base_descriptor(h2b200.synth.device_bases(ctx, n, h2b200.synth.bench_t0()))
    d = h2b200.synth.chacha_scalars_device(ctx, n, h2b200.synth.SEED_SCALARS)
    h = d.cpu().numpy().view(np.uint64)
    ref = None
    for rnd in range(3):
        for v in (va, vb):
            ctx.set_option(knob, v)
            r = eng.msm_with_cached_base(d, desc)
            ref = np.array(r) if ref is None else ref
            same = bool(np.array_equal(ref, np.array(r)))
            ctx.profile_enable(True); ctx.profile_read()
            times = []
            for _ in range(reps):
                eng.msm_with_cached_base(d, desc)
                times.append(ctx.last_call_ms())
            prof = ctx.profile_read(); ctx.profile_enable(False)
            ht = []
            for _ in range(4):
                rh = eng.msm_with_cached_base(h, desc)
                ht.append(ctx.last_call_ms())
            same = same and bool(np.array_equal(ref, np.array(rh)))
            print(f"2^{k} {knob}={v}: device median {np.median(times):.3f} min {min(times):.3f} ms; pageable host median {np.median(ht):.3f} ms; same_point={same}; "
                  + " ".join(f"{nm.replace('msm_', '')} {x['ms'] / reps:.3f}" for nm, x in sorted(prof.items())), flush=True)
    desc.release()
