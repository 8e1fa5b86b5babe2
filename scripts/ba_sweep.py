"""Device-resident MSM under the batched-affine accumulation knobs (msm_ba = 0 / 1 / 2 rounds, msm_ba_regs): total time and the
per-kernel CUDA-event breakdown at each size.   python scripts/ba_sweep.py [sizes=20,21,24] [reps=5]"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import h2b200

sizes = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "20,21,24").split(",")]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
configs = [tuple(int(v) for v in c.split(":")) for c in os.environ.get("BA_CONFIGS", "0:0,1:0,1:1,2:0,2:1").split(",")]
ctx = h2b200.Context([0])
eng = h2b200.B200Engine(ctx)
out = {}
for k in sizes:
    n = 1 << k
    desc = eng.get_base_descriptor(h2b200.synth.device_bases(ctx, n, h2b200.synth.bench_t0()))
    d = h2b200.synth.chacha_scalars_device(ctx, n, h2b200.synth.SEED_SCALARS)
    ref = None
    for ba, regs in configs:
        ctx.set_option("msm_ba", ba)
        ctx.set_option("msm_ba_regs", regs)
        r = eng.msm_with_cached_base(d, desc)
        if ref is None:
            ref = np.array(r)
        same = bool(np.array_equal(ref, np.array(r)))
        ctx.profile_enable(True); ctx.profile_read()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(reps):
            eng.msm_with_cached_base(d, desc)
        torch.cuda.synchronize(); ms = (time.perf_counter() - t0) * 1e3 / reps
        prof = ctx.profile_read(); ctx.profile_enable(False)
        rec = {"ms": round(ms, 3), "same_point": same, "kernels_ms": {nm: round(v["ms"] / reps, 3) for nm, v in prof.items()}}
        out[f"2^{k} ba={ba} regs={regs}"] = rec
        print(f"2^{k} ba={ba} regs={regs}", json.dumps(rec), flush=True)
    desc.release()
print(json.dumps(out))
