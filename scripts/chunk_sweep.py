"""Device-resident MSM of 2^k points (default 21 = one rank's share of the 2^24 MSM on 8 GPUs) for several chunkings of the scalars:
how much of the first chunk's binning stays exposed vs what shorter chunks cost the accumulation.  usage: chunk_sweep.py [k]"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import h2b200
from h2b200 import synth

k = int(sys.argv[1]) if len(sys.argv) > 1 else 21
n = 1 << k
ctx = h2b200.Context([0])
eng = h2b200.B200Engine(ctx)
bases = synth.device_bases(ctx, n, 1)
scal = synth.chacha_scalars_device(ctx, n, synth.SEED_SCALARS)
desc = eng.get_base_descriptor(bases)
out = {}
for setting in ("default", 1, 2, 3, 4, 6, 8):
    ctx.set_option("msm_chunks", None if setting == "default" else setting)
    for _ in range(3):
        eng.msm_with_cached_base(scal, desc)
    ts = []
    for _ in range(20):
        eng.msm_with_cached_base(scal, desc)
        ts.append(ctx.last_call_ms())
    out[str(setting)] = round(float(np.median(ts)), 4)
print(json.dumps({"k": k, "median_device_ms_by_msm_chunks": out}))
