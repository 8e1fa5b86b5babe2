"""First-contact GPU probe: integer-pipe calibration, per-kernel timing of NTT and MSM at a few sizes.
Writes gpurun_out/probe.json.  Usage: python scripts/gpu_probe.py [ks...]"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import h2b200
from oracle import cref as C
from oracle import pyref as R

ks = [int(x) for x in sys.argv[1:]] or [16, 20, 24]
ctx = h2b200.Context()
eng = h2b200.B200Engine(ctx)
res = {"imad": {}, "ntt": {}, "msm": {}}
w, f = ctx.imad_peak()
res["imad"] = {"wide_mads_per_s": w, "field_muls_per_s": f}
print("imad.wide/s %.3e  field muls/s %.3e" % (w, f), flush=True)
ctx.profile_enable(True)
for k in ks:
    n = 1 << k
    a = C.random_fr(n, k)
    d = torch.from_numpy(a.view(np.int64)).cuda()
    dom = h2b200.EvaluationDomain(1, k, ctx)
    dom.coeff_to_lagrange(d)          # builds tables
    ctx.profile_read()
    ts = []
    for _ in range(5):
        dom.coeff_to_lagrange(d)
        ts.append(ctx.last_call_ms())
    prof = ctx.profile_read()
    res["ntt"][k] = {"ms": ts, "prof": prof}
    print("ntt k=%d ms=%s" % (k, ["%.3f" % t for t in ts]), prof, flush=True)
    del d
for k in ks:
    n = 1 << k
    s = C.random_fr(n, k)
    t0 = time.time()
    b = C.gen_bases(n, 1)
    desc = eng.get_base_descriptor(b)
    ds = torch.from_numpy(s.view(np.int64)).cuda()
    for c in ([None] if k < 20 else [None, 14, 18, 20]):
        if c is None:
            os.environ.pop("H2B_MSM_C", None)
        else:
            os.environ["H2B_MSM_C"] = str(c)
        out = eng.msm_with_cached_base(ds, desc)
        ctx.profile_read()
        ts = []
        for _ in range(3):
            t1 = time.time()
            out = eng.msm_with_cached_base(ds, desc)
            ts.append((ctx.last_call_ms(), (time.time() - t1) * 1e3))
        prof = ctx.profile_read()
        res["msm"]["%d_c%s" % (k, c)] = {"ms": ts, "prof": prof}
        print("msm k=%d c=%s (dev ms, wall ms)=%s" % (k, c, ts), flush=True)
        print("   ", {kk: round(v["ms"] / 3, 3) for kk, v in prof.items()}, flush=True)
    os.environ.pop("H2B_MSM_C", None)
    desc.release()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/probe.json", "w"), indent=1)
