"""2^k-point device-resident MSM under the stream-priority / accumulate-variant / binning block-size knobs (each setting in a fresh
context, since the streams are created with the context).  usage: prio_sweep.py [k]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
k = int(sys.argv[1]) if len(sys.argv) > 1 else 24
if len(sys.argv) > 2 and sys.argv[2] == "child":
    import numpy as np
    sys.path.insert(0, ROOT)
    import h2b200
    from h2b200 import synth
    n = 1 << k
    ctx = h2b200.Context([0])
    eng = h2b200.B200Engine(ctx)
    desc = eng.get_base_descriptor(synth.device_bases(ctx, n, 1))
    scal = synth.chacha_scalars_device(ctx, n, synth.SEED_SCALARS)
    for _ in range(3):
        eng.msm_with_cached_base(scal, desc)
    ts = []
    for _ in range(10):
        eng.msm_with_cached_base(scal, desc)
        ts.append(ctx.last_call_ms())
    print("RESULT", float(np.median(ts)))
    sys.exit(0)
out = {}
for name, env in [("default", {}), ("prio_equal", {"H2B_SORT_PRIO": "0"}), ("prio_low", {"H2B_SORT_PRIO": "-1"}),
                  ("v2_96regs", {"H2B_ACC_VARIANT": "2"}), ("v2_prio_equal_tb128", {"H2B_ACC_VARIANT": "2", "H2B_SORT_PRIO": "0", "H2B_MSM_SORT_TB": "128"}),
                  ("v2_prio_low_tb128", {"H2B_ACC_VARIANT": "2", "H2B_SORT_PRIO": "-1", "H2B_MSM_SORT_TB": "128"}),
                  ("v2_prio_high_tb128", {"H2B_ACC_VARIANT": "2", "H2B_MSM_SORT_TB": "128"}),
                  ("tb128", {"H2B_MSM_SORT_TB": "128"}), ("reduce_seg64", {"H2B_MSM_REDUCE_SEG": "64"}),
                  ("reduce_seg128", {"H2B_MSM_REDUCE_SEG": "128"}), ("reduce_seg512", {"H2B_MSM_REDUCE_SEG": "512"}),
                  ("reduce_seg1024", {"H2B_MSM_REDUCE_SEG": "1024"}), ("prio_low_tb128", {"H2B_SORT_PRIO": "-1", "H2B_MSM_SORT_TB": "128"})]:
    r = subprocess.run([sys.executable, __file__, str(k), "child"], env=dict(os.environ, **env), capture_output=True, text=True)
    line = [l for l in r.stdout.split("\n") if l.startswith("RESULT")]
    out[name] = round(float(line[0].split()[1]), 3) if line else r.stderr[-300:]
print(json.dumps({"k": k, "median_device_ms": out}))
