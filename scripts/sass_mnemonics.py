"""Per-kernel SASS mnemonic counts of the built library (cuobjdump -sass): the static evidence for the IMAD.WIDE carry chains of the
field multiplier, the TMA bulk copies (UBLKCP) + mbarrier waits (SYNCS) of the NTT passes, and the atomics of the binning kernels.
usage: python scripts/sass_mnemonics.py > profiles/rN_sass_mnemonics.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "yet-another-halo2-fork_b200", "libh2b200.so")


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    parts = re.split(r"\n\s*Function : ", txt)
    print("SASS mnemonic counts per kernel: `cuobjdump -sass yet-another-halo2-fork_b200/libh2b200.so` (sm_100a), scripts/sass_mnemonics.py")
    print(f"{'kernel':58s} {'instrs':>7s} {'IMAD.WIDE':>9s} {'.U32.X':>7s} {'IADD3':>6s} {'LDG':>5s} {'STG':>5s} {'UBLKCP':>6s} {'SYNCS':>6s} "
          f"{'LDS':>5s} {'STS':>5s} {'ATOM+RED':>8s} {'SHFL':>5s}")
    for p in parts[1:]:
        name = p.split("\n", 1)[0].strip()
        dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
        dem = dem.replace("(anonymous namespace)::", "").replace("h2b::", "").replace("void ", "")
        dem = re.sub(r"\(.*", "", dem)
        if dem.startswith("cub::"):
            continue
        ins = re.findall(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", p)
        c = collections.Counter(ins)
        f = lambda pre: sum(v for k, v in c.items() if k.startswith(pre))
        print(f"{dem[:58]:58s} {len(ins):7d} {f('IMAD.WIDE'):9d} {f('IMAD.WIDE.U32.X'):7d} {f('IADD3'):6d} {f('LDG'):5d} {f('STG'):5d} "
              f"{f('UBLKCP'):6d} {f('SYNCS'):6d} {f('LDS'):5d} {f('STS'):5d} {f('ATOM') + f('RED'):8d} {f('SHFL'):5d}")


if __name__ == "__main__":
    main()
