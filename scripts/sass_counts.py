#!/usr/bin/env python
"""Per-kernel SASS mnemonic counts of libwgs_rasterizer.so (cuobjdump -sass): the instructions that show which hardware
path a kernel uses -- HMMA (tensor pipe), LDSM (ldmatrix), LDGSTS (cp.async), RED/REDG (fire-and-forget atomics, .F32x2
vector form), multimem (NVSwitch reduce / broadcast), UCGABAR / cluster barriers and DSMEM (ld/st.shared::cluster map to
LDS/STS with a cluster address + MAPA), MUFU.

    python scripts/sass_counts.py [lib.so] > profiles/rNN_sass_counts.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "wildgs_slam_blackwell_b200", "libwgs_rasterizer.so")
PATTERNS = [("HMMA", r"^HMMA"), ("LDSM", r"^LDSM"), ("LDGSTS", r"^LDGSTS"), ("REDG/RED", r"^RED"), ("RED .F32x2", r"^RED.*F32x2"),
            ("ATOMG/ATOM", r"^ATOM"), ("ATOMS", r"^ATOMS"), ("multimem (LDGMC / ST..MMEM)", r"^(LDGMC|STGMC|REDGMC|.*MMEM|.*\.MC\b)"),
            ("MAPA", r"^MAPA"), ("UCGABAR (cluster barrier)", r"^UCGABAR"), ("MUFU", r"^MUFU"), ("SHFL", r"^SHFL"), ("VOTE", r"^VOTE"),
            ("REDUX", r"^REDUX"), ("BAR", r"^BAR"), ("FFMA", r"^FFMA"), ("LDG", r"^LDG"), ("STG", r"^STG"),
            ("LDS", r"^LDS(?!M)"), ("STS", r"^STS"), ("ACQBULK / PREEXIT (griddepcontrol)", r"^(ACQBULK|PREEXIT)"), ("UBLKCP/UTMA (TMA)", r"^(UBLKCP|UTMA)")]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, name = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name).replace("wgs::", "").replace("void ", "")
            kernels[name] = []
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Za-z0-9_.]+)", line)
        if m and name:
            kernels[name].append(m.group(1))
    cols = [p[0] for p in PATTERNS]
    print(f"# SASS mnemonic counts per kernel ({os.path.basename(LIB)}, cuobjdump -sass, sm_100a; static counts)\n")
    print("| kernel | instr | " + " | ".join(cols) + " |")
    print("|---|---|" + "---|" * len(cols))
    for k, ins in kernels.items():
        row = [sum(1 for i in ins if re.match(rx, i)) for _, rx in PATTERNS]
        print(f"| {k} | {len(ins)} | " + " | ".join(str(r) if r else "" for r in row) + " |")
    allm = collections.Counter(i.split(".")[0] for ins in kernels.values() for i in ins)
    print("\nAll mnemonic stems in the library: " + ", ".join(f"{k} {v}" for k, v in allm.most_common()))


if __name__ == "__main__":
    main()
