#!/usr/bin/env python
"""sass_cost.py -- static fmaheavy / ALU pipe cost of a kernel from its SASS (cuobjdump -sass), using the per
instruction issue intervals measured on B200 by tools/ubench/int_pipes2.cu (clk per warp-instruction per SMSP):
    IMAD.WIDE  3.1 (addend RZ) / 5.8 (64-bit addend)     IMAD.HI 4.1 (RZ) / 6.1      other IMAD.* 2.05
    ALU class (IADD3, LOP3, SEL, ISETP, SHF, LEA, PRMT, VIADD, MOV ...) 2.0
Usage: python tools/sass_cost.py <lib.so> <kernel-substring> [butterflies-per-thread]
Straight-line kernels only (the NTT passes are fully unrolled); loops are counted once.
"""
import collections
import re
import subprocess
import sys

HEAVY = {"IMAD.WIDE": (3.1, 5.8), "IMAD.HI": (4.1, 6.1)}
ALU = ("IADD3", "LOP3", "SEL", "ISETP", "SHF", "LEA", "PRMT", "VIADD", "MOV", "IABS", "IMNMX", "VIMNMX", "FSEL",
       "PLOP3", "POPC", "FLO", "BREV", "P2R", "R2P", "CS2R", "SGXT", "BMSK")


def main():
    lib, pat = sys.argv[1], sys.argv[2]
    per = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    blocks = re.split(r"\n\s*Function : ", sass)
    for b in blocks[1:]:
        name = b.split("\n", 1)[0].strip()
        if pat not in name:
            continue
        cnt = collections.Counter()
        heavy = alu = 0.0
        n = 0
        for line in b.split("\n"):
            m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_.]+)\s*(.*?);", line)
            if not m:
                continue
            op, args = m.group(1), m.group(2)
            n += 1
            base = op.split(".")[0]
            key = base
            if op.startswith("IMAD.WIDE") or op.startswith("IMAD.HI"):
                k = "IMAD.WIDE" if op.startswith("IMAD.WIDE") else "IMAD.HI"
                rz = args.rstrip().endswith("RZ")
                heavy += HEAVY[k][0 if rz else 1]
                key = k + (".rz" if rz else ".acc")
            elif base == "IMAD":
                heavy += 2.05
                sub = op.split(".")[1] if "." in op else ""
                key = "IMAD." + sub if sub in ("X", "MOV", "SHL", "IADD") else "IMAD"
            elif base in ALU:
                alu += 2.0
            cnt[key] += 1
        print(f"{name}\n  instructions {n}  fmaheavy {heavy:.0f} clk  alu {alu:.0f} clk" +
              (f"  | per butterfly: inst {n / per:.1f} heavy {heavy / per:.1f} alu {alu / per:.1f}" if per else ""))
        print("  " + "  ".join(f"{k}:{v}" for k, v in cnt.most_common(24)))


if __name__ == "__main__":
    main()
