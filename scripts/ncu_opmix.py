#!/usr/bin/env python
"""Dynamic opcode mix per kernel from `ncu -i rep --page source --csv --print-source sass > x.csv`.
Prints warp-instruction counts per opcode class and the pipe demand under the cost model of DESIGN.md section 3
(cycles per warp instruction per SM sub-partition: ALU class 2, IMAD 2, IMAD.WIDE / IMAD.HI 4, FFMA / FADD / FMUL 1,
MUFU 8) as sub-partition cycles summed over the grid; divide by 592 sub-partitions x clock for a time."""
import collections
import csv
import re
import sys

FMA1 = ("FFMA", "FADD", "FMUL", "FFMA2", "FADD2", "FMUL2", "HFMA2", "HADD2", "HMUL2")
FMA2 = ("IMAD",)
ALU = ("LOP3", "PRMT", "SHF", "IADD3", "VIADD", "ISETP", "SEL", "VIMNMX", "VIMNMX3", "VABSDIFF4", "IABS", "LEA", "MOV", "PLOP3",
       "FMNMX", "FSETP", "FSEL", "IDP", "VIADDMNMX", "I2FP", "CS2R", "IMNMX", "LOP", "SGXT", "BMSK", "FMNMX3", "P2R", "R2P", "FLO", "POPC")


def klass(op):
    p = op.split(".")
    if p[0] == "IMAD" and len(p) > 1 and p[1] in ("WIDE", "HI"):
        return "IMAD." + p[1]
    if p[0] in ("MUFU", "LDS", "STS", "LDG", "STG", "LDGSTS"):
        return p[0]
    return p[0]


def main():
    kernel, hdr, mix = None, None, collections.OrderedDict()
    for r in csv.reader(open(sys.argv[1])):
        if not r:
            continue
        if r[0] == "Kernel Name":
            kernel = r[1]
            mix[kernel] = collections.Counter()
            continue
        if r[0] == "Address":
            hdr = r
            ii, si = hdr.index("Instructions Executed"), hdr.index("Source")
            continue
        if hdr is None or len(r) <= ii or not r[ii].isdigit():
            continue
        ins = re.sub(r"^@!?U?P\d+\s+", "", r[si].strip())
        if ins:
            mix[kernel][klass(ins.split()[0])] += int(r[ii])
    for k, c in mix.items():
        tot = sum(c.values())
        alu = sum(v for o, v in c.items() if o in ALU) * 2
        fma = sum(v * (4 if o in ("IMAD.WIDE", "IMAD.HI") else 2 if o in FMA2 else 1 if o in FMA1 else 0) for o, v in c.items())
        xu = c.get("MUFU", 0) * 8
        print("==", k[:90])
        print("   warp instructions %.1f M; pipe demand (M sub-partition cycles): ALU %.1f  FMA %.1f  XU %.1f  issue %.1f" %
              (tot / 1e6, alu / 1e6, fma / 1e6, xu / 1e6, tot / 1e6))
        print("   at 592 sub-partitions x 1.965 GHz: ALU %.3f ms  FMA %.3f ms  XU %.3f ms  issue %.3f ms" %
              tuple(x / 592 / 1.965e6 for x in (alu, fma, xu, tot)))
        print("   " + "  ".join("%s %.1f" % (o, v / 1e6) for o, v in c.most_common(24)))


if __name__ == "__main__":
    main()
