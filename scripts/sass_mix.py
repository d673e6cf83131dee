#!/usr/bin/env python
"""Static SASS opcode mix of one kernel: `sass_mix.py <obj|so> <mangled-name-substring> [--dump]`.
Counts instructions per opcode class over the whole function body (no loop weighting; the fused kernels'
hot loops are fully unrolled so the static mix is close to the dynamic one)."""
import collections
import re
import subprocess
import sys


def functions(path):
    out = subprocess.run(["cuobjdump", "-sass", path], stdout=subprocess.PIPE, text=True).stdout
    cur, body = None, {}
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            body[cur] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and cur:
            body[cur].append(m.group(2).strip())
    return body


def opcode(ins):
    ins = re.sub(r"^@!?U?P\d+\s+", "", ins)
    return ins.split()[0]


def klass(op):
    parts = op.split(".")
    if parts[0] in ("IMAD", "MUFU", "LDS", "STS", "LDG", "STG") and len(parts) > 1:
        return parts[0] + "." + parts[1]
    return parts[0]


def main():
    path, pat = sys.argv[1], sys.argv[2]
    for name, ins in functions(path).items():
        if pat not in name:
            continue
        ops = collections.Counter(klass(opcode(i)) for i in ins)
        print(name, len(ins), "instructions")
        for k, v in ops.most_common():
            print("  %-14s %5d" % (k, v))
        if "--dump" in sys.argv:
            print("\n".join(ins))


if __name__ == "__main__":
    main()
