#!/usr/bin/env python
"""Generates (and proves, by the 0-1 principle on sorted inputs / brute force) the compare-exchange programs of the
separable 7 x 7 median (b2d_median7.cu) and writes them as straight-line code to
image_denoising_b200/csrc/b2d_median7_nets.inc.

The scheme (per 2 x 2 block of outputs, all lists sorted, see b2d_median7.cu for the geometry):
  PAIR   : two sorted 6-columns -> one sorted 12-list                      (merge 6 + 6)
  QUAD   : two 12-lists -> one sorted 24-list                              (merge 12 + 12)
  CORE   : 12-list + 24-list -> ranks 11..24 of the 36 core values         (pruned merge 12 + 24)
  ROW    : 14 survivors + sorted 6 outer-row values -> ranks 6..13 of 20   (pruned merge 14 + 6)
  MID2   : 8 survivors + sorted 6 core values of the private row -> ranks 6, 7 of 14 (pruned merge 8 + 6); the median of
           those 14 and the row's corner value c is min(max(c, m6), m7)
  SORT6  : Bose-Nelson sorting network for 6 values (12 compare-exchanges)
A program is a list of (op, dst, a, b): op 'min' / 'max' on value numbers (SSA), inputs are values 0 .. n_in - 1."""
import itertools
import os
import sys


class Prog:
    def __init__(self, n_in):
        self.n_in = n_in
        self.ops = []          # (kind, a, b) -> value number n_in + index

    def _emit(self, kind, a, b):
        self.ops.append((kind, a, b))
        return self.n_in + len(self.ops) - 1

    def ce(self, a, b):
        return self._emit("min", a, b), self._emit("max", a, b)

    def run(self, inputs):
        v = list(inputs)
        for kind, a, b in self.ops:
            v.append(min(v[a], v[b]) if kind == "min" else max(v[a], v[b]))
        return v

    def pruned(self, outputs):
        """drops every operation no output depends on; returns (new program, new output value numbers)"""
        live = set(outputs)
        for i in range(len(self.ops) - 1, -1, -1):
            if self.n_in + i in live:
                live.add(self.ops[i][1])
                live.add(self.ops[i][2])
        q = Prog(self.n_in)
        remap = {i: i for i in range(self.n_in)}
        for i, (kind, a, b) in enumerate(self.ops):
            if self.n_in + i in live:
                remap[self.n_in + i] = q._emit(kind, remap[a], remap[b])
        return q, [remap[o] for o in outputs]


def bose_nelson(n):
    pairs = []

    def pbracket(i, x, j, y):
        if x == 1 and y == 1:
            pairs.append((i, j))
        elif x == 1 and y == 2:
            pairs.append((i, j + 1))
            pairs.append((i, j))
        elif x == 2 and y == 1:
            pairs.append((i, j))
            pairs.append((i + 1, j))
        else:
            a = x // 2
            b = y // 2 if x % 2 else (y + 1) // 2
            pbracket(i, a, j, b)
            pbracket(i + a, x - a, j + b, y - b)
            pbracket(i + a, x - a, j, b)

    def pstar(i, m):
        if m > 1:
            a = m // 2
            pstar(i, a)
            pstar(i + a, m - a)
            pbracket(i, a, i + a, m - a)

    pstar(0, n)
    return pairs


def sort_net(p, vals):
    w = list(vals)
    for i, j in bose_nelson(len(w)):
        w[i], w[j] = p.ce(w[i], w[j])
    return w


def oe_merge(p, a, b):
    """Batcher's odd-even merge of two sorted lists of value numbers of any lengths (Knuth 5.3.4)"""
    if not a:
        return list(b)
    if not b:
        return list(a)
    if len(a) == 1 and len(b) == 1:
        lo, hi = p.ce(a[0], b[0])
        return [lo, hi]
    v = oe_merge(p, a[0::2], b[0::2])
    w = oe_merge(p, a[1::2], b[1::2])
    out = [v[0]]
    # z = v1, (w1, v2), (w2, v3), ...  (1-based); leftovers of the longer list follow in order
    i = 0
    while i < len(w) and i + 1 < len(v):
        lo, hi = p.ce(w[i], v[i + 1])
        out += [lo, hi]
        i += 1
    out += w[i:] + v[i + 1:]
    return out


def sorted01(n):
    return [[0] * (n - k) + [1] * k for k in range(n + 1)]


def check_merge(prog, outs, sizes, ranks):
    """0-1 principle for merging: every combination of sorted 0-1 inputs"""
    for combo in itertools.product(*[sorted01(s) for s in sizes]):
        flat = [x for c in combo for x in c]
        v = prog.run(flat)
        want = sorted(flat)
        got = [v[o] for o in outs]
        assert got == [want[r] for r in ranks], (sizes, ranks)


def build(sizes, ranks, presort=()):
    """merge sorted lists of the given sizes (lists flagged in presort are sorted first) and keep `ranks`"""
    n = sum(sizes)
    p = Prog(n)
    lists, at = [], 0
    for k, s in enumerate(sizes):
        vals = list(range(at, at + s))
        at += s
        lists.append(sort_net(p, vals) if k in presort else vals)
    cur = lists[0]
    for nxt in lists[1:]:
        cur = oe_merge(p, cur, nxt)
    q, outs = p.pruned([cur[r] for r in ranks])
    return q, outs


def count(prog):
    return len(prog.ops)


def emit(name, prog, outs, f):
    f.write("// %s: %d inputs, %d min / max operations, %d outputs\n" % (name, prog.n_in, len(prog.ops), len(outs)))
    f.write("#define B2D_NET_%s(IN, OUT) do { \\\n" % name)
    for i, (kind, a, b) in enumerate(prog.ops):
        def ref(x):
            return "IN[%d]" % x if x < prog.n_in else "t%d_" % (x - prog.n_in)
        f.write("  const uint32_t t%d_ = %s(%s, %s); \\\n" % (i, "VMIN" if kind == "min" else "VMAX", ref(a), ref(b)))
    for k, o in enumerate(outs):
        f.write("  OUT[%d] = %s; \\\n" % (k, "IN[%d]" % o if o < prog.n_in else "t%d_" % (o - prog.n_in)))
    f.write("} while (0)\n\n")


def main():
    import random
    nets = {}
    # SORT6
    p = Prog(6)
    outs = sort_net(p, list(range(6)))
    for bits in itertools.product((0, 1), repeat=6):
        v = p.run(bits)
        assert [v[o] for o in outs] == sorted(bits)
    nets["SORT6"] = (p, outs)
    specs = {
        "MERGE33": ((3, 3), list(range(6))),   # SORT6 in the kernel = two 3-sorts (min3, max3, sum - min - max) + this
        "PAIR": ((6, 6), list(range(12))),
        "QUAD": ((12, 12), list(range(24))),
        "CORE": ((12, 24), list(range(11, 25))),
        "ROW": ((14, 6), list(range(6, 14))),
        "MID2": ((8, 6), [6, 7]),
    }
    for name, (sizes, ranks) in specs.items():
        q, outs = build(sizes, ranks)
        check_merge(q, outs, sizes, ranks)
        nets[name] = (q, outs)
    for name, (q, outs) in nets.items():
        print("%-6s %3d inputs  %3d ops" % (name, q.n_in, count(q)))
    n = {k: count(v[0]) for k, v in nets.items()}
    per_block = (2 * n["SORT6"] + n["PAIR"]) + n["QUAD"] / 2 + n["CORE"] + 2 * (n["SORT6"] + n["ROW"]) + \
        2 * (n["SORT6"] + 2 * (n["MID2"] + 2)) - n["SORT6"]
    print("min / max operations per 2 x 2 block: %.0f = %.1f per output (forgetful selection: 2 x 181)" %
          (per_block, per_block / 4))
    # whole scheme on random blocks against the plain median
    rnd = random.Random(5)
    for trial in range(300):
        img = [[rnd.randrange(0, 256 if trial % 2 else 4) for _ in range(8)] for _ in range(8)]   # rows y-3..y+4, cols X-3..X+4
        def run(name, vals):
            q, outs = nets[name]
            v = q.run(vals)
            return [v[o] for o in outs]
        cols = [run("SORT6", [img[r][c] for r in range(1, 7)]) for c in range(8)]      # shared rows y-2..y+3 of every column
        pairs = [run("PAIR", cols[c] + cols[c + 1]) for c in (1, 3, 5)]
        quad = run("QUAD", pairs[1] + pairs[2])
        core = run("CORE", pairs[0] + quad)
        for oy in (0, 1):
            outer = run("SORT6", [img[0 if oy == 0 else 7][c] for c in range(1, 7)])
            surv = run("ROW", core + outer)
            for ox in (0, 1):
                pc = 0 if ox == 0 else 7
                m6, m7 = run("MID2", surv + cols[pc])
                med = min(max(img[0 if oy == 0 else 7][pc], m6), m7)
                win = sorted(img[r][c] for r in range(oy, oy + 7) for c in range(ox, ox + 7))
                assert med == win[24], (trial, oy, ox)
    print("scheme verified on 300 random 8 x 8 neighbourhoods")
    nets.update(main5())
    if len(sys.argv) > 1 and sys.argv[1] == "--write":
        with open(INC_PATH, "w") as f:
            f.write(render(nets))
        print("wrote", INC_PATH)
    return nets


def main5():
    """the same scheme for the 5 x 5 median: core 4 x 4 (two row pairs, no 24-list), 16 -> 10 -> (+4) 14 -> 6 -> (+4) two
    middle values and the corner"""
    import random
    nets = {}
    p = Prog(4)
    outs = sort_net(p, list(range(4)))
    for bits in itertools.product((0, 1), repeat=4):
        v = p.run(bits)
        assert [v[o] for o in outs] == sorted(bits)
    nets["SORT4"] = (p, outs)
    specs = {
        "PAIR5": ((4, 4), list(range(8))),
        "CORE5": ((8, 8), list(range(3, 13))),
        "ROW5": ((10, 4), list(range(4, 10))),
        "MID5": ((6, 4), [4, 5]),
    }
    for name, (sizes, ranks) in specs.items():
        q, outs = build(sizes, ranks)
        check_merge(q, outs, sizes, ranks)
        nets[name] = (q, outs)
    for name, (q, outs) in nets.items():
        print("%-6s %3d inputs  %3d ops" % (name, q.n_in, count(q)))
    n = {k: count(v[0]) for k, v in nets.items()}
    per_block = (2 * n["SORT4"] + n["PAIR5"]) + n["CORE5"] + 2 * (n["SORT4"] + n["ROW5"]) + \
        2 * (n["SORT4"] + 2 * (n["MID5"] + 2)) - n["SORT4"]
    print("5 x 5: min / max operations per 2 x 2 block: %.0f = %.1f per output (forgetful selection: 2 x 79)" %
          (per_block, per_block / 4))
    rnd = random.Random(6)
    for trial in range(300):
        img = [[rnd.randrange(0, 256 if trial % 2 else 4) for _ in range(6)] for _ in range(6)]   # rows y-2..y+3, cols X-2..X+3
        def run(name, vals):
            q, outs = nets[name]
            v = q.run(vals)
            return [v[o] for o in outs]
        cols = [run("SORT4", [img[r][c] for r in range(1, 5)]) for c in range(6)]
        pairs = [run("PAIR5", cols[c] + cols[c + 1]) for c in (1, 3)]
        core = run("CORE5", pairs[0] + pairs[1])
        for oy in (0, 1):
            outer = run("SORT4", [img[0 if oy == 0 else 5][c] for c in range(1, 5)])
            surv = run("ROW5", core + outer)
            for ox in (0, 1):
                pc = 0 if ox == 0 else 5
                m4, m5 = run("MID5", surv + cols[pc])
                med = min(max(img[0 if oy == 0 else 5][pc], m4), m5)
                win = sorted(img[r][c] for r in range(oy, oy + 5) for c in range(ox, ox + 5))
                assert med == win[12], (trial, oy, ox)
    print("5 x 5 scheme verified on 300 random 6 x 6 neighbourhoods")
    return nets


INC_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "image_denoising_b200", "csrc",
                        "b2d_median7_nets.inc")


def render(nets):
    import io
    f = io.StringIO()
    f.write("// GENERATED by scripts/gen_median7_networks.py - do not edit.  Merge / selection programs of the separable\n"
            "// 7 x 7 median: every program is proven there (0-1 principle over all sorted inputs) before it is written.\n"
            "// VMIN / VMAX are defined by the including file (u16x2 minimum / maximum).\n\n")
    for name, (q, outs) in nets.items():
        emit(name, q, outs, f)
    return f.getvalue()


if __name__ == "__main__":
    main()
