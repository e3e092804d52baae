#!/usr/bin/env python
"""Opcode histogram of the hottest loop of one kernel (cuobjdump -sass of an object file).

    python tools/sass_loop.py build/obj/lockstep.o ILi2ELi2ELi8ELi2ELb1ELb1ELi128ELb0ELb0E [--dump]

Finds the backward branch with the longest body in the named function and counts instructions by
class: fp64 (DFMA/DMUL/DADD/DSETP/..., two issue cycles each on B200), memory, integer/other.
"""
import collections
import re
import subprocess
import sys


def main():
    obj, key = sys.argv[1], sys.argv[2]
    out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    funcs = out.split("Function : ")
    body = [f for f in funcs if key in f.split("\n", 1)[0]]
    if not body:
        raise SystemExit("no function matches " + key)
    lines = body[0].split("\n")
    ins = []
    for ln in lines:
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    addr_index = {a: i for i, (a, _) in enumerate(ins)}
    best = None
    for i, (a, t) in enumerate(ins):
        m = re.search(r"\bBRA\b.*?(0x[0-9a-f]+)", t)
        if m:
            tgt = int(m.group(1), 16)
            if tgt < a and tgt in addr_index:
                n = i - addr_index[tgt] + 1
                if best is None or n > best[0]:
                    best = (n, addr_index[tgt], i)
    if best is None:
        raise SystemExit("no loop found")
    n, lo, hi = best
    hist = collections.Counter()
    for a, t in ins[lo:hi + 1]:
        t = re.sub(r"^@!?U?P\d+\s+", "", t)
        hist[t.split()[0].split(".")[0]] += 1
    fp64 = sum(v for k, v in hist.items() if k in ("DFMA", "DMUL", "DADD", "DSETP", "DMNMX"))
    print("loop of %d instructions (0x%x..0x%x), total function %d" % (n, ins[lo][0], ins[hi][0], len(ins)))
    print("fp64 %d, other %d  -> issue cycles >= %d" % (fp64, n - fp64, 2 * fp64 + (n - fp64)))
    for k, v in hist.most_common():
        print("  %-10s %d" % (k, v))
    if "--dump" in sys.argv:
        for a, t in ins[lo:hi + 1]:
            print("%06x  %s" % (a, t))


if __name__ == "__main__":
    main()
