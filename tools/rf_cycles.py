#!/usr/bin/env python
"""Static issue-cycle estimate of a SASS range under the register-file rule measured by tools/rf_probe*.cu: an
instruction costs max(1, distinct even source registers, distinct odd source registers) cycles, where a source operand
served by the operand-reuse cache (same register, same slot, flagged .reuse on the previous instruction that used the
slot) is free.  usage: rf_cycles.py file.sass first_line last_line   (1-based line numbers of the cuobjdump dump)"""
import re, sys

def operands(line):
    m = re.search(r"/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?(\S+)\s+(.*?);", line)
    if not m:
        return None
    op, rest = m.group(2), m.group(3)
    parts = [p.strip() for p in rest.split(",")]
    return op, parts

def main(path, a, b):
    lines = open(path).read().splitlines()[a - 1:b]
    reuse = {}          # slot -> register held
    total = n = conflicts = 0
    hist = {}
    for ln in lines:
        o = operands(ln)
        if not o:
            continue
        op, parts = o
        n += 1
        srcs = parts[1:]
        if op.startswith(("LDS", "STS", "LDG", "STG", "BRA", "UI", "UMOV", "ULD", "UISETP", "BAR", "LDC", "S2R")):
            total += 1
            hist[op.split(".")[0]] = hist.get(op.split(".")[0], 0) + 1
            continue
        fetch = set()
        new_reuse = {}
        for slot, s in enumerate(srcs):
            m = re.match(r"[-|~]*R(\d+)(\.reuse)?", s)
            if not m:
                continue
            r = int(m.group(1))
            if reuse.get(slot) != r:
                fetch.add(r)
            if m.group(2):
                new_reuse[slot] = r
        reuse = new_reuse
        ev = sum(1 for r in fetch if r % 2 == 0)
        od = len(fetch) - ev
        c = max(1, ev, od)
        conflicts += c - 1
        total += c
        hist[op.split(".")[0]] = hist.get(op.split(".")[0], 0) + 1
    print(f"instructions {n}, estimated issue cycles {total}, extra cycles from register-bank conflicts {conflicts}; mix {hist}")

if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]), int(sys.argv[3]))
