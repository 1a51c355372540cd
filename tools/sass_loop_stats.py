#!/usr/bin/env python3
"""Static SASS statistics of a kernel's hottest loop, computed here (no GPU): opcode histogram, issue-slot count and the
sum of the per-instruction stall fields (the minimum cycles ONE warp needs per trip when nothing else stalls it).

    python tools/sass_loop_stats.py OBJECT_OR_SO KERNEL_SUBSTRING [--loop N] [--all]

The loop is found as the backward branch with the largest body (--loop N picks the N-th largest).  Control fields follow
the sm_70+ encoding (bits 105-108 stall count, 109 yield, 110-112 write barrier, 113-115 read barrier, 116-121 wait mask).
"""
import collections
import re
import subprocess
import sys


def disasm(path, kernel):
    out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True, check=True).stdout
    funcs, cur = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        if cur is not None:
            funcs[cur].append(line)
    hits = [f for f in funcs if kernel in f]
    if not hits:
        raise SystemExit(f"no kernel matching {kernel!r}; have: " + ", ".join(funcs))
    return hits[0], funcs[hits[0]]


def parse(lines):
    ins = []
    for i, line in enumerate(lines):
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\* 0x([0-9a-f]{16}) \*/", line)
        if not m:
            continue
        hi = 0
        if i + 1 < len(lines):
            m2 = re.match(r"\s*/\* 0x([0-9a-f]{16}) \*/", lines[i + 1])
            if m2:
                hi = int(m2.group(1), 16)
        text = m.group(2).strip()
        pred = ""
        t = text
        if t.startswith("@"):
            pred, t = t.split(None, 1)
        op = t.split()[0].rstrip(";")
        ins.append(dict(addr=int(m.group(1), 16), text=text, op=op, base=op.split(".")[0], stall=(hi >> 41) & 0xF,
                        yld=(hi >> 45) & 1, wbar=(hi >> 46) & 7, rbar=(hi >> 49) & 7, wait=(hi >> 52) & 0x3F))
    return ins


def loops(ins):
    by_addr = {x["addr"]: k for k, x in enumerate(ins)}
    res = []
    for k, x in enumerate(ins):
        if x["base"] == "BRA":
            m = re.search(r"0x([0-9a-f]+)", x["text"])
            if m and int(m.group(1), 16) in by_addr and by_addr[int(m.group(1), 16)] <= k:
                res.append((by_addr[int(m.group(1), 16)], k))
    return sorted(res, key=lambda ab: ab[0] - ab[1])


def main():
    path, kernel = sys.argv[1], sys.argv[2]
    nth = int(sys.argv[sys.argv.index("--loop") + 1]) if "--loop" in sys.argv else 0
    name, lines = disasm(path, kernel)
    ins = parse(lines)
    print(f"# {name}: {len(ins)} SASS instructions")
    ls = loops(ins)
    if "--all" in sys.argv or not ls:
        body = ins
        print("# whole kernel")
    else:
        a, b = ls[nth]
        body = ins[a:b + 1]
        print(f"# loop {nth}: 0x{ins[a]['addr']:x} .. 0x{ins[b]['addr']:x}: {len(body)} instructions (other loops: "
              + ", ".join(str(q - p + 1) for p, q in ls[:6]) + ")")
    hist = collections.Counter(x["base"] for x in body)
    full = collections.Counter(x["op"] for x in body)
    stall = sum(max(x["stall"], 1) for x in body)
    print(f"issue slots {len(body)}, sum of stall fields {stall} cycles ({stall / len(body):.2f} per instruction), "
          f"instructions that wait on a scoreboard: {sum(1 for x in body if x['wait'])}")
    print("opcodes: " + ", ".join(f"{k} {v}" for k, v in hist.most_common()))
    if "--full" in sys.argv:
        print("variants: " + ", ".join(f"{k} {v}" for k, v in full.most_common()))
    if "--dump" in sys.argv:
        for x in body:
            print(f"  {x['addr']:05x} s{x['stall']:<2} y{x['yld']} w{x['wbar']} r{x['rbar']} m{x['wait']:02x}  {x['text']}")


if __name__ == "__main__":
    main()
