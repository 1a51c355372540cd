#!/usr/bin/env python3
"""Source-page summary of one kernel of an `ncu --set full --import-source on` report, computed here (no GPU):

    ncu -i REPORT.ncu-rep --page source --csv --kernel-name regex:NAME > src.csv
    python tools/ncu_source_summary.py src.csv [--launch K] [--top N] > profiles/NAME_source_summary.txt

Prints the stall-reason shares of the warp-state samples, the executed-instruction opcode histogram (warp instructions,
i.e. weighted by how often each SASS line ran), the hot loop (the most executed instructions: count per trip, share of
all executed instructions and of the samples) and the lines with the most samples."""
import argparse
import collections
import csv
import re

ap = argparse.ArgumentParser()
ap.add_argument("csv")
ap.add_argument("--launch", type=int, default=-1, help="which captured launch of the kernel (default: the last)")
ap.add_argument("--top", type=int, default=15)
ap.add_argument("--px", type=float, default=0.0, help="pixels (or other units) per launch: prints lane-instructions per unit")
a = ap.parse_args()

rows = list(csv.reader(open(a.csv)))
heads = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
names = [rows[i - 1][1] if i > 0 and len(rows[i - 1]) > 1 else "?" for i in heads]
h = heads[a.launch]
hdr = rows[h]
end = heads[heads.index(h) + 1] - 1 if heads.index(h) + 1 < len(heads) else len(rows)
body = [r for r in rows[h + 1:end] if len(r) == len(hdr)]
ix = {k: i for i, k in enumerate(hdr)}
num = lambda r, k: int(float(r[ix[k]] or 0))  # noqa: E731
base = int(body[0][0], 16)
stalls = [k for k in hdr if k.startswith("stall_") and "Not Issued" not in k]
total_samples = sum(num(r, "# Samples") for r in body)
total_inst = sum(num(r, "Instructions Executed") for r in body)
thread_inst = sum(num(r, "Thread Instructions Executed") for r in body)
print(f"# source page of {names[heads.index(h)]}")
print(f"# {len(body)} SASS instructions, {total_inst} warp instructions executed, {thread_inst} thread instructions, {total_samples} warp-state samples")
if a.px:
    print(f"# {thread_inst / a.px:.1f} thread instructions per unit ({a.px:.0f} units per launch); {32.0 * total_inst / a.px:.1f} lane slots per unit")
print("\nstall reason, share of all samples (%)")
tot = collections.Counter()
for r in body:
    for k in stalls:
        tot[k] += num(r, k)
for k, v in tot.most_common():
    if v:
        print(f"{k[6:]},{100.0 * v / max(total_samples, 1):.1f}")


def opcode(src):
    t = re.sub(r"^@!?U?P\d+\s+", "", src.strip())
    return t.split()[0].split(".")[0] if t else "?"


ops = collections.Counter()
for r in body:
    ops[opcode(r[ix["Source"]])] += num(r, "Instructions Executed")
print("\nexecuted warp instructions by opcode (share %)")
print(", ".join(f"{k} {100.0 * v / max(total_inst, 1):.1f}" for k, v in ops.most_common(28)))

# the hot loop: instructions whose execution count is within 2 % of the most common large count
counts = collections.Counter(num(r, "Instructions Executed") for r in body)
big = [c for c, n in counts.items() if n >= 50 and c > 0]
if big:
    trip = max(big, key=lambda c: c * counts[c])
    loop = [r for r in body if abs(num(r, "Instructions Executed") - trip) <= 0.02 * trip]
    li = sum(num(r, "Instructions Executed") for r in loop)
    ls = sum(num(r, "# Samples") for r in loop)
    lo = collections.Counter(opcode(r[ix["Source"]]) for r in loop)
    print(f"\nhot loop: {len(loop)} SASS instructions executed {trip} times each = {100.0 * li / total_inst:.1f} % of all warp instructions, "
          f"{100.0 * ls / max(total_samples, 1):.1f} % of the samples")
    print("opcode counts of the hot loop: " + ", ".join(f"{k} {v}" for k, v in lo.most_common(40)))
    lst = collections.Counter()
    for r in loop:
        for k in stalls:
            lst[k] += num(r, k)
    print("stall shares inside the hot loop (%): " + ", ".join(f"{k[6:]} {100.0 * v / max(ls, 1):.1f}" for k, v in lst.most_common(8) if v))

print(f"\ntop {a.top} instructions by samples: samples, share %, offset, SASS, dominant stall")
for r in sorted(body, key=lambda r: -num(r, "# Samples"))[:a.top]:
    dom = max(stalls, key=lambda k: num(r, k))
    print(f"{num(r, '# Samples')},{100.0 * num(r, '# Samples') / max(total_samples, 1):.1f},0x{int(r[0], 16) - base:04x},{r[ix['Source']].strip()[:70]},{dom[6:]}")
