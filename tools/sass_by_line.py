#!/usr/bin/env python
"""Attribute the executed warp-instructions of one kernel in an .ncu-rep to CUDA source lines.

  python tools/sass_by_line.py <ncu --page source --csv dump> <nvdisasm -gi section of the same kernel> [top]

The ncu SASS page and nvdisasm list the same instructions in the same order; nvdisasm -g carries `//## File ... line N`
(+ `inlined at`) markers from -lineinfo. Prints the share of executed warp-instructions per source function-level bucket
(outermost non-inlined frame = line in the kernel body) and per innermost line.
"""
import csv, re, sys, collections, os

rows = list(csv.reader(open(sys.argv[1])))[2:]
execs = [int(r[5]) for r in rows]
thr = [int(r[6]) for r in rows]
samples = [int(r[4]) for r in rows]
inst_loc = []
chain, in_group = [], False
pat = re.compile(r'//## File "([^"]+)", line (\d+)')
for ln in open(sys.argv[2]):
    m = pat.search(ln)
    if m:  # nvdisasm -gi: one marker line per frame, innermost first; a run of marker lines is one location
        if not in_group:
            chain, in_group = [], True
        chain.append((os.path.basename(m.group(1)), int(m.group(2))))
        continue
    if re.search(r'/\*[0-9a-f]{4,}\*/\s', ln):
        inst_loc.append(tuple(chain))
        in_group = False
assert len(inst_loc) == len(execs), (len(inst_loc), len(execs))
total = sum(execs)
inner = collections.Counter(); outer = collections.Counter(); outer_thr = collections.Counter(); outer_smp = collections.Counter()
for loc, e, t, s in zip(inst_loc, execs, thr, samples):
    if not loc:
        loc = (("?", 0),)
    inner[loc[0]] += e
    outer[loc[-1]] += e; outer_thr[loc[-1]] += t; outer_smp[loc[-1]] += s
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
print(f"total executed warp-instructions {total}")
print("-- by kernel-body line (outermost frame): share of instructions, avg active threads, share of stall samples")
ts = sum(samples)
for loc, e in outer.most_common(top):
    print(f"  {loc[0]}:{loc[1]:<5d} {100.0*e/total:6.2f}%  {outer_thr[loc]/max(e,1):5.1f} thr  {100.0*outer_smp[loc]/ts:6.2f}% smp")
print("-- by innermost line")
for loc, e in inner.most_common(top):
    print(f"  {loc[0]}:{loc[1]:<5d} {100.0*e/total:6.2f}%")
