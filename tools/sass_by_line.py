#!/usr/bin/env python
"""Attribute the executed warp-instructions of one kernel in an .ncu-rep to CUDA source lines.

  python tools/sass_by_line.py <capture.ncu-rep> <libcrt.so of the capture> [top]
  python tools/sass_by_line.py <ncu --page source --csv dump> <nvdisasm -gi section of the same kernel> [top]

The first form does the extraction itself: `ncu -i rep --page source --csv`, `cuobjdump -xelf all libcrt.so`,
`nvdisasm -gi crt_api.sm_100a.cubin`, and cuts the `.section .text.<kernel>` block of the kernel the capture names (the
library must be the build that was profiled: rebuild the commit in a scratch worktree if the tree has moved on).
The ncu SASS page and nvdisasm list the same instructions in the same order; nvdisasm -g carries `//## File ... line N`
(+ `inlined at`) markers from -lineinfo. Prints the share of executed warp-instructions per source function-level bucket
(outermost non-inlined frame = line in the kernel body) and per innermost line.
"""
import csv, re, sys, collections, os, subprocess, tempfile


def _extract(rep, lib):
    tmp = tempfile.mkdtemp(prefix="sass_by_line_")
    src = os.path.join(tmp, "source.csv")
    with open(src, "w") as f:
        subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=f, stderr=subprocess.DEVNULL, check=True)
    kernel = next(csv.reader(open(src)))[1]                       # e.g. void k_render_pipe<(int)2048, (bool)1>(PipeArgs)
    m = re.match(r"(?:void )?(\w+)(?:<(.*)>)?\(", kernel)
    name, targs = m.group(1), [a.split(")")[-1].strip() for a in (m.group(2) or "").split(",") if a.strip()]
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL, check=True)
    cubin = max((os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")), key=os.path.getsize)
    sass = subprocess.run(["nvdisasm", "-gi", cubin], capture_output=True, text=True, check=True).stdout.splitlines(keepends=True)
    # mangled template arguments: Li<N>E for ints, Lb<0|1>E for bools
    starts = [i for i, ln in enumerate(sass) if ln.lstrip().startswith(".section") and ".text." in ln]
    def matches(ln):
        if f"{len(name)}{name}" not in ln:
            return False
        nums = re.findall(r"L[ib](\d+)E", ln)
        return not targs or nums == targs
    hit = [i for i in starts if matches(sass[i])]
    if len(hit) != 1:
        raise SystemExit(f"sass_by_line: {len(hit)} sections match {kernel!r}")
    end = next((j for j in starts if j > hit[0]), len(sass))
    out = os.path.join(tmp, "kernel.sass")
    open(out, "w").writelines(sass[hit[0]:end])
    return src, out


if sys.argv[1].endswith(".ncu-rep"):
    sys.argv[1], sys.argv[2] = _extract(sys.argv[1], sys.argv[2])

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
if len(inst_loc) != len(execs):
    raise SystemExit(f"sass_by_line: the library has {len(inst_loc)} instructions in this kernel, the capture {len(execs)}: not the build that was profiled")
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
if len(sys.argv) > 4:  # breakdown of one kernel-body line by the second frame (the function called there) and by innermost line
    want = int(sys.argv[4])
    sub = collections.Counter(); sub_thr = collections.Counter()
    for loc, e, t in zip(inst_loc, execs, thr):
        if loc and loc[-1][1] == want:
            key = (loc[-2] if len(loc) > 1 else loc[-1], loc[0])
            sub[key] += e; sub_thr[key] += t
    print(f"-- inside kernel-body line {want}: (frame below it, innermost line): share of ALL instructions, avg threads")
    for key, e in sub.most_common(top):
        print(f"  {key[0][0]}:{key[0][1]:<5d} {key[1][0]}:{key[1][1]:<5d} {100.0*e/total:6.2f}%  {sub_thr[key]/max(e,1):5.1f} thr")
print("-- by innermost line")
for loc, e in inner.most_common(top):
    print(f"  {loc[0]}:{loc[1]:<5d} {100.0*e/total:6.2f}%")
