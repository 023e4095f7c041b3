#!/usr/bin/env python
"""Summarise an .ncu-rep (one or more captured launches) into the text that goes under profiles/.

  python tools/ncu_summary.py gpurun_out/x.ncu-rep [--hot N] > profiles/rNN_x.txt

Prints, per captured launch: the raw-page metrics the roofline cites (duration, issue utilisation, active threads per
warp instruction, pipe utilisation, occupancy, DRAM/L2 traffic), the warp-stall breakdown summed over the SASS, and the
hottest contiguous SASS segments (share of executed instructions, share of stall samples, average active threads).
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps", "sm__maximum_warps_per_active_cycle_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_active.avg.per_cycle_active",
    "smsp__warps_eligible.avg.per_cycle_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "smsp__thread_inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__thread_inst_executed_pred_on_per_inst_executed.ratio",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__sass_thread_inst_executed_op_fp32_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
    "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_global_ld.sum", "smsp__inst_executed_op_global_st.sum",
    "smsp__inst_executed_op_global_red.sum",
]


def page(rep, which, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", which, "--csv", *extra], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                         text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def raw_summary(rep):
    rows = page(rep, "raw")
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, zip(units, r)))
        print(f"== launch id {d['ID'][1]}: {d['Kernel Name'][1]}  grid {d.get('Grid Size', ('', '?'))[1]} block {d.get('Block Size', ('', '?'))[1]}")
        for k in KEYS:
            if k in d and d[k][1] != "":
                print(f"   {k:68s} {d[k][1]:>18s} {d[k][0]}")


def sass_summary(rep, hot):
    rows = page(rep, "source", ["--print-source", "sass"])
    # the page repeats a 2-line header per kernel
    blocks, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "hdr": None, "data": []}
            blocks.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = r
        elif cur is not None and r:
            cur["data"].append(r)
    for b in blocks:
        hdr, data = b["hdr"], b["data"]
        ix = {k: i for i, k in enumerate(hdr)}
        f = lambda r, k: float(r[ix[k]] or 0)
        ti = sum(f(r, "Instructions Executed") for r in data)
        tt = sum(f(r, "Thread Instructions Executed") for r in data)
        ts = sum(f(r, "# Samples") for r in data)
        print(f"\n== SASS of {b['name']}: {len(data)} instructions, {ti:.0f} warp-instructions executed, "
              f"{tt / max(ti, 1):.2f} active threads per warp-instruction, {ts:.0f} stall samples")
        st = [k for k in hdr if k.startswith("stall_") and "Not Issued" not in k]
        tot = {k: sum(f(r, k) for r in data) for k in st}
        s = sum(tot.values()) or 1
        print("   warp stall samples: " + ", ".join(f"{k[6:]} {v / s * 100:.1f}%" for k, v in sorted(tot.items(), key=lambda x: -x[1]) if v / s > 0.01))
        ops = {}
        for r in data:
            toks = [o for o in r[ix["Source"]].split() if not o.startswith("@")]
            op = toks[0].split(".")[0] if toks else "?"
            ops[op] = ops.get(op, 0) + f(r, "Instructions Executed")
        print("   opcode mix (executed): " + ", ".join(f"{k} {v / ti * 100:.1f}%" for k, v in sorted(ops.items(), key=lambda x: -x[1])[:16]))
        seg, c = [], None
        for i, r in enumerate(data):
            ie = f(r, "Instructions Executed")
            if c and abs(ie - c["ie0"]) <= 0.12 * max(c["ie0"], 1):
                c["n"] += 1; c["ie"] += ie; c["te"] += f(r, "Thread Instructions Executed"); c["s"] += f(r, "# Samples"); c["end"] = i
            else:
                c = {"start": i, "end": i, "n": 1, "ie0": ie, "ie": ie, "te": f(r, "Thread Instructions Executed"), "s": f(r, "# Samples")}
                seg.append(c)
        print("   hottest SASS segments (contiguous runs with equal execution count):")
        for c in sorted(seg, key=lambda c: -c["ie"])[:hot]:
            print(f"     [{c['start']:4d}-{c['end']:4d}] {c['n']:3d} instr x {c['ie0'] / 1e6:8.3f} M  = {c['ie'] / ti * 100:5.1f}% of instructions, "
                  f"{c['s'] / max(ts, 1) * 100:5.1f}% of samples, {c['te'] / max(c['ie'], 1):5.1f} threads active | {data[c['start']][ix['Source']].strip()[:48]}")


if __name__ == "__main__":
    rep = sys.argv[1]
    hot = int(sys.argv[sys.argv.index("--hot") + 1]) if "--hot" in sys.argv else 10
    print(f"# ncu summary of {rep} (tools/ncu_summary.py)")
    raw_summary(rep)
    sass_summary(rep, hot)
