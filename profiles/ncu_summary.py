"""Summarise an ncu report: per-kernel headline metrics + per-source-line hot spots.
usage: python profiles/ncu_summary.py report.ncu-rep [--lines N]"""
import csv, io, subprocess, sys, collections

def run(args):
    return subprocess.run(["ncu", "-i"] + args, capture_output=True, text=True).stdout

def main():
    rep = sys.argv[1]
    nlines = int(sys.argv[sys.argv.index("--lines") + 1]) if "--lines" in sys.argv else 25
    raw = list(csv.reader(io.StringIO(run([rep, "--page", "raw", "--csv"]))))
    hdr = raw[0]
    want = ["Kernel Name", "launch__grid_size", "launch__block_size", "gpu__time_duration.sum", "launch__registers_per_thread",
            "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.max"]
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            print("%-70s %s" % (w, [r[i][:34] for r in raw[2:]]))
    src = list(csv.reader(io.StringIO(run([rep, "--page", "source", "--csv"]))))
    kern, cur = [], None
    for r in src:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            kern.append(cur)
        elif cur is not None:
            cur["rows"].append(r)
    seen = set()
    for k in kern:
        h, data = k["rows"][0], k["rows"][1:]
        ci = {x: i for i, x in enumerate(h)}
        stall = [x for x in h if x.startswith("stall_") and "Not Issued" not in x]
        tot, ops, nsamp, ninst = collections.Counter(), collections.Counter(), 0, 0
        for r in data:
            try:
                s, e = int(r[ci["# Samples"]]), int(r[ci["Instructions Executed"]])
            except Exception:
                continue
            nsamp += s
            ninst += e
            t = r[ci["Source"]].split()
            op = (t[1] if t and t[0].startswith("@") else (t[0] if t else "?")).split(".")[0]
            ops[op] += e
            for c in stall:
                try:
                    tot[c] += int(r[ci[c]])
                except Exception:
                    pass
        key = (k["name"], ninst)
        if key in seen:
            continue
        seen.add(key)
        print("\n=====", k["name"][:90], "| warp instructions", ninst, "| samples", nsamp)
        print("stalls %:", [(c, round(100.0 * v / max(nsamp, 1), 1)) for c, v in tot.most_common(8)])
        print("op mix %:", [(o, round(100.0 * v / max(ninst, 1), 1)) for o, v in ops.most_common(16)])
        top = sorted(data, key=lambda r: -int(r[ci["# Samples"]]) if r[ci["# Samples"]].isdigit() else 0)[:nlines]
        for r in top:
            print("  %7s samples %12s exec  %s" % (r[ci["# Samples"]], r[ci["Instructions Executed"]], r[ci["Source"]][:80]))

if __name__ == "__main__":
    main()
