"""Per-CUDA-source-line sample and instruction totals of one kernel of an ncu report.
usage: python profiles/ncu_lines.py report.ncu-rep <kernel substring> [top N]"""
import csv, io, subprocess, sys, collections

rep, sub = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur_file, cur_fn, take = None, None, False
agg = collections.OrderedDict()
tot_s = tot_i = 0
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        take = sub in r[1].replace("(int)", "").replace(" ", "")
        cur_fn = r[1]
        continue
    if r[0] == "Line No" or not take:
        continue
    if r[0].isdigit():  # a source line with its aggregated metrics
        try:
            s, e = int(r[6]), int(r[7])
        except ValueError:
            continue
        key = (cur_file, int(r[0]))
        a = agg.setdefault(key, [0, 0, r[1].strip()[:110]])
        a[0] += s
        a[1] += e
        tot_s += s
        tot_i += e
print("kernel filter %r: %d samples, %d warp instructions" % (sub, tot_s, tot_i))
for (f, ln), (s, e, src) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% smp %5.1f%% ins  %s:%d  %s" % (100.0 * s / max(tot_s, 1), 100.0 * e / max(tot_i, 1), f, ln, src))

if len(sys.argv) > 4 and sys.argv[4] == "phases":
    # phase totals for kernels_sweep7.cuh (line ranges located by markers in the source)
    import os, re
    srcp = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "acoss_b200", "csrc", "kernels_sweep7.cuh")
    lines = open(srcp).read().split("\n")
    def find(marker):
        for i, l in enumerate(lines):
            if marker in l:
                return i + 1
        return -1
    marks = [("setup", find("// ---- per-item setup")), ("stage1", find("// ---- stage 1 of period q")),
             ("stage2", find("// ---- stage 2 of period p")), ("loop+barriers", find("bool pre_done = false")),
             ("selection", find("// ---- selection: this warp")), ("after-selection", find("S7_TICK(tq3)")),
             ("export/emit", find("if (MODE == 0) {\n") if False else find("if (tid >= s_first && tid <= s_last)")), ("end", 10 ** 9)]
    ph = collections.OrderedDict()
    for (f, ln), (s, e, src) in agg.items():
        if f == "kernels_sweep7.cuh":
            name = "prologue"
            for n, start in marks:
                if start > 0 and ln >= start:
                    name = n
        elif f == "kernels_sweep4.cuh":
            name = "selection(sweep4.cuh)"
        else:
            name = "other:" + f
        a = ph.setdefault(name, [0, 0])
        a[0] += s
        a[1] += e
    for n, (s, e) in ph.items():
        print("%-28s %5.1f%% samples %5.1f%% instructions" % (n, 100.0 * s / tot_s, 100.0 * e / tot_i))
