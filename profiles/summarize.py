#!/usr/bin/env python
"""Turns the ncu outputs brought back in gpurun_out/ into the small text summaries committed under profiles/.

    python profiles/summarize.py launches gpurun_out/launches_r1.csv  > profiles/launches_r1.txt
    python profiles/summarize.py raw      gpurun_out/prof_mlp_r1.ncu-rep > profiles/mlp_r1_metrics.txt
    python profiles/summarize.py source   gpurun_out/prof_mlp_r1.ncu-rep > profiles/mlp_r1_hot_sass.txt
    python profiles/summarize.py stages   gpurun_out/stages_render_r2a.csv > profiles/stages_render_r2_metrics.txt
"""
import collections
import csv
import io
import subprocess
import sys

KEEP = ("gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct", "lts__throughput.avg.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.per_cycle_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "sm__cycles_elapsed.avg",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct", "sm__throughput.avg.pct",
        "smsp__average_warps_issue_stalled", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.avg.per_second",
        "smsp__average_warp_latency_per_inst_issued")


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def launches(path):
    rows = list(csv.reader(open(path)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    H, data = rows[hdr], rows[hdr + 1:]
    ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
    scale = {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}
    agg = collections.OrderedDict()
    for r in data:
        if len(r) <= vi:
            continue
        k = r[ki].split("(")[0]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += float(r[vi].replace(",", "")) * scale[r[ui]]
    tot = sum(v[1] for v in agg.values())
    print("# ncu --metrics gpu__time_duration.sum launch list: %s (cold-cache, serialised: compare shares)" % path)
    print("%-60s %6s %12s %7s" % ("kernel", "n", "total ms", "share"))
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-60s %6d %12.3f %6.1f%%" % (k[:60], v[0], v[1], 100 * v[1] / tot))
    print("%-60s %6d %12.3f" % ("TOTAL", sum(v[0] for v in agg.values()), tot))


def stages(path):
    """Launch list taken with --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum: per kernel the
    number of launches, mean duration and mean DRAM bytes read / written per launch, and the DRAM rate they imply."""
    rows = list(csv.reader(open(path)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    H, data = rows[hdr], rows[hdr + 1:]
    ii, ki, mi, vi, ui = H.index("ID"), H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value"), H.index("Metric Unit")
    tscale = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6}
    bscale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per = collections.OrderedDict()            # launch id -> {kernel, us, rd, wr}
    for r in data:
        if len(r) <= vi:
            continue
        d = per.setdefault(r[ii], {"k": r[ki].split("(")[0], "us": 0.0, "rd": 0.0, "wr": 0.0})
        v = float(r[vi].replace(",", ""))
        if r[mi].startswith("gpu__time_duration"):
            d["us"] = v * tscale[r[ui]]
        elif r[mi].startswith("dram__bytes_read"):
            d["rd"] = v * bscale[r[ui]]
        elif r[mi].startswith("dram__bytes_write"):
            d["wr"] = v * bscale[r[ui]]
    agg = collections.OrderedDict()
    for d in per.values():
        a = agg.setdefault(d["k"], [0, 0.0, 0.0, 0.0])
        a[0] += 1; a[1] += d["us"]; a[2] += d["rd"]; a[3] += d["wr"]
    print("# ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none: %s" % path)
    print("# per launch means (cold-cache, serialised launches); DRAM GB/s = (read + write) / duration")
    print("%-56s %5s %10s %12s %12s %10s" % ("kernel", "n", "us", "dram rd MB", "dram wr MB", "GB/s"))
    for k, (n, us, rd, wr) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-56s %5d %10.2f %12.3f %12.3f %10.1f" % (k[:56], n, us / n, rd / n / 1e6, wr / n / 1e6, (rd + wr) / max(us, 1e-9) / 1e3))


def raw(rep):
    rows = ncu_csv(rep, "raw")
    H, U, data = rows[0], rows[1], rows[2:]
    ki = H.index("Kernel Name")
    print("# ncu --set full --page raw: %s" % rep)
    print("# launches: " + " | ".join("%s grid=%s" % (r[ki].split("(")[0], r[H.index("Grid Size")]) for r in data))
    for i, h in enumerate(H):
        if any(h.startswith(k) for k in KEEP):
            print("%-90s %-10s %s" % (h, U[i], " | ".join(r[i] for r in data)))


def source(rep, top=40):
    rows = ncu_csv(rep, "source", ["--launch-skip", "0", "--launch-count", "1"])
    H, data = rows[1], rows[2:]
    isrc, isamp, iex = H.index("Source"), H.index("# Samples"), H.index("Instructions Executed")
    stall = [i for i, h in enumerate(H) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[isamp]) for r in data) or 1
    print("# ncu source page (first captured launch): %d SASS instructions, %d samples, %d warp-instructions executed"
          % (len(data), tot, sum(int(r[iex]) for r in data)))
    tot_by = collections.Counter()
    for r in data:
        for i in stall:
            tot_by[H[i][6:]] += int(r[i])
    print("# stall samples by reason: " + ", ".join("%s %.1f%%" % (k, 100 * v / tot) for k, v in tot_by.most_common(8)))
    for r in sorted(data, key=lambda r: -int(r[isamp]))[:top]:
        st = sorted(((int(r[i]), H[i][6:]) for i in stall), reverse=True)[:2]
        print("%6.2f%%  exec=%10s  sass#%-5d %-64s %s" % (100 * int(r[isamp]) / tot, r[iex], data.index(r), r[isrc].strip()[:64],
                                                          ", ".join("%s=%d" % (n, c) for c, n in st if c)))


def lines(rep, launch="0", top=40):
    """CUDA source lines by stall samples (needs -lineinfo + --import-source on)."""
    rows = ncu_csv(rep, "source", ["--launch-skip", launch, "--launch-count", "1", "--print-source", "cuda,sass"])
    path, out, H = "?", [], None
    for r in rows:
        if r and r[0] == "File Path":
            path = r[1].split("/")[-1]
        elif r and r[0] == "Function Name":
            fn = r[1]
        elif r and r[0] == "Line No":
            H = r
        elif H and r and r[0]:
            n = r[H.index("# Samples")]
            out.append((int(n) if n.isdigit() else 0, path, r[0], r[1].strip()))
    tot = sum(o[0] for o in out) or 1
    print("# ncu source page, CUDA lines by samples (launch %s: %s): %d samples" % (launch, fn[:70], tot))
    for n, f, ln, src in sorted(out, reverse=True)[:int(top)]:
        print("%6.2f%%  %s:%s  %s" % (100 * n / tot, f, ln, src[:110]))


if __name__ == "__main__":
    {"launches": launches, "stages": stages, "raw": raw, "source": source, "lines": lines}[sys.argv[1]](*sys.argv[2:])
