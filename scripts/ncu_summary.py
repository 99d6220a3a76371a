#!/usr/bin/env python
"""Summaries of ncu output for profiles/ (run here, no GPU needed).
  launches <launches.csv>            per-kernel count / total / average / share of device time
  full <prof.ncu-rep>                selected raw metrics per profiled launch
  traffic <prof.ncu-rep>             JSON {kernel: {dram_bytes, dram_read, dram_write, duration_us, ...}} of the FIRST captured launch of each
                                     kernel (dram__bytes_read.sum + dram__bytes_write.sum) - bench.py's roofline.traffic reads it; also the
                                     issue figures bench.py's roofline.issue reads (warp instructions, IPC, issue-slot utilisation)
  issue <prof.ncu-rep>               for every captured launch: issue-slot utilisation, IPC, per-pipe utilisation (ALU / FMA / XU / LSU /
                                     FP64 / ...) and the warp stall reasons, largest first - every column of the report that matches
  sass <lib.so> <kernel> [lo hi]     SASS of one kernel with an instruction histogram (cuobjdump; no GPU)
"""
import re
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "smsp__inst_executed.sum", "sm__inst_executed_pipe_fp64.sum", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
        "smsp__cycles_active.avg", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        k = row["Kernel Name"].split("(")[0]
        v = float(row["Metric Value"].replace(",", ""))
        a = agg.setdefault(k, [0, 0.0, 0.0])
        a[0] += 1; a[1] += v; a[2] = max(a[2], v)
    tot = sum(a[1] for a in agg.values())
    print("%-28s %7s %12s %10s %10s %7s" % ("kernel", "count", "total_us", "avg_us", "max_us", "share"))
    for k, (n, t, m) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-28s %7d %12.1f %10.2f %10.2f %6.1f%%" % (k[:28], n, t / 1e3, t / n / 1e3, m / 1e3, 100 * t / tot))
    print("total device time of the %d captured launches: %.1f us" % (sum(a[0] for a in agg.values()), tot / 1e3))


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    idx = {h: i for i, h in enumerate(hdr)}
    name_i = idx.get("Kernel Name")
    for r in rows[2:]:
        print("== %s  (ID %s)" % (r[name_i].split("(")[0], r[0]))
        for k in KEYS:
            if k in idx:
                print("   %-70s %s %s" % (k, r[idx[k]], rows[1][idx[k]]))


def traffic(path):
    import json
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    idx = {h: i for i, h in enumerate(rows[0])}
    unit = rows[1]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    res = collections.OrderedDict()
    for r in rows[2:]:
        k = r[idx["Kernel Name"]].split("(")[0]
        if k in res:
            continue
        rd = float(r[idx["dram__bytes_read.sum"]]) * scale[unit[idx["dram__bytes_read.sum"]]]
        wr = float(r[idx["dram__bytes_write.sum"]]) * scale[unit[idx["dram__bytes_write.sum"]]]
        tscale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6, "nsecond": 1e-3}
        res[k] = dict(dram_bytes=rd + wr, dram_read=rd, dram_write=wr,
                      duration_us=float(r[idx["gpu__time_duration.sum"]]) * tscale.get(unit[idx["gpu__time_duration.sum"]], 1.0))
        # the other levels of the hierarchy (BASELINE.json: "% HBM/L2 roofline"): L2 traffic (32 B sectors) and ncu's own
        # percent-of-peak figures for L2, L1/TEX and the SM
        for key, col, mul in (("lts_bytes", "lts__t_sectors.sum", 32.0), ("lts_pct_of_peak", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
                              ("l1tex_pct_of_peak", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
                              ("sm_pct_of_peak", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
                              ("l2_hit_pct", "lts__t_sector_hit_rate.pct", 1.0), ("l1_hit_pct", "l1tex__t_sector_hit_rate.pct", 1.0),
                              ("warp_inst", "smsp__inst_executed.sum", 1.0), ("ipc_active", "sm__inst_executed.avg.per_cycle_active", 1.0),
                              ("ipc_elapsed", "sm__inst_executed.avg.per_cycle_elapsed", 1.0),
                              ("issue_active_pct", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0),
                              ("warps_active_pct", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0)):
            if col in idx and r[idx[col]] not in ("", "n/a"):
                res[k][key] = float(r[idx[col]].replace(",", "")) * mul
    print(json.dumps(dict(source=path.split("/")[-1], how="ncu --set full --clock-control none, first captured launch of each kernel", kernels=res), indent=1))


ISSUE_PATTERNS = [r"issue_active", r"inst_executed\.avg\.per_cycle", r"inst_issued\.avg\.per_cycle", r"inst_executed\.sum$", r"thread_inst_executed",
                  r"sm__inst_executed_pipe_", r"sm__pipe_.*cycles_active", r"pipe_.*pct_of_peak", r"issue_stalled", r"warps_eligible",
                  r"warps_active", r"achieved_occupancy", r"sm__throughput", r"l1tex__throughput", r"lts__throughput", r"gpu__time_duration"]


def issue(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, unit = rows[0], rows[1]
    name_i = hdr.index("Kernel Name")
    cols = [i for i, h in enumerate(hdr) if any(re.search(p, h) for p in ISSUE_PATTERNS)]
    for r in rows[2:]:
        print("== %s  (ID %s)" % (r[name_i].split("(")[0], r[0]))
        stalls, rest = [], []
        for i in cols:
            if r[i] in ("", "n/a"):
                continue
            try:
                v = float(r[i].replace(",", ""))
            except ValueError:
                continue
            (stalls if "issue_stalled" in hdr[i] else rest).append((hdr[i], v, unit[i]))
        for h, v, u in rest:
            print("   %-86s %14.4f %s" % (h, v, u))
        print("   -- warp stall reasons, largest first")
        for h, v, u in sorted(stalls, key=lambda t: -t[1]):
            if v > 0:
                print("   %-86s %14.4f %s" % (h, v, u))


def sass(lib, kernel, lo=None, hi=None):
    """SASS listing of `kernel` (substring of the mangled name) from `lib`, optionally only instruction offsets [lo, hi] (hex), with a
    mnemonic histogram - the evidence behind "instructions per search step / per pixel" in DESIGN.md."""
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    cur = None; lines = []
    for ln in out.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            continue
        if cur and kernel in cur:
            m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", ln)
            if m:
                lines.append((int(m.group(1), 16), m.group(2).strip()))
    lo = int(lo, 16) if lo else 0
    hi = int(hi, 16) if hi else 1 << 30
    sel = [(o, t) for o, t in lines if lo <= o <= hi]
    hist = collections.Counter()
    for o, t in sel:
        op = t.split()[1] if t.startswith("@") else t.split()[0]
        hist[op.split(".")[0]] += 1
        print("/*%04x*/ %s" % (o, t))
    print("# %d instructions in [%#x, %#x] of %s" % (len(sel), lo, min(hi, lines[-1][0] if lines else 0), kernel))
    print("# " + ", ".join("%s %d" % kv for kv in hist.most_common()))


if __name__ == "__main__":
    fn = {"launches": launches, "full": full, "traffic": traffic, "issue": issue, "sass": sass}[sys.argv[1]]
    fn(*sys.argv[2:])
