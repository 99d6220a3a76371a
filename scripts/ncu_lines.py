#!/usr/bin/env python
"""Per-source-line cost of one profiled kernel: joins `ncu --page source` (SASS rows: samples,
instructions executed) with `nvdisasm -g` line info of the same cubin by instruction order.
usage: ncu_lines.py <report.ncu-rep> <kernel regex> <library.so> [top]"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile


def sass_lines(lib, kernel):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=d, capture_output=True)
    cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
    out, cur, infn = [], None, False
    for ln in txt.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            infn = re.search(kernel[:-1] + "E" if kernel.endswith("$") else kernel, m.group(1)) is not None  # mangled: name + 'E'
            continue
        if not infn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4}\*/", ln):
            out.append(cur)
    return out


def main():
    rep, kernel, lib = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
    csvtxt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kernel], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(csvtxt)))
    h = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
    hdr = rows[h]
    si, ii = hdr.index("# Samples"), hdr.index("Instructions Executed")
    tab = []
    for r in rows[h + 1:]:
        if r and r[0] == "Kernel Name":
            break
        if len(r) > ii and r[ii].isdigit():
            tab.append((int(r[si]), int(r[ii]), r[1]))
    lines = sass_lines(lib, kernel)
    print("sass rows %d, lineinfo rows %d" % (len(tab), len(lines)))
    agg = collections.defaultdict(lambda: [0, 0])
    for k, (s, n, _) in enumerate(tab):
        key = lines[k] if k < len(lines) and lines[k] else ("?", 0)
        agg[key][0] += s; agg[key][1] += n
    ts = sum(v[0] for v in agg.values()) or 1
    ti = sum(v[1] for v in agg.values()) or 1
    src = {}
    print("total samples %d, warp-instructions %d" % (ts, ti))
    for (f, l), (s, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        if f not in src:
            try:
                src[f] = open(os.path.join(os.path.dirname(os.path.abspath(lib)), "csrc", f)).read().splitlines()
            except Exception:
                src[f] = []
        text = src[f][l - 1].strip()[:90] if 0 < l <= len(src[f]) else ""
        print("%5.1f%% smp %5.1f%% inst  %s:%d  %s" % (100 * s / ts, 100 * n / ti, f, l, text))


if __name__ == "__main__":
    main()
