#!/bin/bash
# Summaries of one capture set for profiles/ (run here, no GPU): scripts/make_profiles.sh <prefix>     e.g. r3c
# expects gpurun_out/<prefix>_{mapu,track,mapu_early}.ncu-rep, <prefix>_launches.csv, <prefix>_pytest.log (profiles/README.md)
set -e
P=$1
cd "$(dirname "$0")/.."
H=$(python -c "import sys; sys.path.insert(0,'.'); import bench; print(bench.source_hash())")
python scripts/ncu_summary.py launches gpurun_out/${P}_launches.csv > profiles/${P}_launches.txt
(for r in mapu mapu_early track; do echo "### gpurun_out/${P}_$r.ncu-rep"; python scripts/ncu_summary.py full gpurun_out/${P}_$r.ncu-rep; done) > profiles/${P}_full.txt 2>&1
H=$H P=$P python - <<'PY'
import json, os, subprocess
H, P = os.environ["H"], os.environ["P"]
ks = {}
for r in ("mapu", "track"):
    ks.update(json.loads(subprocess.check_output(["python", "scripts/ncu_summary.py", "traffic", "gpurun_out/%s_%s.ncu-rep" % (P, r)]))["kernels"])
out = {"source": "%s_mapu.ncu-rep + %s_track.ncu-rep (gpurun, B200)" % (P, P), "source_hash": H,
       "how": "ncu --set full --clock-control none --import-source on -k regex:<kernel> -s 450 -c 1 python bench.py --steps 12 --warmup 3 "
              "--no-cpu-baseline --no-other-configs --no-strong: the launch of frame 451 of 256 sequences (steady state); scripts/ncu_summary.py traffic",
       "kernels": ks}
json.dump(out, open("profiles/traffic_%s.json" % H, "w"), indent=1)
print(H, {k: (round(v["duration_us"], 1), int(v["dram_bytes"])) for k, v in ks.items()})
PY
(for r in mapu track; do python scripts/ncu_summary.py issue gpurun_out/${P}_$r.ncu-rep; done) \
  | grep -v "l1tex__data_pipe\|l1tex__t_\|\.min\.\|\.max\.\|realtime\|derived__\|_not_issued\|pcsamp" > profiles/issue_$H.txt
grep -E "passed|failed|^FAILED|^ERROR|^FREE-RUN|^TEACHER-FORCED|^VO-VS-GT|^SCHEDULES|^RNG-RUN|differed by more|not bit-identical" gpurun_out/${P}_pytest.log | sort -u > profiles/${P}_pytest_gpu.txt
echo "profiles/${P}_*, traffic_$H.json, issue_$H.txt written"
