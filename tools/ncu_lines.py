"""Per-source-line instruction counts / stall samples from an ncu report.
usage: python tools/ncu_lines.py report.ncu-rep kernel_name [launch-id]"""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", kern]
if len(sys.argv) > 3:
    cmd += ["--launch-skip", sys.argv[3], "--launch-count", "1"]
out = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur_file, hdr = None, None
agg = {}
tot_inst = tot_samp = 0
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        ii, si, ti = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Predicated-On Thread Instructions Executed")
        continue
    if hdr is None or len(r) < len(hdr) or r[2] != "-":
        continue                      # only the per-source-line summary rows (Address == '-')
    try:
        inst, samp, thr = int(r[ii]), int(r[si]), int(r[ti])
    except ValueError:
        continue
    k = (cur_file, int(r[0]), r[1].strip()[:100])
    a = agg.setdefault(k, [0, 0, 0])
    a[0] += inst
    a[1] += samp
    a[2] += thr
    tot_inst += inst
    tot_samp += samp
print(f"total inst {tot_inst}  samples {tot_samp}")
for k, a in sorted(agg.items(), key=lambda x: -x[1][0])[:45]:
    print(f"{a[0]:12d} {100*a[0]/max(tot_inst,1):5.1f}%  samp {100*a[1]/max(tot_samp,1):5.1f}%  lanes {a[2]/max(a[0],1):4.1f}  {k[0]}:{k[1]:<4d} {k[2]}")
