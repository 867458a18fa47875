"""Per-CUDA-source-line instruction counts and stall samples for one launch of an ncu report (needs -lineinfo + --import-source on):
   python tools/ncu_lines.py <rep> <launch-skip> [top]"""
import csv, subprocess, sys, io, os
rep, skip = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda", "--launch-skip", skip, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, cur, data = None, "", []
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": cur = os.path.basename(r[1]); continue
    if len(r) >= 2 and r[0] == "Function Name": print(r[1][:100]); continue
    if "# Samples" in r: hdr = r; continue
    if hdr and len(r) == len(hdr) and r[2] == "-" and r[hdr.index("# Samples")].strip().isdigit():
        data.append((cur, r))
ni, ii = hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot_s = sum(int(r[ni]) for _, r in data); tot_i = sum(int(r[ii] or 0) for _, r in data)
print("total samples", tot_s, "total warp instructions", tot_i)
for f, r in sorted(data, key=lambda fr: -int(fr[1][ii] or 0))[:top]:
    st = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall_cols), reverse=True)[:2]
    print(f"{f[:18]:18s}:{r[0]:>4s} inst {100*int(r[ii] or 0)/max(tot_i,1):5.1f}%  samp {100*int(r[ni])/max(tot_s,1):5.1f}% {st[0][1][:10]:10s} | {r[1].strip()[:100]}")
