"""Summarise the per-instruction stall samples of one launch of an ncu report:
   python tools/ncu_stalls.py <rep> <launch-skip> [top]   -> top SASS instructions by stall samples + totals per stall reason"""
import csv, subprocess, sys, io
rep, skip = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--launch-skip", skip, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][1][:120])
hdr = rows[1]
si, ai, ni = hdr.index("Source"), hdr.index("Address"), hdr.index("# Samples")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = [r for r in rows[2:] if len(r) == len(hdr) and r[ni].strip().isdigit()]
tot = sum(int(r[ni] or 0) for r in data)
agg = {hdr[i]: sum(int(r[i] or 0) for r in data) for i in stall_cols}
print("total samples", tot, {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
for r in sorted(data, key=lambda r: -int(r[ni] or 0))[:top]:
    st = sorted(((int(r[i] or 0), hdr[i]) for i in stall_cols), reverse=True)[:2]
    print(f"{int(r[ni]):7d} {100*int(r[ni])/max(tot,1):5.1f}%  {r[si][:90]:90s} {st}")
