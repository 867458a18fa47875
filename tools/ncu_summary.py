"""Condense an `ncu --set full` report into the per-launch summary CSV kept under profiles/ and (optionally) refresh
profiles/ncu_traffic.json, the per-stage DRAM traffic bench.py quotes in `roofline.traffic`.
   python tools/ncu_summary.py <rep> <out.csv> [--traffic C3] [--render C3]
--render also refreshes profiles/ncu_render.json: issue-active, FMA / ALU / LSU / XU pipe utilisation, shared-memory wavefronts
per clock and SM, registers and warp occupancy of the two tile renderers (what bench.py quotes for the issue-bound stages)."""
import csv, io, json, os, subprocess, sys

COLS = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__grid_size",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
STAGE_OF = [("render_bwd", "render_bwd"), ("render_fwd", "render_fwd"), ("backward_views", "backward_views"), ("preprocess_bwd", "preprocess_bwd"),
            ("preprocess_views", "preprocess_views"), ("preprocess_fwd", "preprocess_fwd")]

rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
ki = hdr.index("Kernel Name")
idx = [hdr.index(c) for c in COLS if c in hdr]
with open(out, "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["id", "Kernel Name"] + [hdr[i] for i in idx])
    w.writerow(["", ""] + [units[i] for i in idx])
    for n, r in enumerate(data):
        w.writerow([n, r[ki].replace(",", ";")[:48]] + [r[i] for i in idx])
print("wrote", out, len(data), "launches")

if "--traffic" in sys.argv:
    wl = sys.argv[sys.argv.index("--traffic") + 1]
    tp = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json")
    tj = json.load(open(tp))
    ri, wi = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per = {}
    for r in data:
        for key, stage in STAGE_OF:
            if key in r[ki]:
                per.setdefault(stage, []).append(float(r[ri]) * scale[units[ri]] + float(r[wi]) * scale[units[wi]])
                break
    tj.setdefault(wl, {}).update({k: int(sum(v) / len(v)) for k, v in per.items()})
    tj["_comment"] = "dram__bytes_read.sum + dram__bytes_write.sum per launch from ncu --set full captures (bytes); key = workload -> stage; source: " + os.path.basename(rep)
    json.dump(tj, open(tp, "w"), indent=1)
    print("traffic", tj[wl])

if "--render" in sys.argv:
    wl = sys.argv[sys.argv.index("--render") + 1]
    rp = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_render.json")
    rj = json.load(open(rp)) if os.path.exists(rp) else {}
    col = lambda name: hdr.index(name) if name in hdr else None
    def mean(rows_, name):
        i = col(name)
        vals = [float(r[i].replace(",", "")) for r in rows_ if i is not None and r[i] not in ("", "n/a")]
        return sum(vals) / len(vals) if vals else None
    out_r = {}
    for key in ("render_fwd", "render_bwd"):
        rows_ = [r for r in data if key in r[ki]]
        if not rows_:
            continue
        cyc = mean(rows_, "sm__cycles_elapsed.max")
        wave = mean(rows_, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")
        sms = 148
        pct = lambda n: (round(mean(rows_, n) / 100.0, 4) if mean(rows_, n) is not None else None)
        out_r[key] = {"issue_active": pct("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                      "pipe_fma": pct("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                      "pipe_alu": pct("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                      "pipe_lsu": pct("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
                      "pipe_xu": pct("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
                      "smem_wavefronts_per_clk_per_sm": round(wave / (cyc * sms), 4) if (wave and cyc) else None,
                      "registers": mean(rows_, "launch__registers_per_thread"),
                      "warp_occupancy": pct("sm__warps_active.avg.pct_of_peak_sustained_active"),
                      "warp_instructions": mean(rows_, "smsp__inst_executed.sum"),
                      "launches_averaged": len(rows_)}
    out_r["_source"] = os.path.basename(rep) + " (ncu --set full, --clock-control none)"
    rj[wl] = out_r
    json.dump(rj, open(rp, "w"), indent=1)
    print("render", json.dumps(out_r))
