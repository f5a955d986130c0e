"""Turns gpurun_out/launches.csv + prof_*.ncu-rep into the text summaries committed under profiles/."""
import collections, csv, subprocess, sys, os

tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
out_dir = "profiles"
os.makedirs(out_dir, exist_ok=True)

def launches(path, out):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
        k = row["Kernel Name"]
        k = k.replace("void ", "").replace("b200rl::<unnamed>::", "b200rl::")[:90]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    ours = sum(v[1] for k, v in agg.items() if "b200rl::" in k)
    with open(out, "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none, `python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu`\n")
        f.write("# (includes the bench's one-off input generation kernels: at::distribution/copy/Fill; per-launch times\n")
        f.write("#  are cold-cache and serialised, compare SHARES not absolutes)\n")
        f.write(f"# total {tot:.1f} us over {sum(v[0] for v in agg.values())} launches; b200rl kernels {ours:.1f} us\n")
        f.write(f"{'us':>12s} {'share':>7s} {'share_of_b200rl':>16s} {'n':>5s}  kernel\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:25]:
            so = f"{100*v[1]/ours:6.2f}%" if "b200rl::" in k else "      -"
            f.write(f"{v[1]:12.1f} {100*v[1]/tot:6.2f}% {so:>16s} {v[0]:5d}  {k}\n")
    print("wrote", out)

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
        "sm__cycles_active.avg", "lts__t_sector_hit_rate.pct",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]

def full(rep, out, note):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on; {note}\n")
        for row in rows[2:]:
            name = row[hdr.index("Kernel Name")]
            f.write(f"\n== {name[:110]}\n")
            for w in WANT:
                if w in hdr:
                    i = hdr.index(w)
                    f.write(f"{w:88s} {row[i]:>16s} {units[i]}\n")
            try:
                t = float(row[hdr.index('gpu__time_duration.sum')].replace(',', ''))
                tu = units[hdr.index('gpu__time_duration.sum')]
                t_s = t * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1}.get(tu, 1e-6)
                rd = float(row[hdr.index('dram__bytes_read.sum')].replace(',', ''))
                wr = float(row[hdr.index('dram__bytes_write.sum')].replace(',', ''))
                mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
                rd *= mul[units[hdr.index('dram__bytes_read.sum')]]
                wr *= mul[units[hdr.index('dram__bytes_write.sum')]]
                f.write(f"{'derived: DRAM traffic (read+write) per launch':88s} {(rd+wr)/1e6:16.1f} MB\n")
                f.write(f"{'derived: DRAM GB/s under ncu (cold, serialised)':88s} {(rd+wr)/t_s/1e9:16.0f} GB/s\n")
            except Exception as e:
                f.write(f"derived: n/a ({e})\n")
    print("wrote", out)

if os.path.exists("gpurun_out/launches.csv"):
    launches("gpurun_out/launches.csv", f"{out_dir}/{tag}_launches_bench.txt")
for name, note in (("prof_h1", "`python scripts/prof_h1.py 2048 h1` (T=2048 rows x V=151936 bf16 = 622 MB per pass)"),
                   ("prof_fused", "`python scripts/prof_h1.py 2048 fused` (fused GRPO step, T=2048 rows x V=151936 bf16; algorithmic HBM bytes 4*V/token = 1244.7 MB)"),
                   ("prof_interval", "`python scripts/prof_interval.py` (cfg5 lists at 1/2 scale: 3.8e9-element bf16 buffer, 5e4 intervals, 3.8 GB copied per pass; slice, set, fused move)"),
                   ("prof_gae", "`python scripts/prof_h1.py 0 gae` (65536 sequences of 64-2048 tokens, 69 M tokens)"),
                   ("prof_gae_team", "`python scripts/prof_h1.py 0 gae_team` (120 sequences of 1024-16384 tokens, gamma = lambda = 1: "
                                     "the longest sequence's chain is the run time; static schedule of the chain = 4.0 cycles per token)"),):
    rep = f"gpurun_out/{name}.ncu-rep"
    if os.path.exists(rep):
        full(rep, f"{out_dir}/{tag}_ncu_{name[5:]}.txt", note)


def traffic_json(out="profiles/traffic.json", rows=2048):
    """dram__bytes_read.sum + dram__bytes_write.sum per [T, V] row of the H1 / fused kernels, from the
    `ncu --set full` captures of this run (T = 2048 rows); bench.py multiplies by the rows of a
    launch for `roofline.traffic`."""
    import json

    mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    res = {}
    for rep, pats in (("gpurun_out/prof_h1.ncu-rep", (("logprob_bwd_kernel<0, 0>", "logprob_bwd_bf16_kernel_bytes_per_row"),
                                                      ("logprob_fwd_kernel<0, 24, 2, 1, 1>", "logprob_fwd_bf16_kernel_bytes_per_row"))),
                      ("gpurun_out/prof_fused.ncu-rep", (("grpo_fused_kernel", "grpo_fused_kernel_bytes_per_row"),))):
        if not os.path.exists(rep):
            continue
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rws = list(csv.reader(raw.splitlines()))
        hdr, units = rws[0], rws[1]
        ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
        for pat, key in pats:
            vals = [float(r[ir].replace(",", "")) * mul[units[ir]] + float(r[iw].replace(",", "")) * mul[units[iw]]
                    for r in rws[2:] if pat in r[ik]]
            if vals:
                res[key] = sum(vals) / len(vals) / rows
    if res:
        res["source"] = (f"scripts/summarize_prof.py {tag}: ncu --set full captures of scripts/prof_h1.py (T={rows} rows x V=151936 bf16), "
                         "dram__bytes_read.sum + dram__bytes_write.sum averaged over the captured launches")
        json.dump(res, open(out, "w"), indent=1)
        print("wrote", out, res)


traffic_json()
