"""Condenses a gpurun visit (tools/gpu_round.sh <tag>) into the tracked files under profiles/:
  profiles/<tag>_launches.md     per-kernel share of the step from the ncu launch list (cold-cache, serialised)
  profiles/<tag>_full_<k>.md     key counters of each `ncu --set full` capture
usage: python tools/ncu_summary.py <tag>
"""
import collections
import csv
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"]


TRAFFIC = {}


def launches(tag):
    path = os.path.join(OUT, f"{tag}_launches.csv")
    if not os.path.exists(path):
        return
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr = rows[hi]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[hi + 1:]:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] in ("ns", "nsecond") else (v * 1e3 if r[ui] in ("ms", "msecond") else v)
        name = r[ki].split("(")[0].replace("void ", "").strip()
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    with open(os.path.join(PROF, f"{tag}_launches.md"), "w") as f:
        f.write(f"# {tag}: ncu launch list of one transcribe pass (whisper-tiny, 64 clips, bf16)\n\n"
                "`ncu --metrics gpu__time_duration.sum --clock-control none` over `python tools/one_step.py 2`, second pass only.\n"
                "Per-launch times are cold-cache and serialised: read the SHARE column, not the absolutes.\n\n"
                f"launches: {sum(v[0] for v in agg.values())}, summed device time {tot / 1e3:.2f} ms\n\n"
                "| kernel | launches | total us | avg us | share |\n|---|---:|---:|---:|---:|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k}` | {v[0]} | {v[1]:.1f} | {v[1] / v[0]:.2f} | {100 * v[1] / tot:.1f}% |\n")
    print("wrote", f"profiles/{tag}_launches.md")


def full(tag):
    for rep in sorted(glob.glob(os.path.join(OUT, f"{tag}_full_*.ncu-rep"))):
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(raw.splitlines()))
        if len(rows) < 3:
            continue
        hdr, units = rows[0], rows[1]
        name = os.path.basename(rep)[:-8]
        with open(os.path.join(PROF, name + ".md"), "w") as f:
            f.write(f"# {name}: `ncu --set full --clock-control none --import-source on` (tools/one_step.py 2)\n\n")
            for r in rows[2:]:
                f.write(f"## `{r[hdr.index('Kernel Name')][:110]}` id {r[0]}\n\n| metric | value |\n|---|---|\n")
                for w in WANT:
                    if w in hdr:
                        i = hdr.index(w)
                        f.write(f"| {w} | {r[i]} {units[i]} |\n")
                f.write("\n")
        print("wrote", f"profiles/{name}.md")
        # DRAM traffic per launch of each captured kernel (read + write), for bench.py's roofline.traffic
        unit_b = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        for r in rows[2:]:
            try:
                ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
                byt = float(r[ir]) * unit_b[units[ir]] + float(r[iw]) * unit_b[units[iw]]
            except (ValueError, KeyError):
                continue
            kname = r[hdr.index("Kernel Name")].split("(")[0].split("<")[0].replace("void ", "").replace("wt::", "").strip()
            TRAFFIC.setdefault(kname, []).append(byt)


if __name__ == "__main__":
    t = sys.argv[1]
    os.makedirs(PROF, exist_ok=True)
    launches(t)
    full(t)
    if TRAFFIC:
        import json
        with open(os.path.join(PROF, "traffic.json"), "w") as f:
            json.dump({"tag": t, "how": "ncu --set full: dram__bytes_read.sum + dram__bytes_write.sum per captured launch "
                       "(tools/one_step.py: whisper-tiny, 64 clips)", "bytes_per_launch": TRAFFIC}, f, indent=1)
        print("wrote profiles/traffic.json")
