"""Turns the raw ncu outputs brought back in gpurun_out/ into the small summaries kept under profiles/.

  python profiles/summarize.py launches <launch_list.csv> <bench.json> <out.md>  [title]
      ncu `--metrics gpu__time_duration.sum` launch list of a bench.py run  vs  the live CUDA-event profile of bench.py
      (shares of the step must agree; ncu's absolutes are cold-cache and serialised)
  python profiles/summarize.py full <capture.ncu-rep> <out.json> <samples> [note]
      one `ncu --set full` capture -> per-kernel DRAM bytes, throughputs, stalls (what bench.py's roofline.traffic reads)
"""
import csv
import io
import json
import re
import subprocess
import sys
from collections import defaultdict


def short(name):
    n = name.replace("void ", "")
    m = re.search(r"(?:^|::)(k_[A-Za-z0-9_]+)(<[^(]*>)?\(", n) or re.search(r"(?:^|::)(k_[A-Za-z0-9_]+)", n)
    if m and "at::" not in n:
        return m.group(1) + (m.group(2) or "" if m.lastindex and m.lastindex >= 2 else "")
    return "torch: " + n[:60]


def launches(csv_path, bench_json, out_md, title="training step"):
    rows = []
    with open(csv_path) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(io.StringIO("".join(lines))):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((int(r["ID"]), short(r["Kernel Name"]), float(r["Metric Value"]) / 1e3))
    bench = json.loads(open(bench_json).read().strip().splitlines()[-1])
    steps = bench["steps"]
    # the timed region and the per-kernel profile pass are the last 2*steps steps of the run: count launches of the
    # once-per-step first kernel backwards to find where they start
    marker = next((m for m in ("k_near_far", "k_step_begin", "k_grid_fwd") if any(n.startswith(m) for _, n, _ in rows)), None)  # first kernel of a step
    begins = [i for i, (_, n, _) in enumerate(rows) if n.startswith(marker)]
    first = begins[-2 * steps] if len(begins) >= 2 * steps else begins[0]
    nsteps = len([b for b in begins if b >= first])
    tot, cnt = defaultdict(float), defaultdict(int)
    for _, n, us in rows[first:]:
        key = n.split("<")[0] if n.startswith("k_") else n
        tot[key] += us
        cnt[key] += 1
    total = sum(tot.values())
    live = {k: v["us_per_launch"] * v["launches_per_step"] for k, v in bench.get("kernels", {}).items()}
    live_total = sum(live.values())
    out = [f"# {title} -- ncu launch list vs live CUDA-event profile", "",
           f"ncu: `{csv_path.split('/')[-1]}`, mean over the last {nsteps} steps of the run (cold-cache, serialised, eager launches).",
           f"live: `bench.py` per-launch CUDA events of the same command ({bench['config'].get('samples_per_step_per_gpu', '?')} samples/step). "
           "Shares must agree, not absolutes.", "",
           "| kernel | launches/step | ncu us/step | ncu share | live us/step | live share |", "|---|---|---|---|---|---|"]
    for k in sorted(tot, key=lambda k: -tot[k]):
        # bench.py names the fused instances of a kernel by role (k_ffmlp_bwd_mp_color ...); the profiler sees one template name
        match = [v for kk, v in live.items() if kk == k or kk.startswith(k + "_")]
        lv = sum(match) if match else None
        out.append(f"| `{k}` | {cnt[k] / nsteps:.2f} | {tot[k] / nsteps:.1f} | {100 * tot[k] / total:.1f} % | "
                   + (f"{lv:.1f} | {100 * lv / live_total:.1f} % |" if lv is not None else " | |"))
    out.append(f"| total | | {total / nsteps:.1f} | | {live_total:.1f} | |")
    open(out_md, "w").write("\n".join(out) + "\n")
    print(out_md)


WANT = {
    "us": ("gpu__time_duration.sum", 1.0),
    "dram_read_bytes": ("dram__bytes_read.sum", None),
    "dram_write_bytes": ("dram__bytes_write.sum", None),
    "sm_throughput_pct": ("sm__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    "lts_throughput_pct": ("lts__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    "l1tex_throughput_pct": ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    "l1tex2xbar_req_active_pct": ("l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed", 1.0),
    "dram_throughput_pct": ("dram__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    # tcgen05-era tensor counters (the hmma ones read n/a on UTCHMMA kernels): pipe busy cycles, executed tensor ops against the
    # pipe's peak (16384 ops/clk/SM = 8192 FMA: what a full-rate M=128 N=256 K=16 instruction sustains), TMEM port busy cycles
    "tensor_pipe_active_pct": ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 1.0),
    "tensor_ops_pct_of_peak": ("sm__ops_path_tensor_src_fp16_dst_fp32.avg.pct_of_peak_sustained_elapsed", 1.0),
    "tensor_ops": ("sm__ops_path_tensor_src_fp16_dst_fp32.sum", 1.0),
    "tmem_cycles_active_pct": ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 1.0),
    "sm_cycles_elapsed": ("sm__cycles_elapsed.avg", 1.0),
    "issue_active_pct": ("smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0),
    "warps_active_pct": ("sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
    "l1_hit_rate_pct": ("l1tex__t_sector_hit_rate.pct", 1.0),
    "l2_hit_rate_pct": ("lts__t_sector_hit_rate.pct", 1.0),
    "lts_sectors": ("lts__t_sectors.sum", 1.0),
    "lts_sectors_from_l1": ("lts__t_sectors_srcunit_tex.sum", 1.0),
    "warp_instructions": ("sm__inst_executed.sum", 1.0),
    "regs": ("launch__registers_per_thread", 1.0),
    "block": ("launch__block_size", 1.0),
    "grid": ("launch__grid_size", 1.0),
}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3}


def full(rep, out_json, samples, note=""):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    stall = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and "not_issued" not in h]
    kernels = defaultdict(list)
    for r in data:
        name = short(r[idx["Kernel Name"]]).split("<")[0]
        e = {}
        for k, (m, scale) in WANT.items():
            if m not in idx or r[idx[m]] in ("", "n/a"):
                continue
            try:
                v = float(r[idx[m]].replace(",", ""))
            except ValueError:
                continue
            if v != v:  # nan: a launch the profiler could not replay
                continue
            e[k] = v * UNIT.get(units[idx[m]], 1.0) if scale is None or k == "us" else v
        st = {h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""): float(r[idx[h]]) for h in stall}
        e["top_stalls_per_issue"] = {k: round(v, 2) for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:4] if k != "selected"}
        kernels[name].append(e)
    json.dump({"source": f"ncu --set full --clock-control none --import-source on; {note}", "samples": int(samples), "kernels": kernels},
              open(out_json, "w"), indent=1)
    print(out_json, {k: len(v) for k, v in kernels.items()})


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(*sys.argv[2:6])
    else:
        full(*sys.argv[2:6])
