#!/usr/bin/env python
"""Turn a capture directory written by profiles/probes/r02_capture.sh (gpurun_out/r02) into the committed summaries:
  profiles/r02_launches.md  per-kernel launch statistics of the bench command and of one batched step of every 3D /
                            detectron workload (ncu --metrics gpu__time_duration.sum, cold-cache, serialised)
  profiles/r02_kernels.md   one row per `ncu --set full` capture: duration, DRAM bytes against the algorithmic bytes,
                            issue / occupancy figures, the three largest stall reasons
  profiles/r02_pre_kernel.json  DRAM traffic of the roofline kernel (read by bench.py for `roofline.traffic`)
Usage: python profiles/probes/r02_summary.py [capture_dir]"""
import csv
import json
import os
import re
import subprocess
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CAP = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "r02")
OURS = re.compile(r"(preprocess|yolo|det_|pp_|rot_|cp_|vox_|pillar_|dyn_|jpeg_|draw_|scan_|pc2_|aa_|eval_|box_iou|class_filter|rotated_|scale_boxes|jsk_)")


def short(name):
    name = name.replace("<unnamed>::", "").replace("void ", "")
    return re.sub(r"\(.*", "", name)


def launch_rows(path):
    rows = list(csv.reader(open(path)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    cols = rows[hdr]
    ki, vi, gi = cols.index("Kernel Name"), cols.index("Metric Value"), cols.index("Grid Size")
    return [(short(r[ki]), r[gi], float(r[vi].replace(",", "")) / 1e3) for r in rows[hdr + 1:] if len(r) > vi]


def table(rows, last_step_only=False):
    """rows: (kernel, grid, us). Our kernels only; share over our kernels."""
    mine = [r for r in rows if OURS.search(r[0])]
    agg = OrderedDict()
    for k, g, us in mine:
        a = agg.setdefault(k, [0, 0.0, g])
        a[0] += 1
        a[1] += us
    total = sum(a[1] for a in agg.values())
    out = ["| kernel | grid | launches | mean µs | share of the listed time |", "|---|---|---:|---:|---:|"]
    for k, (n, t, g) in agg.items():
        out.append(f"| `{k}` | {g} | {n} | {t / n:.1f} | {100 * t / total:.1f} % |")
    return "\n".join(out), total, agg


def raw_metrics(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    return {h: (v, u) for h, u, v in zip(hdr, units, vals)}


def num(m, key):
    v, u = m.get(key, ("0", ""))
    x = float(v.replace(",", "")) if v not in ("", "n/a") else 0.0
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}.get(u, 1.0)
    return x * scale


# algorithmic bytes of the captured launches (what DESIGN.md section 5 states per unit x the units of the launch)
ALGO = {
    "pre": ("`preprocess2_kernel<1,1,0>` (bench, 64 frames 1280x720 -> 640x640 letterbox)", 64 * (3 * 720 * 1280 + 12 * 640 * 640)),
    "filter": ("`yolo_filter_stream_kernel<128>` (bench, 64 x 25200 x 7)", 64 * 25200 * 7 * 4),
    "nms": ("`yolo_nms_kernel` (bench, 128 CTAs)", None),
    "detcand": ("`det_candidates_kernel` (16 RetinaNet images, 201 600 anchors x 80 classes)", 16 * 201600 * 80 * 4 // 1 * 1 // 1),
    "detlevel": ("`det_level_kernel` (80 lists, top-1000 each)", None),
    "detpost": ("`det_post_kernel` (16 images: class-aware NMS of <= 5000 rows in rounds, 100 kept)", None),
    "ppcand": ("`pp_candidates_kernel` (16 frames x 321 408 anchors x 3 classes)", 16 * 321408 * 3 * 4),
    "decorate": ("`pillar_decorate_kernel<1,32,4>` (16 x 12 000 pillars, 32 x 4 -> 32 x 10)", 16 * 12000 * 32 * 14 * 4),
    "voxcell": ("`vox_cell_kernel<1>` (16 sweeps, 1.73 M points)", None),
    "voxplace": ("`vox_place_kernel<1>`", None),
    "voxscatter": ("`vox_scatter_kernel` (also zero-fills 16 x 12 000 x 32 x 4 floats)", 16 * 12000 * 32 * 4 * 4),
    "cphead": ("`cp_head_kernel` (16 frames x 6 heads)", None),
    "rotmask": ("`rot_mask_kernel` (16 frames x 4096 candidates)", None),
    "rotpairs": ("`rot_pairs_kernel`", None),
    "dethist": ("`det_hist_kernel`", None),
    "flat": ("`preprocess_flat_kernel<1,0>` (16 frames 800x1333, same size)", 16 * 15 * 800 * 1333),
    "huff": ("`jpeg_huffman_kernel` (64 frames 1280x720 q90 4:2:0)", None),
}
STALLS = ["long_scoreboard", "short_scoreboard", "barrier", "lg_throttle", "mio_throttle", "membar", "math_pipe_throttle", "wait",
          "no_instruction", "drain", "not_selected", "dispatch_stall", "branch_resolving", "sleeping", "tex_throttle", "imc_miss"]


def main():
    md = ["# Round 2 — ncu launch lists (B200, `--clock-control none`, `--metrics gpu__time_duration.sum`)", "",
          "Written by `profiles/probes/r02_summary.py` from the capture of `profiles/probes/r02_capture.sh` (each command had exited 0",
          "without ncu first). Launches under the profiler are cold-cache and serialised, with no programmatic overlap between them: a",
          "kernel's SHARE of the listed time is what compares with the CUDA-event split of the bench line, not the absolute value.", ""]
    bench_line = None
    bj = os.path.join(CAP, "bench.json")
    if os.path.exists(bj):
        bench_line = json.loads(open(bj).read().strip().splitlines()[-1])
    for tag, title in (("bench", "`python bench.py --steps 4 --warmup 3 --min-seconds 0 --no-configs --no-cpu-baseline` (configs[1], the line's workload)"),
                       ("detectron", "`python profiles/probes/prof_posts.py detectron` (bench.py's configs[2] workload, 3 steps of 16 images)"),
                       ("pointpillars", "`python profiles/probes/prof_posts.py pointpillars` (configs[3], 3 steps of 16 sweeps)"),
                       ("centerpoint", "`python profiles/probes/prof_posts.py centerpoint` (configs[4], 3 steps of 16 clouds)"),
                       ("jpeg", "`python profiles/probes/prof_jpeg.py 2` (64 frames 1280x720, quality 90, no restart intervals)")):
        path = os.path.join(CAP, f"launches_{tag}.csv")
        if not os.path.exists(path):
            continue
        rows = launch_rows(path)
        if tag == "bench":      # the device-resident leg only: the launches up to the first torch kernel after the timed steps
            ours = [i for i, r in enumerate(rows) if OURS.search(r[0])]
            cut = ours[0]
            for a, b in zip(ours, ours[1:]):
                if b - a > 1:
                    break
                cut = b
            rows = rows[:cut + 1]
        t, total, agg = table(rows)
        md += [f"## {title}", "", t, ""]
        if tag == "bench" and bench_line:
            km = bench_line.get("kernel_ms") or {}
            md += [f"Bench line of the same build (CUDA events, warm, launches overlapping): {bench_line['ms_per_step']:.4f} ms per step, "
                   f"stages {json.dumps(km.get('stages', km))}.", ""]
    open(os.path.join(ROOT, "profiles", "r02_launches.md"), "w").write("\n".join(md) + "\n")

    peak = 6556.5
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    km = ["# Round 2 — `ncu --set full` captures, one launch each (B200, `--clock-control none --import-source on`)", "",
          "Written by `profiles/probes/r02_summary.py`. `alg.` = algorithmic bytes of the launch (DESIGN.md section 5); DRAM = `dram__bytes_read.sum` +",
          "`dram__bytes_write.sum` of the same launch; the fraction is algorithmic bytes / duration against the measured copy bandwidth",
          f"({peak} GB/s, MEASURED_PEAKS.json) — a cold, serialised launch, so the bench's CUDA-event figure is the one quoted elsewhere.", "",
          "| kernel | µs | alg. MB | DRAM read / write MB | DRAM / alg. | alg. GB/s (frac) | warp inst M | issue active % | warps active % | regs | top stalls (per issue) |",
          "|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|---|"]
    for fn in sorted(os.listdir(CAP)):
        if not (fn.startswith("full_") and fn.endswith(".ncu-rep")):
            continue
        tag = fn[5:-8]
        m = raw_metrics(os.path.join(CAP, fn))
        if not m:
            continue
        us = num(m, "gpu__time_duration.sum")
        rd, wr = num(m, "dram__bytes_read.sum"), num(m, "dram__bytes_write.sum")
        name, alg = ALGO.get(tag, (f"`{short(m.get('Kernel Name', (tag, ''))[0])}`", None))
        stalls = sorted(((num(m, f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio"), s) for s in STALLS), reverse=True)[:3]
        st = ", ".join(f"{s.replace('_', ' ')} {v:.1f}" for v, s in stalls)
        algs = f"{alg / 1e6:.1f}" if alg else "—"
        ratio = f"{(rd + wr) / alg:.2f}" if alg else "—"
        bw = f"{alg / us / 1e3:.0f} ({alg / us / 1e3 / peak:.2f})" if alg else "—"
        km.append(f"| {name} | {us:.1f} | {algs} | {rd / 1e6:.1f} / {wr / 1e6:.1f} | {ratio} | {bw} | {num(m, 'smsp__inst_executed.sum') / 1e6:.2f} | "
                  f"{num(m, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):.0f} | {num(m, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.0f} | "
                  f"{num(m, 'launch__registers_per_thread'):.0f} | {st} |")
        if tag == "pre":
            json.dump({"kernel": "preprocess2_kernel<1,1,0>", "command": "ncu --set full --clock-control none --import-source on -k regex:preprocess -s 6 -c 1 python bench.py "
                       "--steps 4 --warmup 3 --min-seconds 0 --no-configs --no-cpu-baseline", "duration_us": us, "dram_bytes_read": rd, "dram_bytes_write": wr,
                       "dram_traffic_bytes_per_launch": int(rd + wr), "algorithmic_bytes_per_launch": alg, "round": 2},
                      open(os.path.join(ROOT, "profiles", "r02_pre_kernel.json"), "w"), indent=1)
    open(os.path.join(ROOT, "profiles", "r02_kernels.md"), "w").write("\n".join(km) + "\n")
    print("wrote profiles/r02_launches.md, profiles/r02_kernels.md")


if __name__ == "__main__":
    main()
