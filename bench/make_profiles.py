#!/usr/bin/env python
"""Turns the raw measurement files that gpurun brought back (gpurun_out/, scratch) into the tracked
summaries under profiles/. Re-run after every measurement campaign."""
import csv
import json
import os
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")


def load(name):
    path = os.path.join(G, name)
    if not os.path.exists(path):
        return []
    rows = [json.loads(l) for l in open(path) if l.strip().startswith("{")]
    # a timing without a clock record is not admissible (B200_PROFILING.md): such rows are dropped, loudly
    bad = [r for r in rows if "us" in r and not (r.get("clocks") or {}).get("sm_mhz")]
    if bad:
        print(f"make_profiles: {name}: REFUSING {len(bad)} of {len(rows)} rows that carry no clock record")
    hot = [r for r in rows if set((r.get("clocks") or {}).get("reasons", [])) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}]
    if hot:
        print(f"make_profiles: {name}: REFUSING {len(hot)} rows measured under a thermal / hardware slowdown")
    return [r for r in rows if r not in bad and r not in hot]


def sweep_table(n, fname, note=""):
    rows = load(fname)
    if not rows:
        return
    shutil.copy(os.path.join(G, fname), os.path.join(P, "raw", fname))
    sizes = sorted({r["bytes"] for r in rows if "bytes" in r})
    out = [f"# Allreduce sweep, {str(n).split('_')[0]} x B200, {'bf16' if 'bf16' in str(n) else 'fp32'} SUM, in place on symmetric buffers", "",
           "Device time (CUDA events around back-to-back calls, MAX over ranks / reps); `busbw = bytes/t * 2(N-1)/N`.",
           f"Source: `bench/sweep.py` under torchrun, raw lines in `profiles/raw/{fname}`. {note}", "",
           "| bytes | NCCL us | NCCL busbw | flextree auto us | algo | busbw GB/s | model est us | speed-up vs NCCL |",
           "|---:|---:|---:|---:|---|---:|---:|---:|"]
    for sz in sizes:
        nccl = [r for r in rows if r["impl"] == "nccl" and r["bytes"] == sz]
        auto = [r for r in rows if r.get("requested") == "auto" and r["bytes"] == sz and r.get("mode", "registered") == "registered"]
        if not nccl or not auto:
            continue
        a, c = auto[0], nccl[0]
        algo = f"{a['algo']} {'x'.join(map(str, a['widths']))} (c{a['ctas']}, {a['tile_bytes'] >> 10}K)"
        out.append(f"| {sz} | {c['us']:.1f} | {c['busbw_GBps']:.1f} | {a['us']:.1f} | {algo} | {a['busbw_GBps']:.1f} | "
                   f"{a.get('est_us', 0):.1f} | {c['us'] / a['us']:.2f}x |")
    out += ["", "## Every algorithm / width vector (us measured / us predicted by the link model)", "",
            "| bytes | " + " | ".join(sorted({_label(r) for r in rows if r["impl"] == "flextree" and "error" not in r})) + " |"]
    labels = sorted({_label(r) for r in rows if r["impl"] == "flextree" and "error" not in r})
    out.append("|---:|" + "---:|" * len(labels))
    for sz in sizes:
        cells = []
        for lb in labels:
            m = [r for r in rows if r["impl"] == "flextree" and "error" not in r and r["bytes"] == sz and _label(r) == lb]
            cells.append(f"{m[0]['us']:.1f} / {m[0].get('est_us', 0):.0f}" if m else "-")
        out.append(f"| {sz} | " + " | ".join(cells) + " |")
    open(os.path.join(P, f"sweep_n{n}.md"), "w").write("\n".join(out) + "\n")


def _label(r):
    w = "x".join(map(str, r.get("widths", [])))
    return f"{r.get('requested', r.get('algo'))}:{r.get('algo')}:{w}:{r.get('mode', '')[:3]}"


def tune_table(n, fname, title):
    rows = load(fname)
    if not rows:
        return
    shutil.copy(os.path.join(G, fname), os.path.join(P, "raw", fname))
    sizes = sorted({r["bytes"] for r in rows if "bytes" in r})
    out = [f"# Launch-shape scan, {n} x B200 ({title})", "",
           "us per allreduce for (CTAs per rank, tile bytes per block-chunk); fp32 SUM in place on symmetric buffers.", ""]
    for sz in sizes:
        nccl = [r for r in rows if r["impl"] == "nccl" and r["bytes"] == sz]
        out.append(f"## {sz} bytes   (NCCL: {nccl[0]['us']:.1f} us, {nccl[0]['busbw_GBps']:.0f} GB/s)" if nccl else f"## {sz} bytes")
        for algo in ("stages", "nvls"):
            sel = sorted([r for r in rows if r.get("algo") == algo and r["bytes"] == sz and "error" not in r], key=lambda r: r["us"])
            if sel:
                out.append(f"* **{algo}** best->worst: " + ", ".join(
                    f"(c{r['ctas']}, {r['tile_bytes'] >> 10}K{', t' + str(r['threads']) if 'threads' in r else ''}) {r['us']:.1f}" for r in sel[:10]) +
                    f" ... worst {sel[-1]['us']:.1f}")
        out.append("")
    open(os.path.join(P, f"{os.path.splitext(fname)[0]}.md"), "w").write("\n".join(out) + "\n")


def ncu_summary(rep, out_name, title):
    path = os.path.join(G, rep)
    if not os.path.exists(path):
        return
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    if len(rows) < 3:
        return
    hdr, units = rows[0], rows[1]
    want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
            "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
            "smsp__inst_executed.sum", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
            "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio" if False else "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
            "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__warp_issue_stalled_membar_per_warp_active.pct"]
    out = [f"# ncu --set full: {title}", "", f"Report: `gpurun_out/{rep}` (scratch); command in profiles/README.md. One block per captured launch.", ""]
    for r in rows[2:]:
        out.append("```")
        for w in want:
            if w in hdr:
                i = hdr.index(w)
                out.append(f"{w:<72} {r[i][:100]} {units[i]}")
        out.append("```")
    open(os.path.join(P, out_name), "w").write("\n".join(out) + "\n")
    with open(os.path.join(P, "raw", os.path.splitext(out_name)[0] + "_raw.csv"), "w") as f:
        f.write(raw)


def main():
    os.makedirs(os.path.join(P, "raw"), exist_ok=True)
    sweep_table(8, "sweep_n8.jsonl", "TMA-staged P2P tiles on; NVLS available.")
    sweep_table("8_bf16", "sweep_n8_bf16.jsonl", "bf16 (multimem.ld_reduce with fp32 accumulation in the switch).")
    sweep_table(4, "sweep_n4.jsonl", "TMA-staged P2P tiles on; NVLS available.")
    sweep_table(2, "sweep_n2.jsonl", "TMA-staged P2P tiles on; `sta` columns are private (non-symmetric) buffers.")
    tune_table(8, "tunes_n8.jsonl", "register-staged P2P + NVLS")
    tune_table(2, "tunes_n2_tma.jsonl", "TMA-staged P2P + NVLS")
    tune_table(2, "tunes_n2.jsonl", "register-staged P2P + NVLS")
    tune_table(4, "tunes_n4.jsonl", "TMA-staged P2P + NVLS")
    tune_table(8, "tune_n8.jsonl", "256 MiB, threads x CTAs x tile, register-staged P2P")
    tune_table(2, "tune_n2.jsonl", "256 MiB, threads x CTAs x tile, register-staged P2P")
    ncu_summary("prof_staged.ncu-rep", "ncu_staged_tma_emulated.md",
                "TMA-staged flat allreduce, 8 logical ranks on 1 B200, 64 MiB fp32 (ft_allreduce_kernel<float,float,OpSum,4,5>)")
    for f in ("ddp_n8.jsonl", "bench8.json", "bench4.json", "bench2.json", "bench1.json", "bench_ref1.json", "bench_ref2.json",
              "bench_ref4.json", "bench_ref8.json", "launches.csv"):
        if os.path.exists(os.path.join(G, f)):
            shutil.copy(os.path.join(G, f), os.path.join(P, "raw", f))


if __name__ == "__main__":
    main()
