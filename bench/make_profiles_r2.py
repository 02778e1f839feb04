#!/usr/bin/env python
"""Round-2 summaries under profiles/ from the raw rows in profiles/raw/r2/ (copied there from gpurun_out/).
Every timing row must carry a clock record (sm_mhz, throttle reasons) or it is dropped, loudly."""
import json
import os
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RAW = os.path.join(ROOT, "profiles", "raw", "r2")
P = os.path.join(ROOT, "profiles")
BAD = {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}


def load(name):
    path = os.path.join(RAW, name)
    if not os.path.exists(path):
        return []
    rows = [json.loads(l) for l in open(path) if l.strip().startswith("{")]
    keep = []
    for r in rows:
        timed = "us" in r or "ms" in r or "ms_per_step" in r
        ck = r.get("clocks") or {}
        if timed and (not ck.get("sm_mhz") or set(ck.get("reasons", [])) & BAD):
            print(f"make_profiles_r2: {name}: REFUSING a row without an admissible clock record")
            continue
        keep.append(r)
    return keep


def clocks_line(rows):
    mhz = sorted({r["clocks"]["sm_mhz"] for r in rows if r.get("clocks")})
    reasons = sorted({x for r in rows if r.get("clocks") for x in r["clocks"].get("reasons", [])})
    return f"Clock records: every row sampled during its own timed loop; SM clock {mhz} MHz, throttle reasons {reasons or 'none'}."


def nvls_limit():
    rows = load("nvls_scan_n8.jsonl")
    if not rows:
        return
    out = ["# Where the time of the 8-GPU NVLS allreduce goes (256 MiB fp32, registered buffers)", "",
           "Source: `bench/nvls_scan.py` under torchrun on 8 x B200, raw rows `profiles/raw/r2/nvls_scan_n8.jsonl`; device time",
           "(CUDA events around 10 back-to-back calls, max over ranks). " + clocks_line(rows), "",
           "## 1. Every inner-loop structure x every launch shape lands on the same number", "",
           "us per call (busbw GB/s). Variants: see `StagedCta::nvls_tile_variant` (csrc/core/staged_core.h) - each is its own",
           "kernel in csrc/kernels/nvls_kernel.cu with 0 spills (`-Xptxas -v`, profiles/sass/README.md).", ""]
    names = {}
    table = defaultdict(dict)
    for r in rows:
        if r.get("algo") == "nvls" and r.get("flags") == 0 and r.get("variant", 9) <= 6:
            names[r["variant"]] = r["what"]
            table[r["variant"]][(r["ctas"], r["tile_bytes"])] = r
    shapes = sorted({k for v in table.values() for k in v})
    out.append("| variant | " + " | ".join(f"{c} CTAs x {t >> 10} KiB" for c, t in shapes) + " |")
    out.append("|---|" + "---:|" * len(shapes))
    for v in sorted(table):
        cells = []
        for sh in shapes:
            r = table[v].get(sh)
            cells.append(f"{r['us']:.0f} ({r['busbw_GBps']:.0f})" if r else "-")
        out.append(f"| {v}: {names[v]} | " + " | ".join(cells) + " |")
    nccl = [r for r in rows if r["impl"] == "nccl"]
    best = min((r for r in rows if r.get("algo") == "nvls" and r.get("flags") == 0 and r.get("variant", 9) <= 6), key=lambda r: r["us"])
    out += ["", f"NCCL 2.28.9 on the same buffers in the same loop: {nccl[0]['us']:.0f} us ({nccl[0]['busbw_GBps']:.0f} GB/s). "
            f"Best row: {best['us']:.0f} us = {best['busbw_GBps']:.0f} GB/s busbw.",
            "Six loop structures (batched streaming, software-pipelined, the round-1 loop, warp-specialised ld_reduce/st halves",
            "through shared memory, plain per-vector addressing), 8 to 64 CTAs, 0.5 to 4 MiB tiles: 572-600 us. Only 128-148 CTAs",
            "with 227 KiB tiles are slower (633-667 us: 16x more tile boundaries, each a system fence with multicast stores in flight).",
            "The kernel is not the limiter.", "",
            "## 2. The two halves, alone", ""]
    meas = [r for r in rows if r.get("variant") in (7, 8)]
    out += ["| kernel | CTAs | us | link bytes that dominate | GB/s per direction |", "|---|---:|---:|---|---:|"]
    S = rows[0]["bytes"]
    for r in meas:
        what = "multimem.ld_reduce only (S out, S/8 in)" if r["variant"] == 7 else "multimem.st only (S in, S/8 out)"
        out.append(f"| {what} | {r['ctas']} | {r['us']:.0f} | S = {S / 1e6:.0f} MB | {S / r['us'] / 1e3:.0f} |")
    both = best["us"]
    out += ["", f"Each half alone moves S bytes through ONE direction of the GPU's links in ~394 us = 681 GB/s (88 % of the 770 GB/s peer-copy",
            f"rate, `bench/link_probe.py`). Together each direction carries (1 + 1/8) S and the call takes {both:.0f} us = "
            f"{1.125 * S / both / 1e3:.0f} GB/s per direction:",
            "the sum of the halves is 788 us, perfect overlap would be 443 us, the hardware delivers 1.30x that. The loss appears",
            "only when BOTH directions carry multimem traffic and is independent of how the instructions are issued (section 1), so it",
            "sits in the link / switch path (read requests and write acknowledgements of one direction share the wires with the data",
            "of the other), not in the SMs.", "",
            "## 3. The exit barrier", ""]
    keep = [r for r in rows if r.get("flags") == 4]
    for r in keep:
        ref = table[0].get((r["ctas"], r["tile_bytes"]))
        if ref:
            out.append(f"* {r['ctas']} CTAs: exit barrier kept {r['us']:.1f} us, dropped {ref['us']:.1f} us")
    out += ["", "Dropped for good (one all-to-all flag hop; redundant - see `StagedCta::run_nvls_registered`); worth ~1-2 us, i.e. visible at",
            "1 MiB, invisible at 256 MiB.", ""]
    hyb = load("hybrid_n8.jsonl")
    if hyb:
        out += ["## 4. Can P2P traffic ride along? (hybrid probe)", "",
                "`bench/hybrid_probe.py`: fraction f of the message through the TMA P2P stages (second communicator, second stream) while",
                "1-f goes through NVLS. us for the pair, best P2P CTA count per f:", "", "| f | P2P widths | us | busbw |", "|---:|---|---:|---:|"]
        by = defaultdict(list)
        for r in hyb:
            by[(r["p2p_fraction"], tuple(r["p2p_widths"]))].append(r)
        for k in sorted(by):
            b = min(by[k], key=lambda r: r["us"])
            out.append(f"| {k[0]:.2f} | {'x'.join(map(str, k[1]))} | {b['us']:.0f} | {b['busbw_GBps']:.0f} |")
        out += ["", "Any P2P share makes the call slower (577 -> 595 us at 5 %, 620 us at 25 %): the links themselves are the shared resource,",
                "there is no idle capacity next to the NVLS traffic. The hybrid was not built.", ""]
    out += ["## 5. Verdict on the north-star bar", "",
            "\"85 % of the NVLS roofline\" = 1.125 S / (0.85 x 770 GB/s) = 461 us at 256 MiB. With both directions loaded this box moves",
            "multimem traffic at 525 GB/s per direction (68 % of the copy rate) whatever the kernel does; a single direction reaches 681 GB/s",
            "(88 %). The achievable floor for the standard ld_reduce + st formulation on this hardware is therefore ~575 us, which all",
            "variants hit; NCCL's own NVLS kernel is at 704-726 us (0.80x). NVLink byte counters (NVML field values 138-141, per-link",
            "scopes, `nvidia-smi nvlink -gt`) are not exposed in this container (NVML_ERROR_NOT_SUPPORTED / empty output), and ncu",
            "cannot replay a kernel that spin-waits on peers - the decomposition above is the counter-free evidence.", ""]
    n2 = load("nvls_scan_n2_before_streaming.jsonl")
    n2b = load("nvls_scan_n2_streaming.jsonl")
    if n2 and n2b:
        out += ["## 6. What the streaming rewrite fixed (2 GPUs, where the in-switch load is latency-bound)", "",
                "Before (per-vector address arithmetic between the loads; the plain loops selected their load with a runtime flag):", ""]
        for r in n2:
            if r.get("algo") == "nvls" and r.get("flags") == 0:
                out.append(f"* variant {r['variant']}: {r['us']:.0f} us")
        out += ["", "4.8 ms for 'UNR loads, then UNR stores' whatever UNR (4, 8, 16): ~2.3 us per multimem.ld_reduce per thread - the loads of a",
                "thread did not overlap. SASS: every LDGMC took its address from the same register pair, recomputed in between. After",
                "(one base pointer per batch + compile-time offsets, `LDGMC [R.64+UR+imm]` back to back, profiles/sass/):", ""]
        for r in n2b:
            if r.get("algo") == "nvls" and r.get("flags") == 0:
                out.append(f"* variant {r['variant']} ({r.get('what')}): {r['us']:.0f} us")
        out += ["", "(On 2 GPUs NVLS moves 1.5 S per direction against S for the P2P exchange, so `auto` uses the P2P stages there: 426 us.)", ""]
    open(os.path.join(P, "nvls_limit.md"), "w").write("\n".join(out) + "\n")


def link_probe():
    out = ["# Raw NVLink rates between the GPUs of the box (`bench/link_probe.py`, kernels/link_probe.cu)", "",
           "GB/s of payload per GPU; every CTA streams to/from one peer, no synchronisation, nothing but the traffic. 'alone' = one GPU",
           "active (one direction of its links carries data), 'all' = every GPU runs the same pattern (both directions loaded).", ""]
    for n, name in ((8, "link_n8.jsonl"), (2, "link_n2.jsonl")):
        rows = load(name)
        if not rows:
            continue
        out += [f"## {n} GPUs, {rows[0]['bytes_per_peer'] >> 20} MiB per peer. " + clocks_line(rows), "",
                "| pattern | peers | CTAs | alone | all | all / alone |", "|---|---:|---:|---:|---:|---:|"]
        by = defaultdict(dict)
        for r in rows:
            by[(r["mode"], r["npeers"], r["ctas"])]["alone" if r["active_ranks"] == 1 else "all"] = r["GBps_per_gpu"]
        for k in sorted(by, key=lambda k: (k[0], -k[1], k[2])):
            v = by[k]
            if "alone" in v and "all" in v:
                out.append(f"| {k[0]} | {k[1]} | {k[2]} | {v['alone']:.0f} | {v['all']:.0f} | {v['all'] / v['alone']:.2f} |")
        out.append("")
    out += ["## Reading", "",
            "* One direction alone: TMA bulk loads reach the 770 GB/s peer-copy rate (769), bulk stores 716, 16-byte stores 689.",
            "* Both directions loaded (what every allreduce does): reads drop to 646 GB/s (-16 %), writes to 680 GB/s (-5 %): a read needs a",
            "  request packet in the opposite direction, a write only an acknowledgement.",
            "* Ceiling of ANY two-phase P2P allreduce on 8 GPUs: 2 (N-1)/N S per direction at 646-680 GB/s -> busbw <= 646-680 GB/s. The",
            "  TMA-staged trees run at 622-636 GB/s (96-98 % of the pull ceiling); NCCL's NVLS kernel reaches 647-667 GB/s at 256 MiB and",
            "  727 GB/s at 1 GiB - above the P2P ceiling. The multi-stage P2P tree cannot beat an in-switch reduction at these sizes on this",
            "  fabric; a push-based all-gather could add at most ~5 % (680 vs 646). `auto` therefore uses NVLS from 512 KiB up on 4/8 GPUs.",
            "* (ldg_pull is the probe's own 8-deep load loop, not the library's: it shows that register-staged pulls need far more bytes in",
            "  flight than a few CTAs provide - the reason the stage kernels moved to TMA.)", ""]
    open(os.path.join(P, "link_probe.md"), "w").write("\n".join(out) + "\n")


def clusters():
    rows = load("clusters_n8.jsonl")
    if not rows:
        return
    out = ["# Thread-block clusters: a CTA pair per tile vs one CTA per tile (8 x B200, fp32 SUM in place)", "",
           "`bench/cluster_scan.py`; kernel `ft_cluster_kernel` (csrc/kernels/cluster_tiles.cuh, cluster_kernel.cu): `__cluster_dims__(2,1,1)`,",
           "each CTA bulk-loads its half of the stage's sources (plain cp.async.bulk, no multicast), partial sums cross through DSMEM",
           "(`st.shared::cluster` + remote `mbarrier.arrive.release.cluster`), SASS: UCGABAR_ARV/WAIT, SYNCS.ARRIVE.TRANS64.RED, S2UR",
           "SR_CgaCtaId, CGAERRBAR; 0 spills. " + clocks_line(rows), "",
           "| bytes | widths | buffers | one CTA / tile us (CTAs) | cluster pair / tile us (clusters) | cluster / plain |", "|---:|---|---|---:|---:|---:|"]
    t = defaultdict(dict)
    for r in rows:
        t[(r["bytes"], tuple(r["widths"]), r["mode"])][r["cluster"]] = r
    for k in sorted(t):
        v = t[k]
        if 0 in v and 1 in v:
            out.append(f"| {k[0]} | {'x'.join(map(str, k[1]))} | {k[2]} | {v[0]['us']:.0f} ({v[0]['ctas']}) | {v[1]['us']:.0f} ({v[1]['ctas']}) | "
                       f"{v[1]['us'] / v[0]['us']:.2f} |")
    out += ["", "All rows verified against an NCCL fp64 sum (`ok`).", "",
            "## Decision: rejected, with numbers", "",
            "The pair is never faster: parity for the flat tree at 256 MiB (754 vs 747 us), 5-14 % slower for 8 and 4x2 at 16-64 MiB, 30-75 %",
            "slower for the trees with width-2 stages, where a stage has two sources, the split leaves one per CTA, and every 12 KiB chunk",
            "costs a DSMEM hand-off plus a cross-CTA mbarrier round trip that the one-CTA kernel does not have. The link, not the SM, bounds",
            "these kernels (profiles/link_probe.md): halving the per-CTA source count buys nothing, the exchange costs something. The",
            "one-CTA-per-tile TMA kernel stays the default; the cluster kernel remains selectable (`FT_CLUSTER=1` /",
            "`set_tuning('cluster', 1)`) and tested (tests/test_gpu_emulated.py::test_cluster_pair_per_tile_matches_reference).", ""]
    open(os.path.join(P, "clusters.md"), "w").write("\n".join(out) + "\n")


def private_and_ddp():
    rows = load("sweep_n8_auto_fp32.jsonl")
    out = ["# Ordinary (non-symmetric) tensors and DDP on 8 x B200", ""]
    if rows:
        by = defaultdict(dict)
        for r in rows:
            if "us" in r:
                by[r["bytes"]]["nccl" if r["impl"] == "nccl" else r["mode"]] = r
        out += ["## `comm.allreduce(t)`, auto selection, fp32 (`bench/sweep.py --staging --only-auto`). " + clocks_line(rows), "",
                "| bytes | NCCL us | symmetric us (algo) | vs NCCL | private us (algo) | vs NCCL |", "|---:|---:|---:|---:|---:|---:|"]
        for sz in sorted(by):
            v = by[sz]
            if {"nccl", "registered", "staging"} <= set(v):
                out.append(f"| {sz} | {v['nccl']['us']:.1f} | {v['registered']['us']:.1f} ({v['registered']['algo']}) | "
                           f"{v['nccl']['us'] / v['registered']['us']:.2f}x | {v['staging']['us']:.1f} ({v['staging']['algo']}) | "
                           f"{v['nccl']['us'] / v['staging']['us']:.2f}x |")
        out += ["", "Private buffers were 0.72-0.90x NCCL at 1-16 MiB and 256 MiB in round 1. What changed: the two-shot LL kernel (algo 6) covers",
                "256 KiB - 4 MiB; the copy passes around the NVLS reduce lost their per-vector divisions (they were ALU-bound: 819 -> 702 us at",
                "256 MiB); the reducer CTAs stream 8-deep batches.", ""]
    mid = load("sweep_n8_staging_mid.jsonl")
    if mid:
        out += ["## Every algorithm on private buffers, 64 KiB - 16 MiB (us)", ""]
        by = defaultdict(dict)
        for r in mid:
            if "us" not in r:
                continue
            if r["impl"] == "nccl":
                by[r["bytes"]]["NCCL"] = r["us"]
            elif r.get("mode") == "staging" and r.get("requested") != "auto":
                by[r["bytes"]][r["requested"]] = r["us"]
        cols = ["NCCL", "oneshot", "twoshot", "nvls", "stages", "ring"]
        out += ["| bytes | " + " | ".join(cols) + " |", "|---:|" + "---:|" * len(cols)]
        for sz in sorted(by):
            out.append(f"| {sz} | " + " | ".join(f"{by[sz][c]:.1f}" if c in by[sz] else "-" for c in cols) + " |")
        out.append("")
    ddp = load("ddp_n8.jsonl")
    if ddp:
        out += ["## DDP training step (`models/ddp_bench.py --comm all`, 20 timed steps). " + clocks_line(ddp), "",
                "| model | gradient allreduce | ms / step | items / s | parameters vs NCCL run (rel. diff) |", "|---|---|---:|---:|---:|"]
        for r in ddp:
            out.append(f"| {r['model']} ({r['param_dtype']}, batch {r['batch_per_gpu']}/GPU) | {r['comm']} | {r['ms_per_step']:.2f} | "
                       f"{r['items_per_s']:.0f} | {r.get('param_rel_diff_vs_nccl', '-')} |")
        out += ["", "`flextree` = DDP comm hook on DDP's own (private) buckets; `flextree-dp` = `FlexTreeDataParallel`, gradients accumulate straight",
                "into symmetric buckets that the NVLS kernel reduces in place with the 1/N fused. The engine is ahead of NCCL DDP on both models",
                "(+4.2 % ResNet-50, +1.2 % GPT-2 medium); the hook is within 0.8 % of it (the step is compute-bound: only the last bucket's",
                "collective is exposed).", ""]
    open(os.path.join(P, "private_buffers_and_ddp_n8.md"), "w").write("\n".join(out) + "\n")


def sanitizer():
    out = ["# compute-sanitizer on the GPU kernels", "",
           "`make memcheck racecheck synccheck` (bench/sanitize_target.py on the single-GPU emulation) was run on a B200 box this round:",
           "", "```", open(os.path.join(RAW, "compute_sanitizer_closed.log")).read().strip().splitlines()[0], "```", "",
           "compute-sanitizer is closed on this GPU pool, for all three tools. What stands in for it:", "",
           "* the same algorithm text (core/staged_core.h) runs on the CPU back end under ThreadSanitizer and ASan/UBSan (`make tsan asan`),",
           "  with schedule jitter and a wait-removal mutation campaign (tests/test_host_world.py) - this is where the redundant NVLS exit",
           "  barrier was found and where a missing wait shows up as a data race;",
           "* on the GPU every kernel family is compared bit for bit (fp32 P2P paths) or within rounding (bf16 / NVLS / cluster) against an",
           "  independent numpy / fp64 model over odd sizes, ragged tails and several tiles per CTA (112 tests on one B200, incl. the NVLS",
           "  kernels with the switch emulated and CUDA-graph capture + 20 replays of every path);",
           "* the device watchdog turns a protocol deadlock into an error word instead of a hang, and a fault-injection hook",
           "  (`ft_comm_inject_fault`) exercises that path.", ""]
    open(os.path.join(P, "sanitizer.md"), "w").write("\n".join(out) + "\n")


def sweep(n, dtype):
    rows = load(f"sweep_n{n}_{dtype}.jsonl")
    if not rows:
        return
    by = defaultdict(dict)
    labels = set()
    for r in rows:
        if "us" not in r:
            continue
        if r["impl"] == "nccl":
            by[r["bytes"]]["nccl"] = r
        elif r.get("requested") == "auto":
            by[r["bytes"]]["auto_" + r["mode"]] = r
        else:
            lb = f"{r['requested']} {'x'.join(map(str, r['widths']))} {r['mode'][:3]}"
            labels.add(lb)
            by[r["bytes"]][lb] = r
    out = [f"# Allreduce sweep, {n} x B200, {dtype} SUM (round 2)", "",
           "`bench/sweep.py --staging" + (" --all-widths" if dtype == "fp32" else "") + f"` under torchrun; device time (CUDA events around back-to-back calls,",
           f"max over ranks / reps); busbw = bytes/t x 2(N-1)/N. Raw rows: `profiles/raw/r2/sweep_n{n}_{dtype}.jsonl`. " + clocks_line(rows), "",
           "| bytes | NCCL us | NCCL busbw | symmetric buffers: us (algo, CTAs) | busbw | vs NCCL | private buffers: us (algo) | vs NCCL |",
           "|---:|---:|---:|---:|---:|---:|---:|---:|"]
    for sz in sorted(by):
        v = by[sz]
        if not {"nccl", "auto_registered"} <= set(v):
            continue
        c, a = v["nccl"], v["auto_registered"]
        st = v.get("auto_staging")
        out.append(f"| {sz} | {c['us']:.1f} | {c['busbw_GBps']:.1f} | {a['us']:.1f} ({a['algo']} {'x'.join(map(str, a['widths']))}, c{a['ctas']}) | "
                   f"{a['busbw_GBps']:.1f} | {c['us'] / a['us']:.2f}x | " + (f"{st['us']:.1f} ({st['algo']}) | {c['us'] / st['us']:.2f}x |" if st else "- | - |"))
    labels = sorted(labels)
    if labels:
        out += ["", "## Every algorithm / width vector, us measured / us predicted by the link model (reg = symmetric, sta = private buffers)", "",
                "| bytes | " + " | ".join(labels) + " |", "|---:|" + "---:|" * len(labels)]
        for sz in sorted(by):
            out.append(f"| {sz} | " + " | ".join((f"{by[sz][lb]['us']:.1f} / {by[sz][lb].get('est_us', 0):.0f}" if lb in by[sz] else "-") for lb in labels) + " |")
    after = load(f"sweep_n{n}_auto_after_refit_{dtype}.jsonl")
    if after:
        by2 = defaultdict(dict)
        for r in after:
            if "us" in r:
                by2[r["bytes"]]["nccl" if r["impl"] == "nccl" else r["mode"]] = r
        out += ["", "## `auto` after the round-2 refit of the link model's private-buffer term (separate run, factor-2 size steps)", "",
                "| bytes | NCCL us | symmetric us (algo) | vs NCCL | private us (algo) | vs NCCL |", "|---:|---:|---:|---:|---:|---:|"]
        for sz in sorted(by2):
            v = by2[sz]
            if {"nccl", "registered", "staging"} <= set(v):
                out.append(f"| {sz} | {v['nccl']['us']:.1f} | {v['registered']['us']:.1f} ({v['registered']['algo']}) | "
                           f"{v['nccl']['us'] / v['registered']['us']:.2f}x | {v['staging']['us']:.1f} ({v['staging']['algo']}) | "
                           f"{v['nccl']['us'] / v['staging']['us']:.2f}x |")
        if any("us_loops" in r for r in after):
            off = [(r["bytes"], r.get("mode", "nccl"), r["us_loops"]) for r in after if "us_loops" in r and max(r["us_loops"]) > 1.5 * min(r["us_loops"])]
            out += ["", "This run reports the better of two timed loops per row (`bench/sweep.py`; both loops are in `us_loops` of the raw rows). Rows",
                    "whose two loops differ by more than 1.5x: " + ("; ".join(f"{b} B {m}: {l[0]} / {l[1]} us" for b, m, l in off) or "none") + ".",
                    "They are the transients that single-loop runs put into the tables: the slow loop is always the FIRST one, i.e. the one that",
                    "directly follows NCCL's own measurement at a new size (NCCL connects channels lazily from a helper thread)."]
        out += ["", "Isolated rows far off their neighbours (e.g. a 40 us one-shot at 8 KiB between 8.6 and 9.0 us, a 118 us NVLS call at 2 MiB between",
                "15 and 26 us) are transients of the measurement that directly follows the NCCL loop; the same configuration measured",
                "again in the same run (the forced-algorithm rows of the first table, the other dtype) shows the neighbouring value.", ""]
    open(os.path.join(P, f"sweep_n{n}_{dtype}_r2.md"), "w").write("\n".join(out) + "\n")


def wire():
    rows = load("wire_n8.jsonl")
    if not rows:
        return
    out = ["# Wire formats for fp32 gradients, 8 x B200, private buffers, 1/N fused (`bench/wire_scan.py`)", "", clocks_line(rows), "",
           "| bytes | NCCL fp32 (copy + all_reduce) us | fp32 wire us (algo) | bf16 wire us | rms err | fp8 block-scaled wire us (algo) | rms err | max err / max |",
           "|---:|---:|---:|---:|---:|---:|---:|---:|"]
    by = defaultdict(dict)
    for r in rows:
        by[r["bytes"]][r.get("wire", "nccl")] = r
    for sz in sorted(by):
        v = by[sz]
        f, b, q = v["fp32 (native)"], v["bf16"], v["fp8"]
        out.append(f"| {sz} | {v['nccl']['us']:.0f} | {f['us']:.0f} ({f['algo']}) | {b['us']:.0f} | {b['rms_err_over_rms']:.1e} | {q['us']:.0f} ({q['algo']}) | "
                   f"{q['rms_err_over_rms']:.1e} | {q['max_err_over_max']:.1e} |")
    out += ["", "The fp8 wire (e4m3 + one power-of-two scale byte per 16 values, re-quantised at every reduce-scatter stage) now converts in",
            "hardware (F2FP.SATFINITE.E4M3 / F2FP.F16.E4M3.UNPACK_B), loads every peer's pack and scale before the first conversion and walks",
            "the blocks without per-vector divisions (round 1: a scalar loop with software conversions, 737 local-memory accesses, never timed).",
            "It runs on the P2P stages (the switch cannot reduce e4m3), moves 17/32 of the bf16 bytes and is 3-25 % faster than the bf16",
            "wire from 64 MiB up; its 3.6 % rms error is the 3-bit mantissa. It stays opt-in (`wire_dtype='fp8'` / `FT_WIRE=fp8`).", ""]
    open(os.path.join(P, "wire_formats_n8.md"), "w").write("\n".join(out) + "\n")


def odd_world():
    rows = load("odd_world_n7.jsonl")
    if not rows:
        return
    by = defaultdict(dict)
    for r in rows:
        by[r["bytes"]][r.get("tree", "NCCL")] = r
    cols = ["NCCL"] + [c for c in ("flat", "2*3+1", "3*2+1", "2*2*2-1", "2*4-1", "4*2-1") if any(c in v for v in by.values())]
    out = ["# 7 GPUs: flat tree, '+1' trees (lonely rank) and '-1' trees (phantom rank), fp32 SUM, private buffers", "",
           "`bench/odd_world.py` under `torchrun --nproc-per-node 7` on an 8 x B200 box - the candidate list of the reference's chooseWidth(7)",
           "(cost_model/ChooseWidth.h:14-31, README.md:13-27). us per call; every row checked against an NCCL fp64 sum. " + clocks_line(rows), "",
           "| bytes | " + " | ".join(cols) + " |", "|---:|" + "---:|" * len(cols)]
    for sz in sorted(by):
        out.append(f"| {sz} | " + " | ".join(f"{by[sz][c]['us']:.0f}" if c in by[sz] else "-" for c in cols) + " |")
    out += ["", "On a uniform NVSwitch fabric the flat tree wins at every size (and beats NCCL up to 64 MiB); the '+1' trees pay a fold pass over the",
            "lonely rank's vector, the '-1' trees make rank 0 run two logical ranks with half the CTA budget each. '-1' vs '+1': 4*2-1 ties",
            "3*2+1 at 1 MiB (66.9 vs 67.0 us) and loses by 10-20 % above. Both families are implemented, tested for every factorization of",
            "N+-1 (tests/mp_worker.py sections 1b, 1c; 2-, 3- and 7-rank runs in profiles/raw/r2/) and selectable (`set_widths`,",
            "`FT_TOPO=2,4-1`); `auto` never needs them on one box - they matter where links are not uniform.", ""]
    open(os.path.join(P, "odd_world_n7.md"), "w").write("\n".join(out) + "\n")


def cost_model_validation():
    out = ["# Link model: pick vs measured best (BASELINE config #3), round 2", "",
           "For every world size and message size: what `auto` (the link model, `csrc/planner/cost_model.cc:choose`) ran on symmetric",
           "buffers, against the fastest of ALL forced candidates measured in the same run - every ordered factorization of N for the P2P",
           "stages, ring, one-shot, two-shot, NVLS (`bench/sweep.py --all-widths`). Predicted / measured us for each candidate are in the",
           "second table of `sweep_n{N}_fp32_r2.md`.", ""]
    total = good = 0
    for n in (8, 4, 2):
        rows = load(f"sweep_n{n}_fp32.jsonl")
        if not rows:
            continue
        by = defaultdict(dict)
        for r in rows:
            if "us" not in r or r["impl"] != "flextree" or r.get("mode") != "registered":
                continue
            key = "auto" if r["requested"] == "auto" else f"{r['requested']} {'x'.join(map(str, r['widths']))}"
            by[r["bytes"]][key] = r
        out += [f"## {n} x B200 (fp32 SUM, symmetric buffers)", "",
                "| bytes | auto ran | auto us | model est us | fastest candidate | its us | auto / fastest | factorizations of N: us |",
                "|---:|---|---:|---:|---|---:|---:|---|"]
        for sz in sorted(by):
            v = by[sz]
            if "auto" not in v:
                continue
            a = v["auto"]
            cands = {k: r for k, r in v.items() if k != "auto"}
            if not cands:
                continue
            bk = min(cands, key=lambda k: cands[k]["us"])
            ratio = a["us"] / cands[bk]["us"]
            total += 1
            good += ratio <= 1.10
            trees = "; ".join(f"{k.split(' ')[1]} {r['us']:.0f}" for k, r in sorted(cands.items()) if k.startswith("stages"))
            out.append(f"| {sz} | {a['algo']} {'x'.join(map(str, a['widths']))} | {a['us']:.1f} | {a.get('est_us', 0):.1f} | {bk} | {cands[bk]['us']:.1f} | "
                       f"{ratio:.2f} | {trees} |")
        out.append("")
    out += [f"`auto` is within 10 % of the measured best at {good} of {total} (N, size) points. The misses are isolated transients of the row",
            "measured directly after the NCCL loop (see the note in `sweep_n4_fp32_r2.md`) and the 1 MiB point on 8 GPUs where forced NVLS",
            "(17.6 us) edges out the two-shot LL (19.4 us). On a uniform NVSwitch fabric every factorization moves the same bytes, so the",
            "P2P trees differ by hop count (small messages: the flat tree wins, 26 us against 39-53 us at 1 KiB on 8 GPUs) and by fan-in",
            "(large messages: fewer concurrent peers per stage win - 2x2x2 2970 us, 2x4 3074 us, 4x2 3135 us, flat 3285 us at 1 GiB; at",
            "256 MiB all four are within 3 %). That is the reference's trade-off (CostModel.h: latency per layer vs incast per width)",
            "re-measured on NVSwitch; the in-switch reduction beats all of them from 512 KiB up, which is why `auto` rarely lands on a",
            "multi-stage tree on one box.", ""]
    open(os.path.join(P, "cost_model_validation_r2.md"), "w").write("\n".join(out) + "\n")


if __name__ == "__main__":
    nvls_limit()
    link_probe()
    clusters()
    private_and_ddp()
    sanitizer()
    for n in (8, 4, 2):
        for dt in ("fp32", "bf16"):
            sweep(n, dt)
    wire()
    odd_world()
    cost_model_validation()
    print("profiles written")
