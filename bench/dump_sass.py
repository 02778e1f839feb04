#!/usr/bin/env python
"""Regenerate profiles/sass/*.sass and its README table from the built library (no GPU needed).

  python bench/dump_sass.py [--lib allreduce-over-mpi_b200/lib/libflextree.so]

One file per representative kernel instantiation (`cuobjdump -sass -fun <mangled name>`), plus a table of the
mnemonics that prove the Blackwell/NVLink-native paths (the list in B200_PROFILING.md)."""
import argparse
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
K = "_ZN2ft19ft_allreduce_kernelI{t}NS_5OpSumELi{ve}ELi{algo}EEEvNS_10StagedArgsE"
NV = "_ZN2ft14ft_nvls_kernelI{t}NS_5OpSumELi{v}ELb0EEEvNS_10StagedArgsE"
KERNELS = {
    "f32_sum_oneshot_ll": K.format(t="ff", ve=4, algo=1),
    "f32_sum_twoshot_ll": K.format(t="ff", ve=4, algo=6),
    "f32_sum_stages_ldg": K.format(t="ff", ve=4, algo=2),
    "f32_sum_stages_tma": K.format(t="ff", ve=4, algo=5),
    "f32_sum_stages_cluster_pair": "_ZN2ft17ft_cluster_kernelIffNS_5OpSumEEEvNS_10StagedArgsE",
    # NVLS on registered buffers: the dedicated kernel, default variant (streaming 8) and the alternatives that were measured
    "f32_sum_nvls_registered_stream8": NV.format(t="f", v=0),
    "f32_sum_nvls_registered_stream16": NV.format(t="f", v=1),
    "f32_sum_nvls_registered_stream8_pipelined": NV.format(t="f", v=2),
    "f32_sum_nvls_registered_warp_specialised": NV.format(t="f", v=5),
    "bf16_sum_nvls_registered_stream8": NV.format(t="NS_6bf16_tE", v=0),
    # NVLS on private buffers (copy-in / reduce / copy-out roles) lives in the general kernel
    "f32_sum_nvls_private": K.format(t="ff", ve=4, algo=3),
    "f32_sum_ring": K.format(t="ff", ve=4, algo=4),
    "fp8wire_f32_sum_stages": K.format(t="NS_5fp8_tEf", ve=16, algo=2),
}
COLS = [("UBLKCP (TMA bulk copy)", r"\bUBLKCP"), ("SYNCS (mbarrier)", r"\bSYNCS"), ("LDGMC (multimem.ld_reduce)", r"\bLDGMC"),
        ("LDG.*STRONG.SYS (peer loads/flags)", r"\bLDG\.[A-Z0-9.]*STRONG\.SYS"),
        ("STG.*STRONG.SYS (flags, LL lines, multimem.st)", r"\bSTG\.[A-Z0-9.]*STRONG\.SYS"),
        ("MEMBAR.ALL.SYS", r"\bMEMBAR\.ALL\.SYS"), ("CCTL.IVALL", r"\bCCTL\.IVALL"), ("FENCE.VIEW.ASYNC", r"\bFENCE\.VIEW\.ASYNC"),
        ("UCGABAR (cluster barrier)", r"\bUCGABAR"), ("F2FP e4m3 (hardware fp8 conversion)", r"\bF2FP\.[A-Z0-9._]*E4M3"),
        ("LDL+STL (local-memory accesses)", r"\b(?:LDL|STL)\b"), ("registers", None), ("stack bytes", "STACK")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", default=os.path.join(ROOT, "allreduce-over-mpi_b200", "lib", "libflextree.so"))
    args = ap.parse_args()
    out_dir = os.path.join(ROOT, "profiles", "sass")
    os.makedirs(out_dir, exist_ok=True)
    names = subprocess.run(["cuobjdump", "-res-usage", args.lib], capture_output=True, text=True, check=True).stdout
    usage, stack = {}, {}
    cur = None
    for ln in names.splitlines():
        m = re.search(r"Function (\S+):", ln)
        if m:
            cur = m.group(1)
        m = re.search(r"REG:(\d+) STACK:(\d+)", ln)
        if m and cur:
            usage[cur] = int(m.group(1))
            stack[cur] = int(m.group(2))
            cur = None
    rows = []
    for name, mangled in sorted(KERNELS.items()):
        if not mangled or mangled not in usage:
            print("missing:", name, mangled)
            continue
        sass = subprocess.run(["cuobjdump", "-sass", "-fun", mangled, args.lib], capture_output=True, text=True, check=True).stdout
        # keep the instruction text only: drop the encoding words (half of the bytes, nothing to read)
        lines = []
        for ln in sass.splitlines():
            ln = re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", ln)
            if ln.strip():
                lines.append(ln.rstrip())
        sass = "\n".join(lines) + "\n"
        with open(os.path.join(out_dir, name + ".sass"), "w") as f:
            f.write(sass)
        row = [name + ".sass", str(sass.count("\n"))]
        for _, pat in COLS:
            if pat is None:
                row.append(str(usage[mangled]))
            elif pat == "STACK":
                row.append(str(stack[mangled]))
            else:
                row.append(str(len(re.findall(pat, sass))))
        rows.append(row)
    with open(os.path.join(out_dir, "README.md"), "w") as f:
        f.write("# SASS evidence (cuobjdump -sass of allreduce-over-mpi_b200/lib/libflextree.so, sm_100a)\n\n")
        f.write("Regenerate with `python bench/dump_sass.py` (no GPU needed). Counts of the mnemonics that prove the\n"
                "Blackwell/NVLink-native paths (see B200_PROFILING.md table):\n\n")
        f.write("| kernel | lines | " + " | ".join(c for c, _ in COLS) + " |\n|---|---:|" + "---:|" * len(COLS) + "\n")
        for r in rows:
            f.write("| " + " | ".join(r) + " |\n")
        f.write("\nNo tensor-core mnemonics (UTC*MMA / HMMA) appear: an allreduce has no matrix product; \"Blackwell-native\" "
                "here is the async-copy (TMA), mbarrier and multimem ISA.\n")
        nz = sum(1 for v in stack.values() if v > 64)
        f.write(f"\n{len(usage)} kernels in the library (`cuobjdump -res-usage`): at most {max(usage.values())} registers (128 is the cap of the 512-thread "
                f"kernels, the 256-thread NVLS kernels may use 255), LOCAL:0 everywhere; {len(usage) - nz} keep <= 64 bytes of stack, {nz} keep more (peer-pointer "
                f"tables and, for the 1-byte types at 16 elements per vector, accumulator arrays) - the widest is {max(stack.values())} bytes.\n")
    print(open(os.path.join(out_dir, "README.md")).read())


if __name__ == "__main__":
    main()
