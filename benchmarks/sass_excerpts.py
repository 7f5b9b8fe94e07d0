#!/usr/bin/env python
"""Regenerates profiles/sass_excerpts.txt and profiles/sass_mnemonics.txt from the built
library (no GPU needed): the SASS around the instructions that prove each mechanism."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "parameter_server_distributed_b200", "_native", "libpsb_kernels.so")

SECTIONS = [
    ("gemm2_bf16_kernelILi0E",
     ["UTCHMMA", "UTMALDG", "UGETNEXTWORKID", "UTCBAR", "LDTM", "UTCATOMSWS", "SYNCS.ARRIVE.TRANS64.RED",
      "UTMASTG", "FENCE.VIEW.ASYNC"],
     "CTA-pair tcgen05 GEMM (plain epilogue): TMA loads, UMMA issue, commit, TMEM loads, cluster launch "
     "control, and the staged epilogue (STS into the swizzled piece, proxy fence, UTMASTG)"),
    ("gemm2_bf16_kernelILi2E", ["MUFU.TANH", "LDG.E.128", "UTCHMMA"],
     "same kernel, dX*dGELU epilogue: tanh.approx and the 16-byte loads of Z"),
    ("push_fused_kernelILi3ELi1ELi8E",
     ["LDG.E.128", "LDG.E.NA", "STG.E", "256", "ATOM", "MEMBAR", "LD.E.STRONG.SYS", "ST.E.STRONG.SYS"],
     "K1 register path, AdamW, bf16 gradients, 8 peers"),
    ("push_fused_tma_kernelILi0ELi1E", ["UBLKCP", "SYNCS", "STG.E"], "K1 bulk-copy path (SGD)"),
    ("push_nvls_kernelILi0E", ["LDGMC", "STG.E.128.STRONG.SYS", "MEMBAR", "LD.E.STRONG.SYS"],
     "K1 over NVLS (SGD): multimem.ld_reduce = LDGMC...HPADD.BF16x8 (the switch adds the 8 replicas), "
     "multimem.st = STG.E.128.STRONG.SYS on the multicast address"),
    ("gemm2_bf16_kernelILi1E", ["UTMASTG", "UTMALDG.2D ", "UTMACMDFLUSH", "RED.E", "REDG"],
     "fused all-gather GEMM, gather role: TMA loads on the OWNER's tensor map (UTMALDG.2D without "
     ".2CTA), TMA stores into the local replica (UTMASTG.2D), batched arrival counters"),
    ("colsum_bf16_kernel", ["LDG", "STG", "ATOM", "BAR"], "bias-gradient column sum"),
    ("pull_allgather_kernel", ["LD.E.STRONG.SYS", "LDG.E.128", "LDG.E.NA", "STG.E.128", "CCTL", "MEMBAR",
                               "NANOSLEEP"],
     "K2 pull all-gather: version wait (ld.acquire.sys = LD.E.STRONG.SYS in a NANOSLEEP back-off loop), "
     "16-byte peer loads that bypass L1, 16-byte stores into the local replica; the seqlock variant "
     "re-reads the version after the copy"),
    ("wait_versions_kernel", ["LD.E.STRONG.SYS", "NANOSLEEP", "BPT.TRAP", "LD.E.STRONG"],
     "K9 version wait: one warp spins on the owners' version words (acquire at system scope), polls the "
     "host-mapped abort word, traps when the watchdog expires"),
    ("device_barrier_kernel", ["ST.E.STRONG.SYS", "LD.E.STRONG.SYS", "MEMBAR", "NANOSLEEP"],
     "K9 device barrier over the signal pad: release store of the epoch into every peer's pad, acquire "
     "spin on the own pad"),
    ("push_ready_kernel", ["ST.E.STRONG.SYS", "LD.E.STRONG.SYS", "MEMBAR", "NANOSLEEP"],
     "READY handshake of a bucket taken off the SMs that run K1 (round 2): post the own flag to every "
     "owner, wait for all workers' flags"),
    ("norm_allreduce_kernel", ["ST.E.STRONG.SYS", "LD.E.STRONG.SYS", "MUFU.RSQ", "MUFU.SQRT", "FMNMX",
                               "MEMBAR"],
     "K4 cross-shard gradient norm: every rank posts its partial sum of squares into every pad, spins "
     "for the others, sums in rank order (bit-identical on every rank) and derives the clip coefficient"),
    ("sumsq_partial_kernelILi1E", ["LDG.E.128", "SHFL", "FFMA", "STG"],
     "K4 partial sum of squares of a bf16 buffer (16-byte loads, warp shuffles, one partial per CTA)"),
    ("optim_kernelILi3ELi1ELb1E", ["LDG.E.128", "STG.E.128", "MUFU.RSQ", "MUFU.RCP", "MUFU.SQRT", "F2FP"],
     "K3 stand-alone flat-shard AdamW, bf16 gradients, bf16 emit: 16-byte loads of g / p / m / v, one pass, "
     "packed bf16 conversion of the new weights"),
    ("bcast_slice_nvls_kernel", ["STG.E.128.STRONG.SYS", "ST.E.STRONG.SYS", "MEMBAR", "LDG.E.128"],
     "K2 over NVLS: the owner stores its updated slice once to the multicast address (multimem.st), the "
     "switch replicates it into all replicas"),
    ("async_gate_kernel", ["LD.E.STRONG.SYS", "NANOSLEEP", "ISETP"],
     "async / SSP gate: blocks the stream until the ring slot of the next clock is free on every owner "
     "and (SSP) the slowest worker is within the staleness bound"),
    ("async_plan_kernel", ["LD.E.STRONG.SYS", "ST.E", "STG"],
     "async / SSP owner plan: per bucket, which workers have an unapplied clock (mask) and where their "
     "gradient ring slot is (offset)"),
    ("sparse_pull_kernel", ["LDG.E.NA.128", "STG.E.NA.128", "LDG.E.64", "STG.E.128"],
     "K5 sparse pull: a sub-warp per requested row, 16-byte loads that bypass L1 from the owner's bf16 "
     "shard (a peer address over NVLink), 16-byte stores; zero rows for padding ids"),
    ("sparse_mark_kernel", ["ATOMG", "LDG.E.64.STRONG.SYS", "LDG.E.STRONG.SYS", "STG.E"],
     "K5 sparse push, mark: strong loads of the workers' staged counts / ids, atomicCAS claims the row, "
     "an atomic counter hands out the accumulator slot"),
    ("sparse_accumulate_kernelILi1E", ["REDG", "LDG.E.NA.128", "LDG.E.64.STRONG.SYS"],
     "K5 sparse push, accumulate (bf16 gradients): staged rows read from the worker's heap range, added "
     "into the slot's fp32 row with 16-byte vector reductions (REDG.E.ADD.F32x4)"),
    ("sparse_apply_kernelILi3E", ["LDG.E.128", "STG.E.128", "MUFU.RSQ", "MUFU.SQRT", "F2FP"],
     "K5 sparse push, apply (AdamW): one sub-warp per touched row - accumulator row read and zeroed, "
     "master / m / v updated, bf16 row rewritten, slot released"),
]
MNEMONICS = (r"\b(LDTM|SYNCS|UBLKCP|UCGABAR|UTCATOMSWS|UTCBAR|UTCHMMA|UTMALDG|UTMASTG|UTMACCTL|"
             r"UGETNEXTWORKID|UTMACMDFLUSH|LDGMC|LDG\.E[A-Z0-9.]*256|STG\.E[A-Z0-9.]*256)[A-Za-z0-9_.]*")


def main() -> int:
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    funcs, cur = {}, None
    for line in sass.split("\n"):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
            funcs[cur].append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", line).rstrip())
    out = ["# Excerpts of `cuobjdump -sass parameter_server_distributed_b200/_native/libpsb_kernels.so` (sm_100a):",
           "# the instructions that prove each mechanism, with two lines of context; '...' marks a gap.",
           "# Regenerate with `python benchmarks/sass_excerpts.py`.  Full counts: sass_mnemonics.txt."]
    for sub, pats, title in SECTIONS:
        name = next((k for k in funcs if sub in k), None)
        if name is None:
            print("missing kernel", sub, file=sys.stderr)
            return 1
        ins = funcs[name]
        keep = set()
        for i, l in enumerate(ins):
            if any(p in l for p in pats):
                keep.update(range(max(0, i - 2), min(len(ins), i + 3)))
        out += ["", f"## {title}", f"# {name}  ({len(ins)} instructions)"]
        last = -2
        for n, i in enumerate(sorted(keep)):
            if n >= 140:
                break
            if i != last + 1:
                out.append("        ...")
            out.append(ins[i])
            last = i
    with open(os.path.join(ROOT, "profiles", "sass_excerpts.txt"), "w") as f:
        f.write("\n".join(out) + "\n")
    counts = {}
    for m in re.finditer(MNEMONICS, sass):
        counts[m.group(0)] = counts.get(m.group(0), 0) + 1
    with open(os.path.join(ROOT, "profiles", "sass_mnemonics.txt"), "w") as f:
        for k in sorted(counts):
            f.write(f"{counts[k]:7d} {k}\n")
    # registers / static shared memory / stack of every kernel (dynamic shared memory is set at
    # launch: 197 KB for the CTA-pair GEMM, see gemm_tcgen05_2cta.cu SMEM_BYTES)
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    rows, name = [], None
    for line in res.split("\n"):
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            name = m.group(1)
            continue
        m = re.match(r"\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
        if m and name:
            rows.append((name, *map(int, m.groups())))
            name = None
    demangled = subprocess.run(["c++filt"], input="\n".join(r[0] for r in rows), capture_output=True,
                               text=True, check=True).stdout.split("\n")
    with open(os.path.join(ROOT, "profiles", "kernel_resources.txt"), "w") as f:
        f.write("# `cuobjdump -res-usage` of libpsb_kernels.so (sm_100a): registers per thread, stack bytes,\n"
                "# static shared bytes, local bytes.  Regenerate with `python benchmarks/sass_excerpts.py`.\n")
        f.write(f"{'REG':>4} {'STACK':>6} {'SHARED':>7} {'LOCAL':>6}  kernel\n")
        for (n, reg, stack, shared, local), d in sorted(zip(rows, demangled), key=lambda x: x[1]):
            f.write(f"{reg:4d} {stack:6d} {shared:7d} {local:6d}  {re.sub(r'\(.*', '', d.replace('(anonymous namespace)::', ''))[:100]}\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
