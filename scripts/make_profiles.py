#!/usr/bin/env python
"""Turns what a GPU run left in gpurun_out/ (ncu reports, sweep .jsonl, probe output) and the built
library into the tracked summaries under profiles/.  Runs on the GPU-less build box."""
from __future__ import annotations

import csv
import io
import json
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
GP = os.path.join(ROOT, "gpurun_out")
LIB = os.path.join(ROOT, "akka_allreduce_b200", "_native", "libaar_cuda.so")


def sh(cmd):
    return subprocess.run(cmd, capture_output=True, text=True).stdout


def sass():
    txt = sh(["cuobjdump", "-sass", LIB])
    parts = re.split(r"\n\s*Function : ", txt)
    rows, listing = [], None
    pats = ["UBLKCP.S.G", "UBLKCP.G.S", "SYNCS.ARRIVE.TRANS64", "SYNCS.PHASECHK.TRANS64.TRYWAIT", "LDGMC.E.ADD.F32",
            "LDGMC.E.HPADD.BF16", "STG.E.128", "LDG.E.128.STRONG.SYS", "MEMBAR.SC.SYS", "MEMBAR.ALL.SYS", "MEMBAR.SC.GPU",
            "ATOMG.E.ADD.STRONG.GPU", "ATOMG.E.CAS.STRONG.GPU", "HMMA", "UTCHMMA"]
    for p in parts[1:]:
        name = p.split("\n", 1)[0].strip()
        dem = sh(["c++filt", name]).strip()
        ins = re.findall(r"/\*[0-9a-f]{4,6}\*/\s+(.*?)\s*;", p)
        row = {"kernel": dem, "instructions": len(ins)}
        for pat in pats:
            row[pat] = sum(1 for i in ins if pat in i)
        rows.append(row)
        if "aar_stream_kernel<float, true>" in dem:  # the headline kernel: full listing, addresses + operands
            body = [re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", l).rstrip() for l in p.split("\n")[1:]]
            listing = (dem, [l for l in body if l.strip()])
    with open(os.path.join(OUT, "sass_evidence.md"), "w") as f:
        f.write("# SASS evidence (cuobjdump -sass akka_allreduce_b200/_native/libaar_cuda.so, sm_100a)\n\n")
        f.write("What the mnemonics prove: `UBLKCP.*` = TMA bulk copies (`cp.async.bulk`) staging tiles global->shared->global "
                "(the store side also targets peer GPUs); `SYNCS.ARRIVE.TRANS64` / `SYNCS.PHASECHK...TRYWAIT` = mbarrier "
                "expect_tx / try_wait; `LDGMC.E.ADD.F32x4` / `LDGMC.E.HPADD.BF16` = `multimem.ld_reduce` (sum inside the NVSwitch; the matching "
                "`multimem.st` is emitted as `STG.E.128.STRONG.SYS` on the multicast alias, right after the LDGMC group in "
                "stream_kernel_f32_nvls.sass); in the stream kernel `UBLKCP.G.S` is the TMA fill of the per-element counts; "
                "`*.STRONG.SYS` loads/stores = flag traffic and peer-visible data; `MEMBAR.SC.SYS` only in the publisher warp / "
                "finish path.  No HMMA/UTC*MMA anywhere: the path has no GEMM, so tcgen05/TMEM is not used (north star).\n\n")
        keys = ["kernel", "instructions"] + pats
        f.write("| " + " | ".join(keys) + " |\n|" + "---|" * len(keys) + "\n")
        for r in rows:
            f.write("| " + " | ".join(str(r[k]) for k in keys) + " |\n")
    if listing:
        with open(os.path.join(OUT, "stream_kernel_f32_nvls.sass"), "w") as f:
            f.write(f"// cuobjdump -sass akka_allreduce_b200/_native/libaar_cuda.so, function {listing[0]} (sm_100a)\n"
                    "// (instruction encodings stripped; addresses and operands as printed; regenerate with scripts/make_profiles.py)\n"
                    "// LDGMC.E.ADD.F32x4 = multimem.ld_reduce, STG.E.128.STRONG.SYS on the multicast alias = multimem.st,\n"
                    "// UBLKCP = cp.async.bulk (count fill), MEMBAR.SC.SYS / .GPU = fences, LDG.E.STRONG.SYS + CCTL.IVALL = ld.acquire.sys\n")
            f.write("\n".join(listing[1]) + "\n")


def ncu_summary(rep, out_md, title):
    raw = sh(["ncu", "-i", rep, "--page", "raw", "--csv"])
    rows = list(csv.reader(io.StringIO(raw)))
    if len(rows) < 3:
        return
    hdr, units, vals = rows[0], rows[1], rows[2]
    want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
            "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor",
            "smsp__inst_executed.sum"]
    src = sh(["ncu", "-i", rep, "--page", "source", "--csv"])
    srows = list(csv.reader(io.StringIO(src)))
    with open(out_md, "w") as f:
        f.write(f"# {title}\n\n`ncu --set full --clock-control none --import-source on` (numbers under the profiler are for "
                "shares and traffic, never benchmark values).\n\n| metric | unit | value |\n|---|---|---|\n")
        for i, h in enumerate(hdr):
            if h in want:
                f.write(f"| {h} | {units[i]} | {vals[i]} |\n")
        if len(srows) > 3:
            h2 = srows[1]
            idx = {h: i for i, h in enumerate(h2)}
            data = srows[2:]
            tot = sum(int(r[idx["# Samples"]] or 0) for r in data) or 1
            stalls = [h for h in h2 if h.startswith("stall_") and "Not Issued" not in h]
            agg = sorted(((sum(int(r[idx[s]] or 0) for r in data), s) for s in stalls), reverse=True)[:6]
            f.write("\n## Warp stall samples (source page)\n\n| stall | share |\n|---|---|\n")
            for v, s in agg:
                f.write(f"| {s} | {100 * v / tot:.1f}% |\n")
            f.write("\n## Hottest SASS lines\n\n| samples | share | instruction |\n|---|---|---|\n")
            for r in sorted(data, key=lambda r: -int(r[idx["# Samples"]] or 0))[:10]:
                f.write(f"| {r[idx['# Samples']]} | {100 * int(r[idx['# Samples']]) / tot:.1f}% | `{r[1].strip()}` |\n")
    with open(out_md.replace(".md", "_raw.csv"), "w") as f:
        f.write(raw)


def jsonl_table(path, out_md, title, cols):
    if not os.path.exists(path):
        return
    rows = [json.loads(l) for l in open(path) if l.startswith("{")]
    with open(out_md, "w") as f:
        f.write(f"# {title}\n\n| " + " | ".join(cols) + " |\n|" + "---|" * len(cols) + "\n")
        for r in rows:
            f.write("| " + " | ".join(str(r.get(c, "")) for c in cols) + " |\n")


def fence_probe():
    path = os.path.join(GP, "probe_fence_2gpu.jsonl")
    if not os.path.exists(path):
        return
    rows = [json.loads(l) for l in open(path) if l.startswith("{")]
    cols = ["ctas", "burst_kb_per_cta", "burst_to", "flag_to", "burst_us", "fence_gpu_us", "fence_sys_us", "flag_st_us",
            "flag_ack_fence_sys_us", "st_release_sys_us"]
    with open(os.path.join(OUT, "fence_probe.md"), "w") as f:
        f.write("""# Fence / flag latency probe (2 x B200, benchmarks/probe_fence.py, [median, max] over CTAs, microseconds)

Every CTA writes a burst (to local symmetric memory, to the peer's, or to ordinary local memory), then times a
gpu-scope fence, a system-scope fence, a relaxed flag store, a second system-scope fence (= the flag's ack) and a
st.release.sys.  Findings that shaped the kernel (DESIGN.md 3.4):

* idle: fence.gpu 0.25-0.5 us, fence.sys 1.5-2 us, st.release.sys ~1.75 us; with 32-96 CTAs doing it at once 3-4.5 us;
* after peer bursts fence.gpu grows to 3.5-6 us (it waits for the SM's own NVLink stores to be acknowledged) and fence.sys
  to 7-12 us;
* 96 CTAs x 512 KB of 16-byte peer stores take 61-70 us = 700-780 GB/s: plain LSU stores saturate one NVLink direction;
* inside the streaming allreduce kernel the same fence.sys was measured at up to 56 us (in-kernel trace, 64 MB, 2 and
  8 GPUs): a system-scope fence waits for the NVLink traffic of the whole GPU, not of the issuing SM.  Hence: no
  system-scope fence in data CTAs, one per batch in the publisher warp, and no host-memory polling in the kernel.

| """ + " | ".join(cols) + " |\n|" + "---|" * len(cols) + "\n")
        for r in rows:
            f.write("| " + " | ".join(str(r.get(c, "")) for c in cols) + " |\n")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    sass()
    rep = os.path.join(GP, "prof_twoshot_emu8.ncu-rep")
    if os.path.exists(rep):
        ncu_summary(rep, os.path.join(OUT, "ncu_twoshot_emu8.md"),
                    "ncu: fused two-shot kernel, 8 emulated workers x 16 MB fp32 on one B200 (cooperative launch)")
    if os.path.exists(os.path.join(GP, "launches.csv")):
        open(os.path.join(OUT, "launches_bench_n1.csv"), "w").write(open(os.path.join(GP, "launches.csv")).read())
    cols = ["bytes", "chunk_bytes", "workers", "algo", "zero_copy", "ctas", "ms", "busbw_GBps", "nccl_ms", "nccl_busbw_GBps", "speedup_vs_nccl"]
    for name in sorted(os.listdir(GP)):
        if name.startswith("sweep_") and name.endswith(".jsonl"):
            jsonl_table(os.path.join(GP, name), os.path.join(OUT, name.replace(".jsonl", ".md").replace("--", "_")),
                        f"{name}: fused allreduce vs torch.distributed NCCL all_reduce (device-timed, max over ranks)", cols)
    fence_probe()
    print("profiles written to", OUT)
