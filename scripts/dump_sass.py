"""Regenerates profiles/sass: one SASS listing per kernel family + an opcode census (no GPU needed)."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "distributed-machine-learning-pipeline_b200", "_native", "libdmlp_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")

PICK = {  # listing name -> regex on the mangled kernel name
    "gemm_tcgen05_tf32_fwd.sass": r"gemm_tcgen05_kernelILi4ELb0ELb1ELb0ELi0E",
    "gemm_tcgen05_bf16_wgrad.sass": r"gemm_tcgen05_kernelILi2ELb1ELb1ELb0ELi0E",
    "gemm_tcgen05_bf16_fwd_cta_pair.sass": r"gemm_tcgen05_kernelILi2ELb0ELb1ELb1ELi0E",
    "gemm_tcgen05_bf16_fwd_cta_pair_tma_store.sass": r"gemm_tcgen05_kernelILi2ELb0ELb1ELb1ELi1E",
    "gemm_tcgen05_bf16_wgrad_cta_pair_reduce_scatter.sass": r"gemm_tcgen05_kernelILi2ELb1ELb1ELb1ELi2E",
    "mlp_fused_step_16_8.sass": r"mlp_fused_step_kernelILi16ELi8ELb0E",
    "mlp_fused_step_16_8_global_batch.sass": r"mlp_fused_step_kernelILi16ELi8ELb1E",
    "input_exchange.sass": r"input_exchange_kernel",
    "allreduce_oneshot_f32_sgd.sass": r"allreduce_oneshot_kernelIfLb1E",
    "allreduce_twoshot_f32_sgd.sass": r"allreduce_twoshot_kernelIfLb1E",
    "allreduce_ring_f32_sgd.sass": r"allreduce_ring_kernelIfLb1E",
    "allreduce_nvls.sass": r"allreduce_nvls_kernel",
    "p2p_copy.sass": r"p2p_copy_kernel",
    "sgd.sass": r"sgd_kernel",
    "softmax_ce.sass": r"softmax_ce_kernel",
}
NOTABLE = ["UTCHMMA.2CTA", "UTCHMMA", "LDGSTS", "UTCQMMA", "LDTM", "UTMALDG", "UTMAPF", "UTMASTG", "UBLKCP", "UBLKRED", "UTCBAR", "UTCATOMSWS", "SYNCS", "UCGABAR",
           "LDGMC", "STG.E.128.MMC", "RED", "ATOM", "LDG", "STG", "LDS", "STS", "MEMBAR", "CCTL", "ERRBAR", "R2UR", "BRA.U.ANY", "ELECT"]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    funcs = collections.OrderedDict()
    cur = None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur is not None:
            funcs[cur].append(line)
    os.makedirs(OUT, exist_ok=True)
    rows = []
    for fname, pat in PICK.items():
        name = next((n for n in funcs if re.search(pat, n)), None)
        if name is None:
            rows.append((fname, 0, "(kernel not found)"))
            continue
        body = [l for l in funcs[name] if re.search(r"/\*[0-9a-f]{4,}\*/", l)]
        # keep the instruction lines only, without the 128-bit encodings (the listings are evidence, not input)
        lean = [re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", l).rstrip() for l in body]
        with open(os.path.join(OUT, fname), "w") as f:
            f.write(f"// {name}\n" + "\n".join(lean) + "\n")
        census = collections.Counter()
        for l in body:
            for op in NOTABLE:
                if re.search(r"\b" + re.escape(op), l):
                    census[op] += 1
        note = ", ".join(f"{k}x{v}" for k, v in census.items() if v)
        rows.append((fname, len(body), note))
    with open(os.path.join(OUT, "SUMMARY.md"), "w") as f:
        f.write("# SASS evidence (`cuobjdump -sass` of libdmlp_b200.so, sm_100a) — regenerate with `python scripts/dump_sass.py`\n\n"
                "Full listings of one instantiation per kernel family are in this directory.\n"
                "Mnemonics that prove the Blackwell paths: `UTCHMMA` = tcgen05.mma (`.2CTA` = cta_group::2), `LDTM` = tcgen05.ld, "
                "`UTMALDG` / `UTMASTG` = TMA tensor load / store, `LDGSTS` = cp.async, `UTMAPF` = TMA L2 prefetch, `UBLKCP` / `UBLKRED` = cp.async.bulk / cp.reduce.async.bulk, "
                "`UTCBAR` = tcgen05.commit, `UTCATOMSWS` = TMEM alloc/dealloc, `SYNCS` = mbarrier, `UCGABAR` = cluster barrier, "
                "`LDGMC...ADD` = NVLS multimem.ld_reduce. `BRA.U.ANY` counts per-lane \"waterfall\" loops around uniform-datapath "
                "instructions (0 in the GEMM / fused kernels since the warp-uniform role dispatch, see ../r1_mma_issue_probe.md).\n\n"
                "| listing | instructions | notable opcodes |\n|---|---|---|\n")
        for fname, n, note in rows:
            f.write(f"| `{fname}` | {n} | {note} |\n")
    print(open(os.path.join(OUT, "SUMMARY.md")).read())


if __name__ == "__main__":
    main()
