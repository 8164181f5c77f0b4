"""profiles/r2_kernel_resources.md: registers / shared memory / spills of every kernel in libdmlp_b200.so (no GPU needed).

Sources: `cuobjdump --dump-resource-usage` of the built objects and the `-Xptxas -v` logs kept by build_native.py."""
import collections
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    rows = []
    for obj in sorted(glob.glob(os.path.join(ROOT, "build", "*.o"))):
        txt = subprocess.run(["cuobjdump", "--dump-resource-usage", obj], capture_output=True, text=True).stdout
        cur = None
        for line in txt.splitlines():
            m = re.search(r"Function (\S+):", line)
            if m:
                cur = m.group(1)
                continue
            m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
            if m and cur:
                rows.append((os.path.basename(obj)[:-2], cur) + tuple(int(v) for v in m.groups()))
                cur = None
    spills = {}
    for log in glob.glob(os.path.join(ROOT, "build", "ptxas_*.log")):
        cur = None
        for line in open(log):
            m = re.search(r"Compiling entry function '(\S+)'", line)
            if m:
                cur = m.group(1)
            m = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", line)
            if m and cur:
                spills[cur] = (int(m.group(1)), int(m.group(2)))
    names = demangle(sorted({r[1] for r in rows}))
    seen = collections.OrderedDict()
    for unit, fn, reg, stack, shared, local in rows:
        pretty = re.sub(r"\(.*", "", names.get(fn, fn)).replace("dmlp::", "").replace("void ", "")
        seen[(unit, pretty)] = (reg, stack, shared, spills.get(fn, (0, 0)))
    out = os.path.join(ROOT, "profiles", "r2_kernel_resources.md")
    with open(out, "w") as f:
        f.write("# Kernel resources (sm_100a, `cuobjdump --dump-resource-usage` + `ptxas -v`) — regenerate with "
                "`python scripts/kernel_resources.py`\n\n"
                "`gemm_tcgen05_kernel<EB, A_MN, B_MN, TWO, E2M>`: EB = operand bytes (4 tf32, 2 bf16), MN-major flags, CTA pairs, "
                "epilogue mode (0 generic / 1 TMA store / 2 reduce-scatter + consumer warps). Static shared memory only; the GEMM and "
                "fused-step kernels add their dynamic rings (up to 227 KiB per CTA).\n\n"
                "| unit | kernel | registers | stack B | static smem B | spill st/ld B |\n|---|---|---|---|---|---|\n")
        for (unit, pretty), (reg, stack, shared, sp) in seen.items():
            f.write(f"| {unit} | `{pretty}` | {reg} | {stack} | {shared} | {sp[0]}/{sp[1]} |\n")
    print(open(out).read()[:3000])


if __name__ == "__main__":
    main()
