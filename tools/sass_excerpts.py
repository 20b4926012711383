"""SASS evidence per fused path: mnemonic counts for every kernel of the library plus, for the fused
compute+collective kernels, an EXCERPT of the instruction stream around the instructions that prove the fusion
(tensor-core MMAs and peer / multicast memory traffic in the same kernel).

    cuobjdump -sass distributed_training_compare_jax_b200/lib/libdtc_b200.so > /tmp/lib.sass
    python tools/sass_excerpts.py /tmp/lib.sass profiles/r2/sass

writes  <out>/counts.md  and one  <out>/<kernel>.md  per kernel listed in EXCERPTS."""
import os
import re
import subprocess
import sys
from collections import Counter

KEY = re.compile(r"\b(UTC[A-Z]*MMA[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UTMAREDG[.\w]*|UBLKCP[.\w]*|UTCBAR[.\w]*|"
                 r"UTCATOMSWS[.\w]*|UCGABAR_\w+|LDGMC[.\w]*|STGMC[.\w]*|REDG[.\w]*|RED\.[.\w]*|MUFU\.\w+|HMMA[.\w]*|"
                 r"ST\.E[.\w]*SYS[.\w]*|STG\.E[.\w]*SYS[.\w]*|LDG\.E[.\w]*SYS[.\w]*|LD\.E[.\w]*SYS[.\w]*|MEMBAR[.\w]*|"
                 r"ATOMG[.\w]*|SYNCS[.\w]*)")

# kernel-name substring -> (title, mnemonic regexes to centre an excerpt on, lines of context)
EXCERPTS = {
    "gemm_persistent_kernel<false, true, 128, true, true, false>": (
        "GEMM -> reduce-scatter epilogue (EPI_SCATTER_BF16): tcgen05 tiles stored into the owner rank's slot over NVLink",
        [r"UTCHMMA", r"LDTM", r"STG\.E\.128"], 6),
    "dp_reduce_scatter_ranges_kernel<true>": (
        "DP gradient reduce-scatter: multimem.ld_reduce (in-switch fp32 sum) + squared-norm partials",
        [r"LDGMC"], 8),
    "dp_adamw_allgather_ranges_kernel<true>": (
        "DP sharded AdamW + all-gather: the bf16 result is stored through the multicast address (multimem.st)",
        [r"ST(G)?\.E\.128[.\w]*SYS", r"MUFU"], 6),
    "sp_ln_fwd_kernel": (
        "Sequence-parallel LayerNorm: fp32 sum of the partial slots + bias + residual, output pushed to every rank",
        [r"ST(G)?\.E[.\w]*SYS", r"MUFU\.RSQ"], 6),
    "sp_ln_bwd_kernel": (
        "Sequence-parallel LayerNorm backward: partial-slot reduction, gradient rows pushed to every rank",
        [r"ST(G)?\.E[.\w]*SYS"], 6),
    "gemm_2cta_kernel<false, true, 256, 4, 0>": (
        "CTA-pair GEMM (cta_group::2, 256 x 256 tile, store epilogue compiled in): 2-SM TMA loads, UMMA, slab epilogue with TMA stores",
        [r"UTCHMMA\.2CTA", r"UTMALDG\.2D\.2CTA", r"UTMASTG", r"LDTM"], 4),
    "gemm_2cta_kernel<true, true, 256, 4, 4>": (
        "CTA-pair GEMM, fp32 reduce-add epilogue (weight gradients / split-K): accumulator slabs leave through TMA reduce-add",
        [r"UTMAREDG", r"LDTM", r"STS\.128"], 4),
    "gemm_grouped_wgrad_2cta_kernel<256>": (
        "Grouped weight gradients of a layer on CTA pairs: tile table in the parameter block, TMA reduce-add epilogue",
        [r"UTCHMMA\.2CTA", r"UTMAREDG"], 4),
    "layernorm_bwd_wide_kernel<2>": (
        "LayerNorm backward, 16 warps per CTA: per-warp column partials in shared memory, one red.global.add.v4 per 4 columns",
        [r"RED", r"STS\.128"], 4),
    "flash_attn_fwd_pipe_kernel<32>": (
        "Flash attention forward, pipelined: issue warp (TMA + tcgen05.mma) and softmax warps (LDTM, MUFU.EX2)",
        [r"UTCHMMA", r"MUFU\.EX2", r"LDTM"], 3),
    "flash_attn_bwd_persist_kernel<32, 4>": (
        "Flash attention backward, persistent + pipelined: S / dP, dV / dK / dQ products on tcgen05, dQ via fp32 red.add",
        [r"UTCHMMA", r"REDG", r"LDTM"], 3),
    "symm_barrier_kernel": ("Cross-rank flag barrier (signal pads in symmetric memory)", [r"SYS"], 6),
}


def demangle(names):
    try:
        out = subprocess.run(["cu++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout
        return out.splitlines()
    except Exception:
        return list(names)


def parse(path):
    funcs, cur = {}, None
    for line in open(path, errors="ignore"):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        if cur and re.match(r"\s+/\*[0-9a-f]{4,6}\*/", line):
            addr = re.match(r"\s+/\*([0-9a-f]{4,6})\*/", line).group(1)
            ins = re.sub(r"/\*.*?\*/", "", line)
            funcs[cur].append((addr, " ".join(ins.split()).rstrip(";")))
    return funcs


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(bool\)([01])", lambda m: "true" if m.group(1) == "1" else "false", name)
    name = re.sub(r"\((?:int|unsigned int|long)\)(-?\d+)", r"\1", name)
    name = re.sub(r"\(.*$", "", name)
    return name.replace("dtc::", "")


def main():
    src, out = sys.argv[1], sys.argv[2]
    os.makedirs(out, exist_ok=True)
    funcs = parse(src)
    pretty = dict(zip(funcs, [short(n) for n in demangle(list(funcs))]))
    with open(os.path.join(out, "counts.md"), "w") as f:
        f.write("# SASS mnemonic counts per kernel (`cuobjdump -sass lib/libdtc_b200.so`, sm_100a)\n\n"
                "`UTCHMMA` = tcgen05.mma (`.2CTA` = cta_group::2), `LDTM` = tcgen05.ld, `UTMALDG` / `UTMASTG` = TMA load / store, "
                "`UTCBAR` = tcgen05.commit, `LDGMC…ADD` = multimem.ld_reduce, `REDG…` = red.global (fp32 split-K / dQ, or "
                "multimem.red when `.SYS`), `…STRONG.SYS` stores = peer / multicast stores.\n\n"
                "| kernel | instructions | key mnemonics |\n|---|---:|---|\n")
        for k in sorted(funcs, key=lambda k: pretty[k]):
            c = Counter()
            for _, ins in funcs[k]:
                for m in KEY.findall(ins):
                    c[m] += 1
            if not c:
                continue
            f.write(f"| `{pretty[k]}` | {len(funcs[k])} | " + ", ".join(f"{m}×{n}" for m, n in sorted(c.items())) + " |\n")
        hm = sum(1 for k in funcs for _, ins in funcs[k] if re.search(r"\bHMMA", ins))
        f.write(f"\nStand-alone `HMMA` (legacy mma.sync) instructions in the whole library: **{hm}**.\n")
    for sub, (title, pats, ctx) in EXCERPTS.items():
        hit = [k for k in funcs if sub in pretty[k]]
        if not hit:
            print(f"[sass_excerpts] no kernel matches `{sub}`")
            continue
        k = hit[0]
        ins = funcs[k]
        keep = set()
        for p in pats:
            idx = [i for i, (_, t) in enumerate(ins) if re.search(p, t)]
            for i in idx[:3]:  # first three occurrences of each pattern
                keep.update(range(max(0, i - ctx), min(len(ins), i + ctx + 1)))
        fname = re.sub(r"[^A-Za-z0-9_]+", "_", sub).strip("_") + ".md"
        with open(os.path.join(out, fname), "w") as f:
            f.write(f"# {title}\n\nKernel `{pretty[k]}` — {len(ins)} SASS instructions; excerpt around "
                    + ", ".join(f"`{p}`" for p in pats) + " (addresses in hex, `...` = lines skipped).\n\n```\n")
            last = -2
            for i in sorted(keep):
                if i != last + 1:
                    f.write("        ...\n")
                f.write(f"  /*{ins[i][0]}*/  {ins[i][1]}\n")
                last = i
            f.write("```\n")
    print(f"[sass_excerpts] wrote {out}/")


if __name__ == "__main__":
    main()
