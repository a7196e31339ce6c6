"""SASS evidence for the dominant kernels (no GPU needed: cuobjdump on the built objects).

    python tools/sass_excerpt.py > profiles/r02_sass_excerpts.md

Per kernel: mnemonic counts that prove the design (UBLKCP = cp.async.bulk / TMA bulk copies, UBLKPF = bulk prefetch into L2, SYNCS = mbarrier,
LDS = shared-memory visits, LDC = kernel-parameter (constant-bank) tree levels, LDG.E.256 = two-level block
gathers) and the first lines of the unrolled traversal loop."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "tahoe_b200", "build")
KERNELS = [
    ("config 3: k_smem_forest<12, 5, ids=0, staged=0, top_levels=2, SINGLE tiles (one TMA buffer per warp, next tile prefetched into L2)>", "kernels_smem_depth_12.o", "_ZN5tahoe13k_smem_forestILi12ELi5ELb0ELb0ELi2ELb1EEEvNS_10SmemParamsENS_8TopWordsE"),
    ("config 2: k_smem_forest<10, 5, ids=0, staged=0, top_levels=2 (levels 0-1 from kernel parameters), SINGLE tiles>", "kernels_smem_depth_10.o", "_ZN5tahoe13k_smem_forestILi10ELi5ELb0ELb0ELi2ELb1EEEvNS_10SmemParamsENS_8TopWordsE"),
    ("config 4: k_smem_forest_deep<6, 6, ids=0, pieces=8, K=2, pair=1>", "kernels_smem_depth_6.o", "_ZN5tahoe18k_smem_forest_deepILi6ELi6ELb0ELi8ELi2ELb1EEEvNS_10SmemParamsENS_8TopWordsE"),
    ("config 1: k_shared_data<TMA=1, G=2, blocked=1>", "kernels_shared_data.cu.o", "_ZN5tahoe13k_shared_dataILb1ELi2ELb1EEEvNS_16SharedDataParamsE"),
    ("k_quantize<from_tiles=1> (config 3: quantises the transposed float tiles in place)", "kernels_smem.cu.o", "_ZN5tahoe10k_quantizeILb1EEEvNS_11QuantParamsE"),
]
for title, obj, sym in KERNELS:
    out = subprocess.run(["cuobjdump", "-sass", "-fun", sym, os.path.join(OBJ, obj)], capture_output=True, text=True).stdout
    ins = []
    for line in out.splitlines():
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m:
            ins.append((m.group(1), m.group(2).strip()))
    print(f"## {title}\n\n`{sym}` in `tahoe_b200/build/{obj}` ({len(ins)} SASS instructions, sm_100a)\n")
    if not ins:
        print("(symbol not found in this build)\n"); continue
    cnt = collections.Counter()
    for _, t in ins:
        t = re.sub(r"^@!?U?P\d+\s+", "", t)
        op = t.split()[0]
        key = op if op.startswith(("UBLKCP", "UBLKPF", "SYNCS", "LDG", "STG", "LDC", "BAR", "UTMA")) else op.split(".")[0]
        cnt[key] += 1
    keys = [k for k in cnt if k.startswith(("UBLKCP", "UBLKPF", "SYNCS", "LDS", "STS", "LDG", "STG", "LDC", "BAR", "UTMA", "ISETP", "LOP3", "SEL", "IADD3"))]
    print("| mnemonic | count |\n|---|---|")
    for k in sorted(keys, key=lambda k: -cnt[k]):
        print(f"| `{k}` | {cnt[k]} |")
    # first stretch of the unrolled walk: the longest run that is dense in LDS
    best, best_i = 0, 0
    for i in range(0, max(1, len(ins) - 48), 8):
        n = sum(1 for _, t in ins[i:i + 48] if " LDS" in " " + t or t.startswith("LDS"))
        if n > best:
            best, best_i = n, i
    print(f"\nTraversal loop (48 instructions from {ins[best_i][0]}; {best} of them `LDS`):\n\n```")
    for a, t in ins[best_i:best_i + 48]:
        print(f"/*{a}*/ {t} ;")
    print("```\n")
