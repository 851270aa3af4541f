"""Writes trimmed SASS listings of the hand-written kernels to profiles/sass/<kernel>.sass.

    python profiles/make_sass.py            # needs cuobjdump (CUDA toolkit) and the built library

One file per kernel: every instruction of the classes that prove what the kernel is made of (tensor-core MMA,
tensor-memory loads, TMA / bulk copies, mbarrier traffic, peer / multicast / system-scope memory operations,
reductions), with its address, in program order; plus the instruction count of the whole kernel.  The PTX names
never appear in SASS: tcgen05.mma = UTCHMMA, tcgen05.ld = LDTM, cp.async.bulk.tensor = UTMALDG / UTMASTG,
cp.async.bulk = UBLKCP, mbarrier = SYNCS, multimem.ld_reduce = LDGMC...(.ADD), multimem.st = a plain STG...SYS to
the multicast address, red.global.add.v4.f32 = REDG / RED.
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "paddle_operator_b200", "lib", "libpdjob_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEEP = re.compile(r"\b(UTC[A-Z]*MMA|UTCBAR|UTCCP|LDTM|STTM|UTMALDG|UTMASTG|UTMAPF|UBLKCP|UBLKRED|SYNCS|LDGMC|MULTIMEM|"
                  r"RED[G]?|ATOM[GS]?|LDG\.E[^ ]*\.(SYS|STRONG)[^ ]*|STG\.E[^ ]*\.(SYS|STRONG)[^ ]*|LD\.E[^ ]*SYS|ST\.E[^ ]*SYS|"
                  r"ELECT|MEMBAR|FENCE|HMMA|CCTL|ERRBAR|NANOSLEEP)\b")
WANT = ["allreduce_nvls", "allreduce_twoshot", "allreduce_oneshot", "allreduce_sgd", "alltoall", "broadcast_kernel",
        "barrier_kernel", "emb_lookup_tma", "emb_lookup_kernel", "emb_lookup_batched", "emb_lookup_shadow", "emb_push", "emb_apply",
        "wd_head", "colsum_w", "linear_tcgen05", "wgrad_mn",
        "conv1x1_tcgen05", "stem_wgrad", "wgrad3x3", "wgrad_reduce", "bn_stats_tma", "bn_bwd_reduce_tma"]


def demangle(names):
    try:
        out = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
        return out if len(out) == len(names) else names
    except OSError:
        return names


def main():
    if not os.path.exists(LIB):
        sys.exit(f"{LIB} missing: build first")
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    os.makedirs(OUT, exist_ok=True)
    blocks = re.split(r"\n\s*Function : ", txt)[1:]
    written = {}
    for b in blocks:
        name = b.split("\n", 1)[0].strip()
        key = next((w for w in WANT if w in name), None)
        if key is None:
            continue
        lines = [l for l in b.splitlines() if re.match(r"\s*/\*[0-9a-f]{4,}\*/", l)]
        kept = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip() for l in lines if KEEP.search(l)]
        pretty = demangle([name])[0]
        if key in ("allreduce_twoshot", "allreduce_oneshot", "allreduce_sgd") and "<8" not in pretty:
            continue  # one instance per rank count: the 8-rank one is listed
        tag = key if key not in written else f"{key}_{written[key]}"
        written[key] = written.get(key, 0) + 1
        with open(os.path.join(OUT, f"{tag}.sass"), "w") as f:
            f.write(f"// {pretty}\n// {len(lines)} SASS instructions in the kernel; {len(kept)} shown (cuobjdump -sass, sm_100a)\n")
            shown = kept if len(kept) <= 220 else kept[:200] + [f"        ... {len(kept) - 200} more of the same classes ..."]
            f.write("\n".join(shown) + "\n")
    print(f"wrote {sum(written.values())} listings to {OUT}")


if __name__ == "__main__":
    main()
