"""Turns the scratch captures in gpurun_out/ into the committed summaries in profiles/.

  python profiles/summarize.py            # reads gpurun_out/*.ncu-rep, gpurun_out/launches*.csv

* <name>.ncu.md      key metrics of every kernel instance in an `ncu --set full` report
* <name>.launches.md per-kernel totals of an `ncu --metrics gpu__time_duration.sum,...` launch list
* sass_evidence.txt  counts of the SASS mnemonics that prove tcgen05 / TMA / multimem code paths
"""
import collections
import csv
import glob
import io
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
SCRATCH = os.path.join(ROOT, "gpurun_out")

METRICS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram rd"),
    ("dram__bytes_write.sum", "dram wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %peak"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor inst"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps act %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__shared_mem_per_block_dynamic", "dyn smem"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
]


def short(name: str) -> str:
    name = name.replace("void ", "").replace("<unnamed>::", "").replace("(anonymous namespace)::", "")
    return re.sub(r"\(.*", "", name)[:70]


def ncu_raw(rep: str):
    r = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True)
    rows = list(csv.reader(io.StringIO(r.stdout)))
    if len(rows) < 3:
        return None, None, []
    return rows[0], rows[1], rows[2:]


def summarize_rep(rep: str, title: str) -> None:
    hdr, units, rows = ncu_raw(rep)
    if hdr is None:
        return
    name = os.path.splitext(os.path.basename(rep))[0]
    cols = [(hdr.index(m), lab, units[hdr.index(m)]) for m, lab in METRICS if m in hdr]
    ki = hdr.index("Kernel Name")
    with open(os.path.join(OUT, name + ".ncu.md"), "w") as f:
        f.write(f"# {title}\n\nSource: `ncu --set full --clock-control none --import-source on` on a B200 "
                f"(report `{os.path.basename(rep)}`, kept in scratch).  One row per profiled kernel instance.\n\n")
        f.write("| kernel | " + " | ".join(f"{lab} [{u}]" if u else lab for _, lab, u in cols) + " |\n")
        f.write("|---|" + "---|" * len(cols) + "\n")
        for row in rows:
            f.write("| `" + short(row[ki]) + "` | " + " | ".join(row[i] for i, _, _ in cols) + " |\n")


def summarize_launches(path: str, title: str) -> None:
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        a = agg[short(row["Kernel Name"])]
        m = row["Metric Name"]
        if m == "gpu__time_duration.sum":
            a[0] += 1
            a[1] += v
        elif m == "dram__bytes_read.sum":
            a[2] += v
        elif m == "dram__bytes_write.sum":
            a[3] += v
    tot = sum(a[1] for a in agg.values()) or 1.0
    name = os.path.splitext(os.path.basename(path))[0]
    ours = ("bn_", "maxpool", "sgd_momentum", "softmax_xent", "preprocess", "emb_", "linear_tcgen05", "colsum", "wd_head", "conv1x1_tcgen05", "stem_wgrad", "wgrad3x3", "wgrad_reduce", "wgrad_mn", "conv_tcgen05", "s2d_", "pool_", "allreduce_", "alltoall_", "broadcast_kernel", "barrier_kernel",
            "allreduce", "alltoall", "broadcast", "barrier_kernel")
    with open(os.path.join(OUT, name + ".launches.md"), "w") as f:
        f.write(f"# {title}\n\nEvery kernel of one training step, serialised under ncu (cold caches: compare shares, "
                f"not absolutes).  `*` = hand-written kernel of this repository.\n\n")
        f.write(f"Total: {tot / 1e6:.3f} ms over {sum(a[0] for a in agg.values())} launches; hand-written kernels: "
                f"{sum(a[1] for k, a in agg.items() if k.startswith(ours)) / tot * 100:.1f}% of the time, "
                f"{sum(a[0] for k, a in agg.items() if k.startswith(ours))} launches.\n\n")
        f.write("| share | launches | time [us] | dram read [MB] | dram write [MB] | kernel |\n|---|---|---|---|---|---|\n")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
            star = "*" if k.startswith(ours) else ""
            f.write(f"| {a[1] / tot * 100:.1f}% | {a[0]} | {a[1] / 1e3:.1f} | {a[2] / 1e6:.1f} | {a[3] / 1e6:.1f} | `{k}`{star} |\n")


def sass_evidence() -> None:
    lib = os.path.join(ROOT, "paddle_operator_b200", "lib", "libpdjob_b200.so")
    if not os.path.exists(lib):
        return
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    cur, per = None, collections.defaultdict(collections.Counter)
    pat = re.compile(r"\b(UTCHMMA|UTCQMMA|UTMALDG\.2D|UTMASTG\.2D|UBLKCP\S*|LDTM\S*|STTM\S*|UTCBAR\S*|UTCATOMSWS\S*|SYNCS\.\S+|"
                     r"HMMA\S*|RED\.E\S*|REDG\S*|ATOMG\S*|LDG\.E\.128\S*|MULTIMEM\S*|ST\.E\.\S*SYS\S*|LD\.E\.\S*SYS\S*)")
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        if cur:
            for mm in pat.findall(line):
                per[cur][mm.split(" ")[0]] += 1
    ptx = subprocess.run(["cuobjdump", "-ptx", lib], capture_output=True, text=True).stdout
    with open(os.path.join(OUT, "sass_evidence.txt"), "w") as f:
        f.write("SASS mnemonic counts per kernel of libpdjob_b200.so (cuobjdump -sass, sm_100a).\n"
                "UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA tensor load/store, UTCBAR = tcgen05.commit,\n"
                "SYNCS = mbarrier, RED/ATOM = peer reductions; multimem.* appears in SASS as *.MULTIMEM / MMLS forms.\n\n")
        for fn, cnt in sorted(per.items()):
            if not cnt:
                continue
            dem = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip()
            f.write(f"{short(dem)}\n    " + "  ".join(f"{k}={v}" for k, v in sorted(cnt.items())) + "\n")
        keys = ("tcgen05.mma", "tcgen05.ld", "tcgen05.alloc", "tcgen05.commit", "cp.async.bulk.tensor", "mbarrier.",
                "multimem.ld_reduce", "multimem.st", "red.relaxed.sys", "st.release.sys", "ld.acquire.sys")
        if any(ptx.count(k) for k in keys):
            f.write("\nPTX instruction families present (cuobjdump -ptx | grep -c):\n")
            for key in keys:
                f.write(f"  {key:24s} {ptx.count(key)}\n")
        else:
            # built with -gencode arch=compute_100a,code=sm_100a: SASS only, no PTX is embedded
            srcs = sorted(glob.glob(os.path.join(ROOT, "paddle_operator_b200", "csrc", "*.cu*")))
            text = "".join(open(p).read() for p in srcs)
            f.write("\nThe library embeds sm_100a SASS only (no PTX).  Inline-PTX instruction families in csrc/ "
                    "(occurrences in the sources):\n")
            for key in keys:
                f.write(f"  {key:24s} {text.count(key)}\n")


def main() -> int:
    reps = {"prof_bn": "Fused BatchNorm kernels (ResNet-50 step, batch 256)",
            "prof_gemm": "tcgen05 + TMA fused linear kernel (profiles/gemm_bench.py shapes)",
            "prof_emb": "Sharded-embedding kernels and bias colsum (Wide&Deep step, batch 32768, 16.7 M rows)",
            "prof_conv": "tcgen05 convolution kernels (1x1 + BN statistics, dgrad + skip add, stem forward / wgrad) and "
                         "TMA-streamed BN reductions on ResNet-50 layer shapes, batch 256 (profiles/run_conv_kernels.py)",
            "prof_wdhead": "Wide&Deep fused head: output layer + wide part + BCE loss (forward) and the masked gradient of the "
                           "top layer (backward), batch 32768 (profiles/run_wd_step.py)",
            "prof_collectives": "Fused all-reduce / all-to-all kernel bodies, 8 ranks emulated on one GPU "
                                "(cooperative launch; bandwidths here are HBM-local, NVLink numbers are in r1/sweep_n8.json)",
            "prof_wgrad3x3": "tcgen05 3x3 weight-gradient kernel (halo tile + row-shifted descriptors) on the four ResNet-50 "
                             "stages, batch 256 (profiles/run_wgrad3x3.py); round 2",
            "prof_collectives2": "Round 2: fused collectives incl. K1b (reduce-scatter + SGD-momentum + master / parameter "
                                 "broadcast, allreduce_sgd_kernel) and the unrolled all-to-all, 8 ranks emulated on one GPU "
                                 "(HBM-local; ncu may not wrap a multi-rank run, NVLink numbers: r2/sweep_n*.json)",
            "prof_wd2": "Round 2: Wide&Deep step kernels at full size (batch 32768, 16.7 M rows, one GPU): TMA-fed lookup "
                        "writing the padded tower input, push, apply, tcgen05 tower layers (profiles/run_wd_step.py)",
            "prof_gemm2": "Fused linear 8192 x 4096 x 4096 (bias + ReLU): one CTA per tile, then CTA pairs (tcgen05 cta_group::2, "
                          "M = 256 tiles), twice each (profiles/run_gemm_pair.py); round 2",
            "prof_wd4": "Round 2, final: every hand-written kernel of the Wide&Deep step at full size (batch 32768, 16.7 M rows, one "
                        "GPU): read-view lookup, batched push, 512-row-window apply, tower GEMMs (weights as stored, mask tile "
                        "by TMA), MN-major weight gradients, weighted column sums, head (profiles/run_wd_step.py)",
            "prof_wd3": "Round 2, second half: Wide&Deep step kernels at full size (batch 32768, 16.7 M rows, one GPU) with one "
                        "record per table row, the vectorised mask read, the MN-major tower weight gradient and the "
                        "weighted column-sum kernel (profiles/run_wd_step.py)"}
    for n, t in reps.items():
        p = os.path.join(SCRATCH, n + ".ncu-rep")
        if os.path.exists(p):
            summarize_rep(p, t)
    lists = {"launches3": "ResNet-50 step launch list (batch 256, one B200) - before the 1x1-convolution GEMM",
             "launches4": "ResNet-50 step launch list (batch 256, one B200) - first version of the 1x1-convolution GEMM",
             "launches5": "ResNet-50 step launch list (batch 256, one B200) - before the TMA-streamed BN reductions",
             "launches6": "ResNet-50 step launch list (batch 256, one B200)",
             "launches_wd2": "Wide&Deep step launch list (batch 32768, one B200) - 256-thread embedding blocks",
             "launches_wd3": "Wide&Deep step launch list (batch 32768, one B200) - before the fused head / loss kernel",
             "launches_wd4": "Wide&Deep step launch list (batch 32768, one B200)",
             "launches7": "ResNet-50 step launch list (batch 256, one B200) - round 2: tcgen05 3x3 weight gradients, "
                          "warp-uniform MMA issue",
             "launches_wd5": "Wide&Deep step launch list (batch 32768, one B200) - round 2: lookup writes the padded tower "
                             "input in place, zero-copy gradient push",
             "launches8": "ResNet-50 step launch list (batch 256, one B200) - round 2, final: 1x1 / linear weight gradients "
                          "on the MN-major tcgen05 kernel",
             "launches_wd6": "Wide&Deep step launch list (batch 32768, one B200) - round 2, final: MN-major tcgen05 weight "
                             "gradient for the tower, one record per table row",
             "launches9": "ResNet-50 step launch list (batch 256, one B200) - round 2, final: + own backward for 1x1 convolutions "
                          "without a skip gradient, gradient sinks (no gradient copies)",
             "launches_wd8": "Wide&Deep step launch list (batch 32768, one B200) - round 2, final: bf16 read view in the table "
                             "records, batched lookup / push / apply, weights read as stored by the dgrad GEMMs, mask tile by "
                             "TMA, gradient sinks, one column-sum kernel (side-stream weight gradients do not show: ncu "
                             "serialises)",
             "launches_wd7": "Wide&Deep step launch list (batch 32768, one B200) - round 2, final: vectorised mask read in "
                             "the masked dgrad, gradient sinks (no gradient copies), one weighted column-sum kernel for "
                             "the bias and head gradients"}
    for n, t in lists.items():
        p = os.path.join(SCRATCH, n + ".csv")
        if os.path.exists(p):
            summarize_launches(p, t)
    sass_evidence()
    return 0


if __name__ == "__main__":
    sys.exit(main())
