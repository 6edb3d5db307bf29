"""K1 ASCII kernel alone (tnf_count_ascii2_kernel): resident input, CUDA-event time per launch, HBM roofline fraction;
single-line records (every unit on the register-classified fast path) and the same records wrapped at 60 columns
(about half the units hold a newline and take the byte-by-byte path).  Counts are checked against the packed kernel
and, on a prefix, against the oracle."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
from rosella_b200 import gpu, synth  # noqa: E402


def timed(ctx, reps=8, warm=3):
    ms = []
    for i in range(warm + reps):
        ctx.tnf_count_resident()
        if i >= warm:
            ms.append(ctx.timings()["tnf_count_ms"])
    return float(np.mean(ms)), float(np.min(ms))


def main():
    gbp = float(sys.argv[1]) if len(sys.argv) > 1 else 4.0
    out = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/tnf_ascii_kernel.json"
    peaks = json.load(open("MEASURED_PEAKS.json")) if os.path.exists("MEASURED_PEAKS.json") else {}
    hbm = float(peaks.get("hbm_gbs", peaks.get("hbm_gbps", 6534.8)) or 6534.8)
    ctx = gpu.Context(0)
    seq, offs = synth.contigs(int(gbp * 1e9), threads=16)
    n, bases = len(offs) - 1, int(offs[-1])
    res = {"bases": bases, "contigs": n, "hbm_peak_gbs": hbm}
    ctx.tnf_load_ascii(seq, offs)
    mean, best = timed(ctx)
    c, nk, _ = ctx.tnf_fetch()
    algo = bases + n * (136 * 4 + 8 + 8)
    res["single_line"] = {"ms": mean, "ms_min": best, "tbp_s": bases / mean / 1e9, "gbs": algo / mean / 1e6,
                          "frac_hbm": algo / mean / 1e6 / hbm}
    codes, valid, woff = gpu.pack_host(seq, offs, threads=16)
    ctx.tnf_load_packed(codes, valid, woff)
    mean_p, best_p = timed(ctx)
    c2, nk2, _ = ctx.tnf_fetch()
    res["packed"] = {"ms": mean_p, "ms_min": best_p, "tbp_s": bases / mean_p / 1e9}
    res["ascii_equals_packed"] = bool(np.array_equal(c, c2) and np.array_equal(nk, nk2))
    k = int(np.searchsorted(offs, 30_000_000))
    want = oracle.tnf_count(seq[: int(offs[k])], offs[: k + 1], fast=True)
    res["prefix_equals_oracle"] = bool(np.array_equal(c[:k], want[0]) and np.array_equal(nk[:k], want[1]))
    # 60-column FASTA lines on a 1 Gbp prefix (fasta_wrap is a Python loop per record)
    k = int(np.searchsorted(offs, min(bases, 1_000_000_000)))
    wrapped, woffs = synth.fasta_wrap(seq[: int(offs[k])], offs[: k + 1])
    ctx.tnf_load_ascii(wrapped, woffs)
    mean_w, best_w = timed(ctx)
    cw, nkw, _ = ctx.tnf_fetch()
    wb = int(offs[k])
    res["wrapped_60"] = {"bases": wb, "bytes": int(woffs[-1]), "ms": mean_w, "tbp_s": wb / mean_w / 1e9,
                         "gbs": int(woffs[-1]) / mean_w / 1e6, "frac_hbm": int(woffs[-1]) / mean_w / 1e6 / hbm,
                         "equals_single_line": bool(np.array_equal(cw, c[:k]) and np.array_equal(nkw, nk[:k]))}
    ctx.close()
    os.makedirs(os.path.dirname(out) or ".", exist_ok=True)
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
