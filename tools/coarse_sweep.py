#!/usr/bin/env python
"""BASELINE.json configs[4]: coarse-correlation sweep N0 = N1 from 1k to 64k tokens, C = 256.

For every N: the fused statistics kernel (tcgen05 GEMM + dual-softmax sums, the matrix never
materialised) timed alone with CUDA events -> TFLOP/s and fraction of the measured bf16 peak;
the whole fm_coarse_match call; and, where the matrices fit, the reference's own formulation in
eager PyTorch ON THE SAME GPU (einsum -> / temperature -> softmax x 2 -> product -> threshold /
mutual-NN masks, coarse_matching_new.py:64-110) in fp32 and under bf16 inputs, as the
like-for-like GPU baseline (library kernels, not ours).  One JSON line per N on stdout.

    python tools/coarse_sweep.py [--max-n 65536] [--ref-max-n 32768]
"""
import argparse
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from featurematching_b200 import _lib, modules, synth  # noqa: E402


def peaks():
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return p.get("bf16_tflops_sustained", p["bf16_tflops"]), "measured (sustained bf16)"
    return 1400.0, "fallback"


def grid_for(n):
    h = 1
    while h * h < n:
        h *= 2
    return (h, n // h) if h * (n // h) == n else (1, n)


def time_events(fn, iters):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def reference_eager(f0, f1, thr=0.2, temperature=0.1):
    """coarse_matching_new.py:64-110 verbatim in eager torch (border mask left out: O(N) work)."""
    C = f0.shape[-1]
    a, b = f0 / C ** .5, f1 / C ** .5
    sim = torch.einsum("nlc,nsc->nls", a, b) / temperature
    conf = torch.softmax(sim, 1) * torch.softmax(sim, 2)
    mask = conf > thr
    mask = mask * (conf == conf.max(dim=2, keepdim=True)[0]) * (conf == conf.max(dim=1, keepdim=True)[0])
    return torch.where(mask)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-n", type=int, default=65536)
    ap.add_argument("--ref-max-n", type=int, default=32768)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    lib = _lib.load()
    peak, peak_src = peaks()
    n = 1024
    while n <= args.max_n:
        # batch so that every point does comparable work (>= one wave of work units)
        B = max(1, min(32, (8192 // n) ** 2 * 2)) if n < 8192 else 1
        hw = grid_for(n)
        sigma = 1.0 if n <= 8192 else (1.15 if n <= 32768 else 1.25)
        f0, f1, _ = synth.coarse_features(B, n, n, 256, 0.1, "planted", 5, device=dev, sigma=sigma)
        cm = modules.CoarseMatching({"thr": 0.2, "border_rm": 2, "dsmax_temperature": 0.1}).eval()
        data = {"hw0_i": (hw[0] * 8, hw[1] * 8), "hw1_i": (hw[0] * 8, hw[1] * 8), "hw0_c": hw, "hw1_c": hw}
        cm(f0, f1, data)
        M = int(data["b_ids"].numel())
        ws = cm._ws.buf
        ms = ctypes.c_float(0.0)
        iters = max(3, min(50, int(2e12 / (2.0 * n * n * 256 * B))))
        rc = lib.fm_profile_coarse_stats(_lib.ptr(f0), _lib.ptr(f1), B, n, n, 256, 0.1, 0.2, _lib.ptr(ws), ws.numel(),
                                         cm.cand_per_row, iters, ctypes.byref(ms), _lib.stream_ptr(dev))
        _lib.check(rc, "fm_profile_coarse_stats")
        flops = 2.0 * n * n * 256 * B
        ms_call = time_events(lambda: cm.match_raw(f0, f1, data["hw0_i"], hw, hw), iters)
        line = {"N": n, "B": B, "C": 256, "matches": M, "stats_kernel_ms": ms.value,
                "stats_kernel_tflops": flops / (ms.value * 1e-3) / 1e12,
                "frac_of_peak": flops / (ms.value * 1e-3) / 1e12 / peak, "peak_tflops": peak, "peak_source": peak_src,
                "fm_coarse_match_ms": ms_call, "pairs_per_s_coarse": B / (ms_call * 1e-3)}
        if n <= args.ref_max_n:
            try:
                g0, g1 = f0[:1].float(), f1[:1].float()
                ms_ref = time_events(lambda: reference_eager(g0, g1), 3)
                ref_idx = reference_eager(g0, g1)
                line["reference_eager_fp32_ms_per_pair"] = ms_ref
                line["reference_eager_fp32_tflops"] = 2.0 * n * n * 256 / (ms_ref * 1e-3) / 1e12
                line["speedup_vs_reference_eager_fp32"] = ms_ref / (ms_call / B)
                sel = data["b_ids"] == 0
                line["match_set_equal_reference_eager"] = bool(
                    ref_idx[1].numel() >= int(sel.sum()) and
                    set(zip(data["i_ids"][sel].tolist(), data["j_ids"][sel].tolist())) <=
                    set(zip(ref_idx[1].tolist(), ref_idx[2].tolist())))
                h0, h1 = f0[:1], f1[:1]
                line["reference_eager_bf16_ms_per_pair"] = time_events(lambda: reference_eager(h0, h1), 3)
                del g0, g1, ref_idx
            except torch.OutOfMemoryError:
                line["reference_eager_fp32_ms_per_pair"] = "out of memory"
            torch.cuda.empty_cache()
        print(json.dumps(line), flush=True)
        del f0, f1, cm, ws
        torch.cuda.empty_cache()
        n *= 2


if __name__ == "__main__":
    main()
