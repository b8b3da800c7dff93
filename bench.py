#!/usr/bin/env python
"""Benchmark of the matching head (BASELINE.json metric: matched frame pairs/s at 640x512).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

One step = one pass of the hot path as network/net.py:75-83 runs it (coarse dual-softmax mutual-NN match ->
fine window gather -> fine LocalFeatureTransformer -> fine dual-correlate match -> match-record pack
[+ cross-GPU gather to rank 0 when N > 1]) over one batch of synthetic frame pairs: BASELINE.json configs[1],
32 planted 640x512-shaped pairs per GPU, bf16 features.  The fine transformer is INSIDE the step since round 2
(net.py always runs it when M > 0); --without-fine-transformer restores round 1's step on both arms.
The coarse transformer (the producer of the coarse tokens) is timed beside the step on
the step's tokens and reported under kernels.coarse_transformer (with random-init weights it would scramble the
planted correspondences, so it cannot sit in front of a step that must produce matches).
Beside the step, on rank 0 at N = 1: `cpu_baseline` (the oracle port on the host cores, on the GPU's own pair 0,
whose result also serves as the parity self-check of what was timed), `gpu_eager_baseline` (the same oracle
formulation in eager PyTorch on the same B200, fp32, TF32 off, stage by stage) and `e2e_from_images` (the whole
matcher from host images, backbone in eager PyTorch -- a measurement, not a kernel claim).
Prints ONE JSON line (rank 0).  `--impl reference` times the CPU restatement of the reference
(oracle/, the reference itself cannot travel to the GPU box) on the host cores instead.
`--config seq10k` is BASELINE configs[3] as one job: a 10 000-frame 320x320 sequence block-sharded over the
ranks with a one-frame halo, each rank streaming its pairs in steps of 64.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "matched frame pairs/sec at 640x512"
UNIT = "pairs/s"
# BASELINE.json configs that fit this bench: pairs per GPU per step, image size, and whether the
# pairs are consecutive frames of one sequence (image 1 of pair k is image 0 of pair k + 1)
CONFIGS = {
    "640x512": {"pairs": 32, "height": 512, "width": 640, "sequence": False, "sigma": 1.0,
                "name": "BASELINE configs[1]"},
    # sigma 1.15: at sigma 1 the planted confidences sit at thr = 0.2 against 20 480-token denominators
    "1280x1024": {"pairs": 4, "height": 1024, "width": 1280, "sequence": False, "sigma": 1.15,
                  "name": "BASELINE configs[2]"},
    "seq320": {"pairs": 64, "height": 320, "width": 320, "sequence": True, "sigma": 1.0,
               "name": "BASELINE configs[3]"},
    # the whole configs[3] job: --frames consecutive frames, sharded over the ranks (pipeline.shard_sequence)
    "seq10k": {"pairs": 64, "height": 320, "width": 320, "sequence": True, "sigma": 1.0,
               "name": "BASELINE configs[3], whole 10k-frame job", "job": True},
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="640x512", choices=sorted(CONFIGS),
                    help="BASELINE.json workload: 640x512 = configs[1] (the metric's), 1280x1024 = configs[2], "
                         "seq320 = configs[3] (consecutive frames of a 320x320 sequence)")
    ap.add_argument("--pairs", type=int, default=None, help="frame pairs per GPU per step (default: the config's)")
    ap.add_argument("--height", type=int, default=None)
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--noise", type=float, default=0.1)
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-coarse-transformer", action="store_true",
                    help="skip the coarse LocalFeatureTransformer stage (timed beside the step, kernels.coarse_transformer)")
    ap.add_argument("--without-fine-transformer", action="store_true",
                    help="leave the fine LocalFeatureTransformer (net.py:79-80) OUT of the step (round 1's step: "
                         "coarse match + gather + fine match), on the GPU and CPU arms alike; it is then timed "
                         "beside the step and reported under kernels.fine_transformer with in_step = false")
    ap.add_argument("--with-fine-transformer", action="store_true", help="(default since round 2; kept for old command lines)")
    ap.add_argument("--frames", type=int, default=10000, help="--config seq10k: frames of the sequence")
    ap.add_argument("--seed-rank", type=int, default=None, help=argparse.SUPPRESS)
    ap.add_argument("--exchange-every", type=int, default=8,
                    help="--config seq10k: steps whose records are exchanged together (one exchange per E steps)")
    ap.add_argument("--gather", default="root", choices=["root", "all"],
                    help="N > 1: match records go to rank 0 only (grouped send/recv) or to every rank (all_gather)")
    ap.add_argument("--no-gpu-eager-baseline", action="store_true")
    ap.add_argument("--no-from-images", action="store_true")
    a = ap.parse_args()
    a.with_fine_transformer = not a.without_fine_transformer
    a.job = bool(CONFIGS[a.config].get("job"))
    preset = CONFIGS[a.config]
    for k in ("pairs", "height", "width"):
        if getattr(a, k) is None:
            setattr(a, k, preset[k])
    a.sequence = preset["sequence"]
    a.sigma = preset["sigma"]
    return a


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"hbm_gbs": p["hbm_gbs"], "tflops_burst": p["bf16_tflops"],
                "tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "tflops_burst": 1590.0, "tflops_sustained": 1400.0, "source": "fallback"}


def shapes(args):
    H, W = args.height, args.width
    hw_i, hw_c, hw_f = (H, W), (H // 8, W // 8), (H // 2, W // 2)
    return hw_i, hw_c, hw_f, hw_c[0] * hw_c[1]


# ------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi sampling DURING the timed region (B200_PROFILING.md's clocks line)."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.tmp = index, None, None

    def start(self):
        try:
            self.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "20"], stdout=self.tmp, stderr=subprocess.DEVNULL)
            time.sleep(0.15)   # let the first samples land before the region starts
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.tmp.flush()
        self.tmp.seek(0)
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.tmp.read().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        self.tmp.close()
        os.unlink(self.tmp.name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(power)}


# ------------------------------------------------------------------------------ CPU arm
FINE_CFG = {"d_model": 64, "nhead": 8, "layer_names": ["self", "cross"], "attention": "linear"}   # demo/net_config.py:40-44
COARSE_CFG = {"d_model": 256, "nhead": 8, "layer_names": ["self", "cross"] * 4, "attention": "linear"}   # demo/net_config.py:20-25
HEAD_CFG = {"fine_window_size": 7, "fine_concat_coarse_feat": True, "coarse": {"d_model": 256}, "fine": FINE_CFG,
            "match_coarse": {"thr": 0.2, "border_rm": 2, "dsmax_temperature": 0.1}}


def fine_transformer_module():
    """The B200 LocalFeatureTransformer with seeded random weights (no checkpoint exists offline);
    the CPU arms read the same weights through its state dict."""
    import torch
    from featurematching_b200 import modules
    torch.manual_seed(77)
    m = modules.LocalFeatureTransformer(FINE_CFG).eval()
    m.reuse_packed_weights = True   # the bench owns these weights and never writes them: pack once (eval-mode serving)
    return m


def coarse_transformer_module():
    """The coarse LocalFeatureTransformer with seeded random-init weights."""
    import torch
    from featurematching_b200 import modules
    torch.manual_seed(78)
    m = modules.LocalFeatureTransformer(COARSE_CFG).eval()
    m.reuse_packed_weights = True
    return m


def oracle_head(args, device="cpu"):
    """run(fc0, fc1, ff0, ff1) -> the oracle's whole head (network/net.py:75-83) on `device`, with the bench's
    seeded hot-path parameters and fine-transformer weights."""
    import torch
    from featurematching_b200 import synth
    from oracle import matcher_oracle as mo, reference_shims
    cfg = reference_shims.default_config()
    hw_i, hw_c, hw_f, _ = shapes(args)
    p = synth.hot_path_params(seed=99, device=device)
    ft = None
    if args.with_fine_transformer:
        sd = {k: v.to(device) for k, v in fine_transformer_module().state_dict().items()}
        layers = mo.transformer_params_from_state_dict(sd, len(FINE_CFG["layer_names"]))
        ft = lambda w0, w1: mo.local_feature_transformer(w0, w1, layers, FINE_CFG["layer_names"], FINE_CFG["nhead"])

    def run(fc0, fc1, ff0, ff1):
        with torch.no_grad():
            return mo.match_head(fc0, fc1, ff0, ff1, hw_i, hw_c, hw_c, hw_f, cfg["match_coarse"], p, p, fine_transformer=ft)

    return run


def synthetic_pair(args, k):
    """Pair k of the workload, generated on the host (the reference arm's input)."""
    from featurematching_b200 import synth
    _, _, hw_f, N = shapes(args)
    fc0, fc1, _ = synth.coarse_features(1, N, N, 256, args.noise, "planted", 1234 + k, sigma=args.sigma)
    ff0, ff1 = synth.fine_features(1, 64, hw_f[0], hw_f[1], 4321 + k, channels_last=False)
    return fc0, fc1, ff0, ff1


def cpu_baseline(args, pair0=None, gpu0=None):
    """The oracle port on the host cores, on ONE pair of the step repeated for a bounded time.  `pair0` = the GPU
    arm's own pair 0 (copied back) and `gpu0` = what the GPU returned for it: the oracle's result then doubles as
    the parity self-check of the timed step (ids under the tie rule, 1e-3 relative confidence, 1e-3 px)."""
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    run = oracle_head(args)
    inp = pair0 if pair0 is not None else synthetic_pair(args, 0)
    ref = run(*inp)   # warm-up; also the parity reference
    t0 = time.perf_counter()
    n = 0
    while True:
        run(*inp)
        n += 1
        el = time.perf_counter() - t0
        if el >= args.cpu_seconds or n >= args.pairs:
            break
    matches = int(ref["b_ids"].numel())
    out = {"value": n / el, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": f"pair 0 of the step's {args.pairs} pairs ({matches} matches), {n} repetitions, oracle match head"
                     f"{' with the fine transformer' if args.with_fine_transformer else ''} (torch CPU fp32, {cores} threads), {el:.1f} s"}
    parity = None
    if gpu0 is not None:
        parity = parity_check(args, inp, ref, gpu0)
    if not args.no_coarse_transformer:
        # the coarse transformer stage beside it: the oracle's restatement on ONE pair of the same token count
        from oracle import matcher_oracle as mo
        _, _, _, N = shapes(args)
        layers = mo.transformer_params_from_state_dict(coarse_transformer_module().state_dict(), len(COARSE_CFG["layer_names"]))
        f0, f1 = mo.seeded_tokens(5, 1, N, N, 256)
        with torch.no_grad():
            t0 = time.perf_counter()
            mo.local_feature_transformer(f0, f1, layers, COARSE_CFG["layer_names"], COARSE_CFG["nhead"])
            ct = time.perf_counter() - t0
        out["coarse_transformer"] = {"value": 1.0 / ct, "unit": UNIT, "ms_per_pair": 1e3 * ct,
                                     "sample": f"1 pair of 2 x {N} tokens, 8 layers, oracle (torch CPU fp32, {cores} threads)"}
    return out, parity


def synth_params():
    from featurematching_b200 import synth
    return synth.hot_path_params(seed=99)


def parity_check(args, inp, ref, gpu0):
    """GPU pair-0 records (numpy RECORD_DTYPE) against the oracle's result for the same pair."""
    import numpy as np
    import torch
    from oracle import compare, matcher_oracle as mo
    hw_i, hw_c, _, _ = shapes(args)
    ours = {"b_ids": torch.zeros(len(gpu0), dtype=torch.int64), "i_ids": torch.from_numpy(gpu0["i"].astype(np.int64)),
            "j_ids": torch.from_numpy(gpu0["j"].astype(np.int64)), "mconf": torch.from_numpy(gpu0["conf"].copy())}
    mk0 = torch.from_numpy(np.stack([gpu0["x0"], gpu0["y0"], gpu0["std0"]], 1))
    mk1 = torch.from_numpy(np.stack([gpu0["x1"], gpu0["y1"], gpu0["std1"]], 1))
    so = {(int(a), int(b)) for a, b in zip(ours["i_ids"], ours["j_ids"])}
    sr = {(int(a), int(b)) for a, b in zip(ref["i_ids"], ref["j_ids"])}
    diff = so ^ sr
    ties = 0
    if diff:
        o64 = mo.coarse_match(inp[0], inp[1], hw_i, hw_c, hw_c, dtype=torch.float64, return_conf=True)["conf_matrix"]
        ties = sum(1 for (a, b) in diff if compare._ambiguous(o64, 0, a, b, HEAD_CFG["match_coarse"]["thr"]))
    common = sorted(so & sr)
    if not common:   # nothing to compare coordinates on: that is a failed check, not a crash
        return {"parity_checked": True, "parity_ok": False, "pair": 0, "matches_gpu": len(so), "matches_oracle": len(sr),
                "id_differences": len(diff), "of_which_fp64_ties": ties, "note": "GPU and oracle share no match"}
    io = {t: k for k, t in enumerate(zip(ours["i_ids"].tolist(), ours["j_ids"].tolist()))}
    ir = {t: k for k, t in enumerate(zip(ref["i_ids"].tolist(), ref["j_ids"].tolist()))}
    ko, kr = torch.tensor([io[t] for t in common]), torch.tensor([ir[t] for t in common])
    conf_rel = float(((ours["mconf"][ko].double() - ref["mconf"][kr].double()).abs() / ref["mconf"][kr].double()).max())
    px = max(float((mk0[ko][:, :2].double() - ref["mkpts0_f"][kr][:, :2].double()).abs().max()),
             float((mk1[ko][:, :2].double() - ref["mkpts1_f"][kr][:, :2].double()).abs().max()))
    std_rel = max(float(((mk0[ko][:, 2].double() - ref["mkpts0_f"][kr][:, 2].double()).abs() / (ref["mkpts0_f"][kr][:, 2].double().abs() + 1e-2)).max()),
                  float(((mk1[ko][:, 2].double() - ref["mkpts1_f"][kr][:, 2].double()).abs() / (ref["mkpts1_f"][kr][:, 2].double().abs() + 1e-2)).max()))
    # the std bar of oracle/compare.py: |err| <= STD_ATOL + STD_RTOL * |ref|, reported as the worst ratio,
    # against the fp64 evaluation of the reference's formula on the common matches (its fp32 E[g^2] - E[g]^2 cancels
    # for peaked heat maps: compare.check_fine); the fp32 oracle's own ratio against fp64 is reported beside it
    sel = kr.tolist()
    with torch.no_grad():
        w0, w1 = mo.fine_preprocess(inp[2], inp[3], inp[0], inp[1], ref["b_ids"][sel], ref["i_ids"][sel], ref["j_ids"][sel],
                                    hw_c, hw_c, shapes(args)[2], 7, synth_params(), torch.float64)
        if args.with_fine_transformer:
            layers = mo.transformer_params_from_state_dict(fine_transformer_module().state_dict(), len(FINE_CFG["layer_names"]))
            w0, w1 = mo.local_feature_transformer(w0, w1, layers, FINE_CFG["layer_names"], FINE_CFG["nhead"], dtype=torch.float64)
        e0, e1 = mo.fine_match(w0, w1, ref["mkpts0_c"][sel], ref["mkpts1_c"][sel], synth_params(), hw_i, shapes(args)[2],
                               dtype=torch.float64)
    std_bar, std_bar_f32 = 0.0, 0.0
    for o, r, x in ((mk0[ko][:, 2], ref["mkpts0_f"][kr][:, 2], e0[:, 2]), (mk1[ko][:, 2], ref["mkpts1_f"][kr][:, 2], e1[:, 2])):
        tol = compare.STD_ATOL + compare.STD_RTOL * x.abs()
        std_bar = max(std_bar, float(((o.double() - x).abs() / tol).max()))
        std_bar_f32 = max(std_bar_f32, float(((r.double() - x).abs() / tol).max()))
    ok = (len(diff) == ties) and conf_rel <= compare.CONF_RTOL and px <= compare.COORD_TOL_PX and std_bar <= 1.0
    return {"parity_checked": True, "parity_ok": bool(ok), "pair": 0, "matches_gpu": len(so), "matches_oracle": len(sr),
            "id_differences": len(diff), "of_which_fp64_ties": ties, "max_conf_rel_err": conf_rel, "max_px_err": px,
            "max_std_rel_err": std_rel, "max_std_err_over_bar": std_bar, "std_reference": "oracle fp64",
            "fp32_oracle_std_err_over_bar": std_bar_f32,
            "bars": {"conf_rel": compare.CONF_RTOL, "px": compare.COORD_TOL_PX, "std_rel": compare.STD_RTOL,
                     "std_abs": compare.STD_ATOL}}


def run_reference(args):
    """--impl reference: the CPU restatement of the reference path on all host cores; one step =
    one pair of the workload (bounded sample).  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    run = oracle_head(args)
    inputs = [synthetic_pair(args, k) for k in range(min(4, args.pairs))]
    for k in range(args.warmup):
        run(*inputs[k % len(inputs)])
    t0 = time.perf_counter()
    for k in range(args.steps):
        run(*inputs[k % len(inputs)])
    el = time.perf_counter() - t0
    v = args.steps / el
    hw_i, hw_c, hw_f, N = shapes(args)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args, 1) + "; reference arm: 1 pair per step", "N0": N, "N1": N, "C": 256},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "1 pair per step, oracle match head on all host cores (the reference "
                                   "itself is Python under /root/reference and cannot travel to the GPU box)"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_name(args, world):
    return (f"{args.pairs} planted {args.width}x{args.height} pairs/GPU/step ({CONFIGS[args.config]['name']}"
            + (", consecutive frames of one sequence" if args.sequence else "") + "): coarse match + fine gather + "
            + ("fine transformer + " if args.with_fine_transformer else "") + "fine match + record pack"
            + ((" + match records to rank 0 (grouped send/recv" if args.gather == "root" else " + all_gather of match records (")
               + ", side stream, under the next step's fine kernels)" if world > 1 else ""))


# ------------------------------------------------------------------------------ eager PyTorch on the same GPU
def gpu_eager_baseline(args, dev, fc0, fc1, ff0, ff1, pairs=2):
    """The like-for-like GPU baseline SURVEY.md 8(d) asks for: the reference formulation (the oracle port, which
    restates the reference's einsum / softmax / unfold-gather / Linear ops) in eager PyTorch on the SAME B200, fp32,
    TF32 off, stage by stage, on `pairs` pairs of the step's own inputs.  Library kernels (cuBLAS, ATen)."""
    import torch
    from featurematching_b200 import synth
    from oracle import matcher_oracle as mo, reference_shims
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    cfg = reference_shims.default_config()
    hw_i, hw_c, hw_f, N = shapes(args)
    p = synth.hot_path_params(seed=99, device=dev)
    a0, a1 = fc0[:pairs].float(), fc1[:pairs].float()
    g0, g1 = ff0[:pairs].float().contiguous(), ff1[:pairs].float().contiguous()
    sd = {k: v.to(dev) for k, v in fine_transformer_module().state_dict().items()}
    fl = mo.transformer_params_from_state_dict(sd, 2)
    sdc = {k: v.to(dev) for k, v in coarse_transformer_module().state_dict().items()}
    cl = mo.transformer_params_from_state_dict(sdc, 8)

    def timed(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            out = fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps, out

    with torch.no_grad():
        # one pair at a time for the coarse stage: the reference keeps ~6 N0 x N1 fp32 temporaries per pair
        def coarse():
            return [mo.coarse_match(a0[b:b + 1], a1[b:b + 1], hw_i, hw_c, hw_c) for b in range(pairs)]
        ms_coarse, cm = timed(coarse)
        b_ids = torch.cat([torch.full_like(c["b_ids"], b) for b, c in enumerate(cm)])
        i_ids = torch.cat([c["i_ids"] for c in cm]); j_ids = torch.cat([c["j_ids"] for c in cm])
        mk0 = torch.cat([c["mkpts0_c"] for c in cm]); mk1 = torch.cat([c["mkpts1_c"] for c in cm])
        ms_gather, (w0, w1) = timed(lambda: mo.fine_preprocess(g0, g1, a0, a1, b_ids, i_ids, j_ids, hw_c, hw_c, hw_f, 7, p))
        ms_ft, (t0, t1) = timed(lambda: mo.local_feature_transformer(w0, w1, fl, FINE_CFG["layer_names"], 8))
        ms_fm, _ = timed(lambda: mo.fine_match(t0, t1, mk0, mk1, p, hw_i, hw_f))
        ms_ct, _ = timed(lambda: mo.local_feature_transformer(a0, a1, cl, COARSE_CFG["layer_names"], 8), reps=2)
    per = lambda ms: ms / pairs
    head = per(ms_coarse) + per(ms_gather) + per(ms_fm) + (per(ms_ft) if args.with_fine_transformer else 0.0)
    return {"unit": "ms per pair", "pairs_sampled": pairs, "matches_per_pair": int(b_ids.numel()) // pairs,
            "coarse_match": per(ms_coarse), "fine_gather": per(ms_gather), "fine_transformer": per(ms_ft),
            "fine_match": per(ms_fm), "coarse_transformer": per(ms_ct), "head_pairs_per_s": 1e3 / head,
            "how": "oracle port (reference formulation) in eager PyTorch on cuda, fp32, TF32 off, CUDA events; "
                   "coarse stage one pair per call (N0 x N1 temporaries), gather by index arithmetic on matched "
                   "cells only (cheaper than the reference's full F.unfold, fine_preprocess.py:43-46)"}


def e2e_from_images(args, dev, pairs=8, steps=3):
    """SURVEY.md 8(f) rank 4, step one -- a MEASUREMENT, no kernel claim: the whole matcher from HOST images
    (featurematching_b200.matcher.Matcher: Swin-UNet backbone in eager PyTorch = library kernels, then the B200
    coarse transformer + matching head), pinned images -> H2D -> matcher -> D2H of mkpts.  Random-init weights give
    M ~ 0 at thr 0.2, so this times backbone + coarse transformer + coarse matching; the fine
    stage sees next to nothing.  H2D is 2 x 3 x H x W fp32 per pair (7.9 MB at 640x512) instead of the 26 MB of
    features per pair the `e2e` boundary ships."""
    import torch
    from featurematching_b200 import synth
    from featurematching_b200.matcher import Matcher
    from featurematching_b200.matcher import default_config
    cfg = default_config()
    m = Matcher(cfg)
    m.load_state_dict(synth.seeded_state_dict(m, 2027))
    m = m.to(dev)
    im0, im1 = synth.images(pairs, args.height, args.width, 7)
    im0, im1 = im0.pin_memory(), im1.pin_memory()
    stage = {}

    def once(timing=False):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if timing else None
        data = {"image0": im0.to(dev, non_blocking=True), "image1": im1.to(dev, non_blocking=True)}
        if timing:
            ev[0].record()
            with torch.no_grad():
                c, f = m.backbone(torch.cat([data["image0"], data["image1"]]), emit_bf16=True)
            ev[1].record()
            t0 = c[:pairs].permute(0, 2, 3, 1).reshape(pairs, -1, 256)
            t1 = c[pairs:].permute(0, 2, 3, 1).reshape(pairs, -1, 256)
            m.coarse(t0, t1)
            ev[2].record()
            torch.cuda.synchronize()
            stage["backbone_ms"] = ev[0].elapsed_time(ev[1])
            stage["coarse_transformer_ms"] = ev[1].elapsed_time(ev[2])
        m(data)
        out = (data["mkpts0_f"].cpu(), data["mkpts1_f"].cpu(), data["mconf"].cpu())
        return out

    once(); once()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        out = once()
    el = time.perf_counter() - t0
    once(timing=True)
    return {"value": pairs * steps / el, "unit": UNIT, "pairs_per_call": pairs, "calls": steps, "ms_per_call": 1e3 * el / steps,
            "h2d_bytes_per_call": int(2 * im0.numel() * 4), "matches_per_call": int(out[0].shape[0]),
            "stage_ms_per_call": stage,
            "api": "featurematching_b200.matcher.Matcher(data) on pinned host images (network/net.py:40-92 wiring)",
            "note": "backbone = eager PyTorch (library kernels), no kernel credit claimed; random-init weights, M ~ 0"}


# ------------------------------------------------------------------------------ GPU arm
def init_gpu(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the matching head has no CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # keep stdout to the one JSON line: NCCL prints its version there at NCCL_DEBUG=VERSION
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        # the exchange is 5.8 MB per rank per step: a couple of CTAs carry it; do not let NCCL take more SMs from
        # the persistent 148-CTA kernels it runs beside
        os.environ.setdefault("NCCL_MAX_CTAS", "4")
        dist.init_process_group("nccl", device_id=dev)
    return rank, world, local, dev


def step_loop(args, head, ftm, og, lib, rank, B, inputs, shapes_, K, stage_events=True, pair_offset=None, group=None):
    """K steps of the head on `inputs` (a function k -> (fc0, fc1, ff0, ff1)), the same kernel sequence as
    MatchHead.run with stage boundaries marked by CUDA events.  Returns (result, per-step event rows)."""
    import torch
    from featurematching_b200 import _lib
    hw_i, hw_c, hw_f = shapes_
    cm, pre, fin = head.coarse_matching, head.fine_preprocess, head.fine_matching
    dev = head.device
    rows = []
    r = None
    for k in range(K):
        fc0, fc1, ff0, ff1 = inputs(k)
        nb = fc0.shape[0]
        cap = head.default_capacity(B, hw_c, hw_c, cm.border_rm)
        r = head._buffers(cap, 49, 64)
        e = [torch.cuda.Event(enable_timing=True) for _ in range(6)] if stage_events else None
        if e: e[0].record()
        r.coarse = cm.match_raw(fc0, fc1, hw_i, hw_c, hw_c, capacity=r.capacity, out=r.coarse)
        c = r.coarse
        if e: e[1].record()
        if og:
            og.launch()   # the previous step's records leave now, behind the tensor-core kernel
        pre.gather(ff0, ff1, fc0, fc1, c.b_ids, c.i_ids, c.j_ids, hw_c, hw_c, hw_f, count=c.status,
                   capacity=r.capacity, out=(r.win0, r.win1))
        if e: e[2].record()
        if args.with_fine_transformer:
            ftm.transform(r.win0, r.win1, count=c.status, out=(r.win0, r.win1))
        if e: e[3].record()
        fin.refine(r.win0, r.win1, c.mkpts0_c, c.mkpts1_c, float(hw_i[0]) / hw_f[0], count=c.status,
                   out=(r.mkpts0_f, r.mkpts1_f))
        if e: e[4].record()
        if og:
            og.launch()
        off = rank * B if pair_offset is None else pair_offset(k)
        if group is not None:
            # the sequence job exchanges every E steps: E steps' records land behind one another in one buffer
            # (fm_pack_matches_append), which is staged for the exchange when the group is complete
            E, big, counts, ci, last = group["every"], group["records"], group["counts"], group["index"], group["last"]
            rc = lib.fm_pack_matches_append(_lib.ptr(c.b_ids), _lib.ptr(c.i_ids), _lib.ptr(c.j_ids), _lib.ptr(c.mconf),
                                            _lib.ptr(r.mkpts0_f), _lib.ptr(r.mkpts1_f), _lib.ptr(c.status), r.capacity,
                                            off, _lib.ptr(counts), ci, E * r.capacity, _lib.ptr(big), _lib.stream_ptr(dev))
            _lib.check(rc, "fm_pack_matches_append")
            if og and (ci == E - 1 or last):
                og.stage(big, E * r.capacity)
        else:
            rc = lib.fm_pack_matches(_lib.ptr(c.b_ids), _lib.ptr(c.i_ids), _lib.ptr(c.j_ids), _lib.ptr(c.mconf),
                                     _lib.ptr(r.mkpts0_f), _lib.ptr(r.mkpts1_f), _lib.ptr(c.status), r.capacity,
                                     off, _lib.ptr(r.records), _lib.stream_ptr(dev))
            _lib.check(rc, "fm_pack_matches")
            if og:
                og.stage(r.records, r.capacity)
        if e:
            e[5].record()
            rows.append(e)
    return r, rows


def run_b200(args):
    import torch
    import torch.distributed as dist
    import ctypes
    from featurematching_b200 import _lib, pipeline, synth

    rank, world, local, dev = init_gpu(args)
    lib = _lib.load()
    pk = peaks()
    hw_i, hw_c, hw_f, N = shapes(args)
    B = args.pairs

    # synthetic inputs, resident in HBM (seeds differ per rank and pair)
    if args.sequence:
        # this rank's block of the sequence: B pairs = B + 1 frames (one-frame halo, shard_sequence)
        frames, _ = synth.sequence_features(B + 1, N, 256, args.noise, 1234 + rank, device=dev)
        fc0, fc1 = frames[:-1], frames[1:]
        fmaps, _ = synth.fine_features(B + 1, 64, hw_f[0], hw_f[1], 4321 + rank, device=dev)
        ff0, ff1 = fmaps[:-1], fmaps[1:]
    else:
        fc0, fc1, _ = synth.coarse_features(B, N, N, 256, args.noise, "planted", 1234 + rank * B, device=dev,
                                             sigma=args.sigma)
        ff0, ff1 = synth.fine_features(B, 64, hw_f[0], hw_f[1], 4321 + rank, device=dev)
    ftm = fine_transformer_module().to(dev)
    head = pipeline.MatchHead(HEAD_CFG, dev, fine_transformer=ftm if args.with_fine_transformer else None)
    head.load_params(synth.hot_path_params(seed=99))
    in_bytes = sum(t.numel() * t.element_size() for t in (fc0, fc1, ff0, ff1))
    dst = 0 if args.gather == "root" else None
    og = pipeline.OverlappedGather(dev, dst=dst) if world > 1 else None
    inputs = lambda k: (fc0, fc1, ff0, ff1)
    shp = (hw_i, hw_c, hw_f)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    r, _ = step_loop(args, head, ftm, og, lib, rank, B, inputs, shp, max(args.warmup, 3), stage_events=False)
    if og:
        og.finish()
    M = head.check_status(r)   # also validates: no overflow flags

    # ---- timed region: K steps, device timers, stage events inside
    K = args.steps
    t_ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    sampler = ClockSampler(local) if rank == 0 else None
    lib.fm_profile_enable(1)
    if sampler:
        sampler.start()
    barrier()
    launches0 = lib.fm_launch_count()
    t_ev[0].record()
    r, rows = step_loop(args, head, ftm, og, lib, rank, B, inputs, shp, K)
    gathered = og.finish() if og else None   # the last exchange is inside the timed region
    t_ev[1].record()
    barrier()
    launches = lib.fm_launch_count() - launches0
    clocks = sampler.stop() if sampler else None
    ms_total = t_ev[0].elapsed_time(t_ev[1])
    ms_stats, n_stats = ctypes.c_float(0), ctypes.c_int(0)
    _lib.check(lib.fm_profile_collect(ctypes.byref(ms_stats), ctypes.byref(n_stats)), "fm_profile_collect")
    lib.fm_profile_enable(0)
    if world > 1:
        t = torch.tensor([ms_total], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    M_end = head.check_status(r)
    assert M_end == M, "match count changed between steps"
    if og and gathered is not None:   # the consumer rank holds every rank's list
        lists = pipeline.unpack_records(gathered.cpu(), r.capacity, world)
        assert len(lists) == world and len(lists[rank]) == M and all(len(x) > 0 for x in lists), "bad gathered records"
    # pair 0 of this rank as the GPU returned it (for the parity self-check against the oracle)
    rec_all = r.records.cpu().numpy().view(pipeline.RECORD_DTYPE)
    rec_all = rec_all[1:1 + int(rec_all[0]["pair"])]
    gpu0 = rec_all[rec_all["pair"] == rank * B].copy()

    def span(a, b):
        return sum(e[a].elapsed_time(e[b]) for e in rows) / K
    stage_ms = [span(0, 1), span(1, 2), span(3, 4), span(4, 5)]
    ft_ms = span(2, 3)
    if not args.with_fine_transformer:
        # timed beside the step, on the step's own windows (separate output buffers)
        o = (torch.empty_like(r.win0), torch.empty_like(r.win1))
        for _ in range(2):
            ftm.transform(r.win0, r.win1, count=r.coarse.status, out=o)
        fe = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        fe[0].record()
        for _ in range(5):
            ftm.transform(r.win0, r.win1, count=r.coarse.status, out=o)
        fe[1].record()
        torch.cuda.synchronize()
        ft_ms = fe[0].elapsed_time(fe[1]) / 5
        del o
    # coarse transformer (the producer of fc0 / fc1 in the full network): timed beside the step on fp32
    # copies of the step's tokens -- with random-init weights it would scramble the planted matches
    ct_ms = None
    if not args.no_coarse_transformer:
        ctm = coarse_transformer_module().to(dev)
        t0f, t1f = fc0.float().contiguous(), fc1.float().contiguous()
        o = (torch.empty_like(t0f), torch.empty_like(t1f))
        for _ in range(2):
            ctm.transform_tokens(t0f, t1f, out=o)
        fe = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        fe[0].record()
        for _ in range(3):
            ctm.transform_tokens(t0f, t1f, out=o)
        fe[1].record()
        torch.cuda.synchronize()
        ct_ms = fe[0].elapsed_time(fe[1]) / 3
        del o, t0f, t1f, ctm
    # pose from matches, beside the step on the step's own match list: one batched RANSAC
    # over the B pairs (the synthetic matches carry no two-view geometry, so this is the no-consensus worst case:
    # every hypothesis is scored against every match, every polish runs all its rounds)
    pose_ms = None
    try:
        from featurematching_b200 import pose as fm_pose
        solver = fm_pose.PoseFromMatches(num_hypotheses=4096)
        Kmat = torch.tensor([[520.0, 0, args.width / 2], [0, 520.0, args.height / 2], [0, 0, 1]], device=dev).expand(B, 3, 3).contiguous()
        for _ in range(2):
            solver.estimate(r.coarse.b_ids, r.mkpts0_f, r.mkpts1_f, Kmat, B, count=r.coarse.status, capacity=r.capacity)
        fe = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        fe[0].record()
        for _ in range(5):
            solver.estimate(r.coarse.b_ids, r.mkpts0_f, r.mkpts1_f, Kmat, B, count=r.coarse.status, capacity=r.capacity)
        fe[1].record()
        torch.cuda.synchronize()
        pose_ms = fe[0].elapsed_time(fe[1]) / 5
    except Exception as ex:
        pose_ms = None
        sys.stderr.write(f"pose stage failed: {ex!r}\n")
    ms_step = ms_total / K
    value = world * B * K / (ms_total * 1e-3)

    # ---- end to end through the host-buffer entry point (H2D of the features, D2H of the records)
    e2e = None
    if args.e2e_steps > 0:
        host = [t.cpu().pin_memory() for t in (fc0, fc1, ff0, ff1)]   # strides (channels-last) are kept
        for _ in range(2):
            rec, h2d, d2h = head.match_host(*host, hw_i, hw_c, hw_c, hw_f, pair_offset=rank * B)
        barrier()
        t0 = time.perf_counter()
        ee = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ee[0].record()
        # a stream of batches through the public API: batch k + 1 is submitted before batch k is collected (two
        # tickets outstanding at most), so the PCIe link never idles behind a batch's last group; EVERY step's
        # inputs cross from pinned host memory and EVERY step's records and status words are read back and checked
        ticket = head.match_host_submit(*host, hw_i, hw_c, hw_c, hw_f, pair_offset=rank * B)
        for k in range(args.e2e_steps):
            nxt = (head.match_host_submit(*host, hw_i, hw_c, hw_c, hw_f, pair_offset=rank * B)
                   if k + 1 < args.e2e_steps else None)
            rec, h2d, d2h = head.match_host_result(ticket)
            assert len(rec) == M, "e2e step returned a different match count"
            ticket = nxt
        ee[1].record()
        barrier()
        ms_e2e = max(ee[0].elapsed_time(ee[1]), (time.perf_counter() - t0) * 1e3)   # host-synchronous API
        if world > 1:
            t = torch.tensor([ms_e2e], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_e2e = float(t.item())
        assert len(rec) == M
        e2e = {"value": world * B * args.e2e_steps / (ms_e2e * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": args.e2e_steps,
               "ms_per_step": ms_e2e / args.e2e_steps,
               "api": "featurematching_b200.pipeline.MatchHead.match_host_submit / match_host_result (pinned host features "
                      "-> match records; batch k + 1 submitted before batch k is collected)",
               "floor": "PCIe: this boundary ships FEATURES (26 MB per pair) from the host; at ~53 GB/s per GPU that is "
                        f"{h2d / 53e9 * 1e3:.1f} ms per step whatever the kernels do (e2e_from_images moves the boundary to images)"}
        del host

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- rooflines: ALGORITHMIC work of SURVEY.md 8(d) / CUDA-event duration, measured peaks
    flops = 2.0 * N * N * 256 * B
    ms_k1 = ms_stats.value / max(n_stats.value, 1)
    tr = lambda name: load_traffic(name)

    def hbm_entry(name, ms, bytes_alg, bytes_variant, variant_note):
        t = tr(name)
        d = {"bound": "hbm", "ms": ms, "achieved": bytes_alg / (ms * 1e-3) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
             "bytes_per_match": bytes_alg / max(M, 1), "traffic": t,
             "frac_shipped_layout": bytes_variant / (ms * 1e-3) / 1e9 / pk["hbm_gbs"], "shipped_layout": variant_note}
        if t:
            d["frac_dram"] = t / (ms * 1e-3) / 1e9 / pk["hbm_gbs"]
            d["traffic_over_algorithmic"] = t / bytes_alg
        return d

    kernels = {
        # burst peak: the kernel runs ~0.4 ms at full clock inside a short region (not the seconds-long sustained regime)
        "coarse_stats_tc": {"bound": "tensor", "ms": ms_k1, "achieved": flops / (ms_k1 * 1e-3) / 1e12,
                            "peak": pk["tflops_burst"], "unit": "TFLOP/s", "peak_kind": "burst bf16 (MEASURED_PEAKS bf16_tflops)",
                            "frac_of_sustained": flops / (ms_k1 * 1e-3) / 1e12 / pk["tflops_sustained"]},
        # the whole coarse stage: prep + statistics kernel (+ the gated exact pass) + resolve launches
        "coarse_stage": {"bound": "tensor", "ms": stage_ms[0], "achieved": flops / (stage_ms[0] * 1e-3) / 1e12,
                         "peak": pk["tflops_burst"], "unit": "TFLOP/s", "peak_kind": "burst bf16",
                         "frac_of_sustained": flops / (stage_ms[0] * 1e-3) / 1e12 / pk["tflops_sustained"]},
        # 26 112 B/match (bf16 windows); what ships writes fp32 windows: 12 544 + 1 024 + 25 088 B
        "fine_gather": hbm_entry("fine_gather", stage_ms[1], M * 26112.0, M * 38656.0,
                                 "fp32 windows (the reference's dtype): 38 656 B/match incl. 12 544 B of map reads L2 mostly absorbs"),
        # 12 584 B/match (bf16 windows) / 25 128 B/match (fp32 windows, shipped)
        "fine_match": hbm_entry("fine_match", stage_ms[2], M * 12584.0, M * 25128.0, "fp32 windows: 25 128 B/match (SURVEY 8(d) lists it)"),
    }
    # fine transformer: 4 encoder-layer applications per match (self x 2 sides, cross x 2), each
    # 49 rows x (q,k,v,merge 4 x 64x64 + mlp 128x128 + 64x128) MACs = 4.01 MFLOP; fp32-grade GEMMs
    # run as 3 bf16 tensor-core terms, so the ALGORITHMIC flop below is a third of what the pipe executes
    ft_flops = M * 4 * 49 * 2.0 * (4 * 64 * 64 + 128 * 128 + 64 * 128)
    kernels["fine_transformer"] = {"bound": "tensor", "ms": ft_ms, "achieved": ft_flops / (ft_ms * 1e-3) / 1e12,
                                   "peak": pk["tflops_sustained"], "unit": "TFLOP/s", "peak_kind": "sustained bf16",
                                   "in_step": bool(args.with_fine_transformer), "traffic": tr("fine_transformer"),
                                   "tflops_bf16_issued": 3 * ft_flops * (128 / 98.0) / (ft_ms * 1e-3) / 1e12,
                                   "note": "fp32-equivalent flop of the reference's GEMMs; executed as 3-term bf16 splits on "
                                           "128-lane tiles holding 98 rows; a serial GEMM/SIMT chain per tile (latency-bound)"}
    if ct_ms is not None:
        # 8 encoder layers x 2 sides x B x N rows; the reference's GEMMs are 2 x (4 x 256^2 + 512^2 + 256 x 512) flop
        # per row.  ALGORITHMIC bytes: read x and the source once, write x once per half-pass = 3 KB per row
        n_half = len(COARSE_CFG["layer_names"]) * 2
        ct_flops = n_half * B * N * 2.0 * (4 * 256 * 256 + 512 * 512 + 256 * 512)
        ct_bytes_alg = n_half * B * N * 3.0 * 1024
        t = tr("coarse_transformer")
        kernels["coarse_transformer"] = {"bound": "tensor", "ms": ct_ms, "achieved": ct_flops / (ct_ms * 1e-3) / 1e12,
                                         "peak": pk["tflops_sustained"], "unit": "TFLOP/s", "peak_kind": "sustained bf16",
                                         "in_step": False, "traffic": t, "algorithmic_bytes": ct_bytes_alg,
                                         "traffic_over_algorithmic": (t / ct_bytes_alg) if t else None,
                                         "tflops_bf16_issued": 3 * (ct_flops + n_half * B * N * 2.0 * 256 * 256) / (ct_ms * 1e-3) / 1e12,
                                         "note": "fp32-equivalent flop of the reference's GEMMs (3.44 TFLOP per 32 pairs at 640x512); "
                                                 "16 half-passes of 6 fused-epilogue CTA-pair GEMMs + the K'^T V reduction; every "
                                                 "intermediate makes an HBM round trip, hence traffic_over_algorithmic"}
    for kname, kv in kernels.items():
        kv["frac"] = kv["achieved"] / kv["peak"]
        kv["share_of_step"] = kv["ms"] / ms_step
    top = max((n for n in kernels if kernels[n].get("in_step", True) and n != "coarse_stage"), key=lambda n: kernels[n]["ms"])
    roofline = dict(kernels[top], kernel=top, peak_source=pk["source"] + " (" + kernels[top].get("peak_kind", "copy bandwidth") + ")")
    # BASELINE.json's metric has a second half -- "coarse-stage % bf16 TC peak": carried inside `roofline` (the driver
    # keeps this object), as fractions of the burst bf16 peak of MEASURED_PEAKS.json, beside the dominant kernel's
    for name, key in (("coarse_stats_tc", "coarse_stats_tc_frac"), ("coarse_stage", "coarse_stage_frac")):
        roofline[key] = kernels[name]["achieved"] / kernels[name]["peak"]
        roofline[key + "_of_sustained"] = kernels[name]["frac_of_sustained"]
    line = {
        "metric": METRIC if (args.width, args.height) == (640, 512) else
                  f"matched frame pairs/sec at {args.width}x{args.height}", "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(args, world),
                   "N0": N, "N1": N, "C": 256, "fine_map": [64, hw_f[0], hw_f[1]], "window": 7,
                   "matches_per_step": M, "thr": 0.2, "border_rm": 2, "temperature": 0.1,
                   "fine_transformer_in_step": bool(args.with_fine_transformer),
                   "l2": f"inputs ({in_bytes / 1e6:.0f} MB/step) exceed the 126 MB L2; no explicit flush",
                   "windows": "fp32 (what the reference hands to its fine transformer)",
                   "inputs": "resident, bf16, fine maps channels-last (what the B200 coarse transformer / backbone emit); the "
                             "fp32 -> bf16 / NCHW -> channels-last conversions of a foreign producer are outside `value`"},
        "roofline": roofline, "kernels": kernels,
        "stage_ms": {"coarse_transformer": ct_ms, "coarse_total": stage_ms[0], "fine_gather": stage_ms[1],
                     "fine_transformer": ft_ms, "fine_match": stage_ms[2], "pack_and_gather": stage_ms[3],
                     "pose_from_matches_beside_step": pose_ms},
        "e2e": e2e, "gpu_launches": int(launches),
        "gpu_launches_note": "fm_launch_count() difference over the timed region: kernels of libfm_b200.so only (NCCL's are not counted)",
        "clocks": clocks,
    }
    if world == 1:
        pair0 = tuple(t[:1].cpu() for t in (fc0, fc1, ff0, ff1))
        if not args.no_cpu_baseline:
            line["cpu_baseline"], parity = cpu_baseline(args, pair0, gpu0)
            if parity:
                line.update({"parity_checked": parity["parity_checked"], "parity_ok": parity["parity_ok"],
                             "max_px_err": parity["max_px_err"], "parity": parity})
        if not args.no_gpu_eager_baseline:
            try:
                line["gpu_eager_baseline"] = gpu_eager_baseline(args, dev, fc0, fc1, ff0, ff1)
                g = line["gpu_eager_baseline"]
                per_pair = {"coarse_match": stage_ms[0] / B, "fine_gather": stage_ms[1] / B, "fine_transformer": ft_ms / B,
                            "fine_match": stage_ms[2] / B, "coarse_transformer": (ct_ms / B) if ct_ms else None}
                g["speedup_vs_eager"] = {k: (g[k] / v) if v else None for k, v in per_pair.items()}
            except Exception as ex:   # a baseline must never take the bench line down
                line["gpu_eager_baseline"] = {"error": repr(ex)[:200]}
        if not args.no_from_images and (args.width, args.height) == (640, 512):
            try:
                del fc0, fc1, ff0, ff1
                torch.cuda.empty_cache()
                line["e2e_from_images"] = e2e_from_images(args, dev)
            except Exception as ex:
                line["e2e_from_images"] = {"error": repr(ex)[:200]}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------ configs[3] as one job
def run_job(args):
    """BASELINE configs[3]: consecutive-frame matching of ONE --frames-long 320x320 sequence (the loop of
    network/pose_estimate/test.py:79-89 over pairs (k, k + 1), datasets/scared_new2.py:76-83), block-sharded over the
    ranks with a one-frame halo (pipeline.shard_sequence); every rank streams ITS pairs through the head in steps of
    --pairs pairs, match records go to rank 0 behind each step.  Timed region = the whole job; value = pairs / time
    (strong scaling: the job is fixed)."""
    import torch
    import torch.distributed as dist
    from featurematching_b200 import _lib, pipeline, synth
    rank, world, local, dev = init_gpu(args)
    lib = _lib.load()
    hw_i, hw_c, hw_f, N = shapes(args)
    B = args.pairs
    begin, end, (f_lo, f_hi) = pipeline.shard_sequence(args.frames, rank, world)
    n_pairs, n_frames = end - begin, f_hi - f_lo
    # this rank's frames (synthetic: every rank's block is its own planted chain; a real loader would read frames
    # f_lo .. f_hi - 1), generated in chunks to bound the fp32 transients
    srank = rank if args.seed_rank is None else args.seed_rank   # (debug) another rank's synthetic block
    toks, maps = [], []
    for c0 in range(0, n_frames, 512):
        c1 = min(c0 + 512 + 1, n_frames)   # one frame of overlap keeps the chain continuous across chunks
        F, _ = synth.sequence_features(c1 - c0, N, 256, args.noise, 1234 + 7919 * srank + c0, device=dev)
        G, _ = synth.fine_features(c1 - c0, 64, hw_f[0], hw_f[1], 4321 + 7919 * srank + c0, device=dev)
        last = c1 == n_frames
        toks.append(F if last else F[:-1]); maps.append(G if last else G[:-1])
        if last:
            break
    frames = torch.cat(toks); fmaps = torch.cat(maps).contiguous(memory_format=torch.channels_last)
    del toks, maps
    assert frames.shape[0] == n_frames
    ftm = fine_transformer_module().to(dev)
    head = pipeline.MatchHead(HEAD_CFG, dev, fine_transformer=ftm if args.with_fine_transformer else None)
    head.load_params(synth.hot_path_params(seed=99))
    og = pipeline.OverlappedGather(dev, dst=0 if args.gather == "root" else None) if world > 1 else None
    steps = (n_pairs + B - 1) // B

    def inputs(k):
        lo, hi = k * B, min((k + 1) * B, n_pairs)
        return frames[lo:hi], frames[lo + 1:hi + 1], fmaps[lo:hi], fmaps[lo + 1:hi + 1]

    shp = (hw_i, hw_c, hw_f)
    E = max(1, min(args.exchange_every, 64))
    cap = head.default_capacity(B, hw_c, hw_c, head.coarse_matching.border_rm)
    grp = {"every": E, "records": torch.zeros((E * cap + 1) * _lib.RECORD_BYTES, dtype=torch.uint8, device=dev),
           "counts": torch.zeros(_lib.MAX_CHUNKS, dtype=torch.int32, device=dev)}
    step_loop(args, head, ftm, og, lib, rank, B, inputs, shp, min(max(args.warmup, 3), steps), stage_events=False)
    if og:
        # warm-up of the exchange itself, at the timed region's message size and through BOTH staging slots: NCCL
        # connects peers (and sizes its buffers) lazily at the first send / receive -- ~150 ms once, which at
        # N > 1 otherwise lands inside a job that takes 300-700 ms in all
        og.finish()
        for _ in range(2):
            og.stage(grp["records"], E * cap)
            og.launch()
        og.finish()
    torch.cuda.synchronize()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()   # BEFORE the barrier: it sleeps 150 ms, and a rank that enters the job late makes every
                          # other rank wait at its first exchange (r02: the whole "rank 1 is 27 % slower" mystery)
    if world > 1:
        dist.barrier()
    counts = torch.zeros(steps, dtype=torch.int32, device=dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    launches0 = lib.fm_launch_count()
    torch.cuda.synchronize()
    t_host0 = time.perf_counter()
    ev[0].record()
    step_ev = [] if os.environ.get("FM_BENCH_STEP_TRACE") else None   # (debug) per-step device times of every rank
    for k in range(steps):
        grp["index"], grp["last"] = k % E, k == steps - 1
        r, _ = step_loop(args, head, ftm, og, lib, rank, B, lambda _k, kk=k: inputs(kk), shp, 1, stage_events=False,
                         pair_offset=lambda _k, kk=k: begin + kk * B, group=grp)
        counts[k:k + 1].copy_(r.coarse.status[:1], non_blocking=True)   # device-side bookkeeping, no sync
        if step_ev is not None:
            step_ev.append(torch.cuda.Event(enable_timing=True)); step_ev[-1].record()
    if og:
        og.finish()
    ev[1].record()
    host_ms = (time.perf_counter() - t_host0) * 1e3   # host-side enqueue time of the whole job (no sync inside)
    torch.cuda.synchronize()
    if step_ev:
        ts = [ev[0].elapsed_time(e) for e in step_ev]
        sys.stderr.write(f"rank {rank} step ms: " + " ".join(f"{b - a:.2f}" for a, b in zip([0.0] + ts[:-1], ts)) + "\n")
    ms = ev[0].elapsed_time(ev[1])
    launches = lib.fm_launch_count() - launches0
    per_rank = [[ms, host_ms, float(counts.sum().item())]]
    if world > 1:
        allr = torch.zeros(world, 3, device=dev)
        allr[rank, 0], allr[rank, 1], allr[rank, 2] = ms, host_ms, float(counts.sum().item())
        dist.all_reduce(allr)
        per_rank = allr.tolist()
        ms = max(r[0] for r in per_rank)
        dist.barrier()
    clocks = sampler.stop() if sampler else None
    flags = int(r.coarse.status[_lib.ST_FLAGS].item())
    matches = torch.tensor([int(counts.sum().item()), n_pairs], device=dev, dtype=torch.int64)
    if world > 1:
        dist.all_reduce(matches)
    assert flags == 0, f"status flags {flags:#x}"
    if rank == 0:
        total_pairs = args.frames - 1
        assert int(matches[1]) == total_pairs
        line = {"metric": f"matched frame pairs/sec at {args.width}x{args.height}, {args.frames}-frame sequence",
                "value": total_pairs / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms / steps, "job_ms": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "bf16", "data": "synthetic",
                "config": {"workload": f"{CONFIGS[args.config]['name']}: {args.frames} frames -> {total_pairs} consecutive pairs, "
                                       f"block-sharded over {world} rank(s) with a one-frame halo (pipeline.shard_sequence), "
                                       f"{B} pairs per step per rank" + (", fine transformer in the step" if args.with_fine_transformer else "")
                                       + (f", records to rank 0 every {E} steps" if world > 1 else ""),
                           "N0": N, "N1": N, "C": 256, "fine_map": [64, hw_f[0], hw_f[1]], "pairs_rank0": n_pairs,
                           "matches_total": int(matches[0]), "l2": "every step reads new frames (no reuse between steps)"},
                "gpu_launches": int(launches), "clocks": clocks,
                "per_rank_ms": [round(r[0], 2) for r in per_rank],
                "per_rank_host_enqueue_ms": [round(r[1], 2) for r in per_rank],
                "per_rank_matches": [int(r[2]) for r in per_rank]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def load_traffic(kernel):
    """DRAM bytes per launch from the committed ncu summary (profiles/traffic.json), or None."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get(kernel)
    return None


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    elif a.job:
        run_job(a)
    else:
        run_b200(a)
