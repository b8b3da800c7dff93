#!/usr/bin/env python
"""Benchmark of the matching head (BASELINE.json metric: matched frame pairs/s at 640x512).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

One step = one pass of the hot path (coarse dual-softmax mutual-NN match -> fine window gather ->
fine dual-correlate match -> match-record pack [+ cross-GPU gather when N > 1]) over one batch of
synthetic frame pairs: BASELINE.json configs[1], 32 planted 640x512-shaped pairs per GPU, bf16
features.  The fine transformer is timed beside the step and reported
under kernels.fine_transformer; --with-fine-transformer puts it inside the step on both arms.
The coarse transformer (8(f) rank 2, the producer of the coarse tokens) is timed beside the step
on the step's tokens and reported under kernels.coarse_transformer.
Prints ONE JSON line (rank 0).  `--impl reference` times the CPU restatement of the reference
(oracle/, the reference itself cannot travel to the GPU box) on the host cores instead.
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
# kernels of libfm_b200.so per step: coarse prep, stats, finish_lse, cand_best, cand_select,
# row_blocksum, blocksum_scan, emit; gather fold, ctx_gemm, window_merge; fine_match; pack
LAUNCHES_PER_STEP = 13
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
    ap.add_argument("--with-fine-transformer", action="store_true",
                    help="put the fine LocalFeatureTransformer (SURVEY.md 8(f) rank 1, net.py:79-80) INSIDE the step, "
                         "between gather and fine match (GPU and CPU arms alike); by default it is timed beside "
                         "the step and reported under kernels.fine_transformer with in_step = false")
    a = ap.parse_args()
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


def fine_transformer_module():
    """The B200 LocalFeatureTransformer with seeded random weights (no checkpoint exists offline);
    the CPU arms read the same weights through its state dict."""
    import torch
    from featurematching_b200 import modules
    torch.manual_seed(77)
    return modules.LocalFeatureTransformer(FINE_CFG).eval()


COARSE_CFG = {"d_model": 256, "nhead": 8, "layer_names": ["self", "cross"] * 4, "attention": "linear"}   # demo/net_config.py:20-25


def coarse_transformer_module():
    """The coarse LocalFeatureTransformer with seeded random-init weights."""
    import torch
    from featurematching_b200 import modules
    torch.manual_seed(78)
    return modules.LocalFeatureTransformer(COARSE_CFG).eval()


def cpu_head_factory(args):
    """The oracle's match head on one synthetic pair (same generator as the GPU arm)."""
    import torch
    from featurematching_b200 import synth
    from oracle import matcher_oracle as mo, reference_shims
    cfg = reference_shims.default_config()
    hw_i, hw_c, hw_f, N = shapes(args)
    p = synth.hot_path_params(seed=99)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)

    def make_pair(k):
        fc0, fc1, _ = synth.coarse_features(1, N, N, 256, args.noise, "planted", 1234 + k, sigma=args.sigma)
        ff0, ff1 = synth.fine_features(1, 64, hw_f[0], hw_f[1], 4321 + k, channels_last=False)
        return fc0, fc1, ff0, ff1

    ft = None
    if args.with_fine_transformer:
        layers = mo.transformer_params_from_state_dict(fine_transformer_module().state_dict(), len(FINE_CFG["layer_names"]))
        ft = lambda w0, w1: mo.local_feature_transformer(w0, w1, layers, FINE_CFG["layer_names"], FINE_CFG["nhead"])

    def run_pair(inp):
        with torch.no_grad():
            out = mo.match_head(*inp, hw_i, hw_c, hw_c, hw_f, cfg["match_coarse"], p, p, fine_transformer=ft)
        return int(out["b_ids"].numel())

    return make_pair, run_pair, cores


def cpu_baseline(args):
    make_pair, run_pair, cores = cpu_head_factory(args)
    inp = make_pair(0)
    run_pair(inp)   # warm-up
    t0 = time.perf_counter()
    n, matches = 0, 0
    while True:
        matches += run_pair(inp)
        n += 1
        el = time.perf_counter() - t0
        if el >= args.cpu_seconds or n >= args.pairs:
            break
    out = {"value": n / el, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": f"{n} of the {args.pairs} pairs of one step ({matches // n} matches/pair), oracle "
                     f"match head (torch CPU fp32, {cores} threads), {el:.1f} s"}
    if not args.no_coarse_transformer:
        # the coarse transformer stage beside it: the oracle's restatement on ONE pair of the same token count
        import torch
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
    return out


def run_reference(args):
    """--impl reference: the CPU restatement of the reference path on all host cores; one step =
    one pair of the workload (bounded sample).  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    make_pair, run_pair, cores = cpu_head_factory(args)
    inputs = [make_pair(k) for k in range(min(4, args.pairs))]
    for k in range(args.warmup):
        run_pair(inputs[k % len(inputs)])
    t0 = time.perf_counter()
    for k in range(args.steps):
        run_pair(inputs[k % len(inputs)])
    el = time.perf_counter() - t0
    v = args.steps / el
    hw_i, hw_c, hw_f, N = shapes(args)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.pairs} planted {args.width}x{args.height} pairs/GPU/step (configs[1]); "
                               "reference arm: 1 pair per step", "N0": N, "N1": N, "C": 256},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "1 pair per step, oracle match head on all host cores (the reference "
                                   "itself is Python under /root/reference and cannot travel to the GPU box)"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------ GPU arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    import ctypes
    from featurematching_b200 import _lib, pipeline, synth

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
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    pk = peaks()
    hw_i, hw_c, hw_f, N = shapes(args)
    B = args.pairs
    cfg = {"fine_window_size": 7, "fine_concat_coarse_feat": True, "coarse": {"d_model": 256},
           "fine": FINE_CFG,
           "match_coarse": {"thr": 0.2, "border_rm": 2, "dsmax_temperature": 0.1}}

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
    head = pipeline.MatchHead(cfg, dev, fine_transformer=ftm if args.with_fine_transformer else None)
    head.load_params(synth.hot_path_params(seed=99))
    in_bytes = sum(t.numel() * t.element_size() for t in (fc0, fc1, ff0, ff1))

    # N > 1: step k's record all_gather runs on a side stream under step k + 1's kernels
    og = pipeline.OverlappedGather(dev) if world > 1 else None

    def step():
        if og:
            og.before_pack()   # warm-up only: head.run packs at its end, so wait up front
        r = head.run(fc0, fc1, ff0, ff1, hw_i, hw_c, hw_c, hw_f, pair_offset=rank * B)
        if og:
            step.gathered = og.launch(r.records, r.capacity)
        return r

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        r = step()
    M = head.check_status(r)   # also validates: no overflow flags

    # ---- timed region: K steps, device timers, stage events inside
    K = args.steps
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(1)]
    stage_ev = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(K)]
    sampler = ClockSampler(local) if rank == 0 else None
    lib.fm_profile_enable(1)
    if sampler:
        sampler.start()
    barrier()
    ev[0][0].record()
    cm, pre, fin = head.coarse_matching, head.fine_preprocess, head.fine_matching
    for k in range(K):
        # the same sequence as MatchHead.run, with stage boundaries marked by events
        e = stage_ev[k]
        e[0].record()
        r.coarse = cm.match_raw(fc0, fc1, hw_i, hw_c, hw_c, capacity=r.capacity, out=r.coarse)
        c = r.coarse
        e[1].record()
        pre.gather(ff0, ff1, fc0, fc1, c.b_ids, c.i_ids, c.j_ids, hw_c, hw_c, hw_f, count=c.status,
                   capacity=r.capacity, out=(r.win0, r.win1))
        e[2].record()
        if args.with_fine_transformer:
            ftm.transform(r.win0, r.win1, count=c.status, out=(r.win0, r.win1))
        e[5].record()
        fin.refine(r.win0, r.win1, c.mkpts0_c, c.mkpts1_c, float(hw_i[0]) / hw_f[0], count=c.status,
                   out=(r.mkpts0_f, r.mkpts1_f))
        e[3].record()
        if og:
            og.before_pack()
        rc = lib.fm_pack_matches(_lib.ptr(c.b_ids), _lib.ptr(c.i_ids), _lib.ptr(c.j_ids), _lib.ptr(c.mconf),
                                 _lib.ptr(r.mkpts0_f), _lib.ptr(r.mkpts1_f), _lib.ptr(c.status), r.capacity,
                                 rank * B, _lib.ptr(r.records), _lib.stream_ptr(dev))
        _lib.check(rc, "fm_pack_matches")
        if og:
            step.gathered = og.launch(r.records, r.capacity)
        e[4].record()
    if og:
        og.finish()   # the last exchange is inside the timed region
    ev[0][1].record()
    barrier()
    clocks = sampler.stop() if sampler else None
    ms_total = ev[0][0].elapsed_time(ev[0][1])
    ms_stats, n_stats = ctypes.c_float(0), ctypes.c_int(0)
    _lib.check(lib.fm_profile_collect(ctypes.byref(ms_stats), ctypes.byref(n_stats)), "fm_profile_collect")
    lib.fm_profile_enable(0)
    if world > 1:
        t = torch.tensor([ms_total], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    M_end = head.check_status(r)
    assert M_end == M, "match count changed between steps"
    if og:   # the overlapped exchange delivered every rank's list
        lists = pipeline.unpack_records(step.gathered.cpu(), r.capacity, world)
        assert len(lists) == world and len(lists[rank]) == M and all(len(x) > 0 for x in lists), "bad gathered records"
    def span(a, b):
        return sum(stage_ev[k][a].elapsed_time(stage_ev[k][b]) for k in range(K)) / K
    stage_ms = [span(0, 1), span(1, 2), span(5, 3), span(3, 4)]
    ft_ms = span(2, 5)
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
        for _ in range(args.e2e_steps):
            rec, h2d, d2h = head.match_host(*host, hw_i, hw_c, hw_c, hw_f, pair_offset=rank * B)
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
               "api": "featurematching_b200.pipeline.MatchHead.match_host (pinned host features -> match records)"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- rooflines (algorithmic work / CUDA-event duration; SURVEY.md 8d figures, fp32 windows)
    flops = 2.0 * N * N * 256 * B
    ms_stats_avg = ms_stats.value / max(n_stats.value, 1)
    gather_bytes = M * (2 * 49 * 64 * 2 + 2 * 256 * 2 + 2 * 49 * 64 * 4)
    match_bytes = M * (2 * 49 * 64 * 4 + 16 + 24)
    kernels = {
        "coarse_stats_tc": {"bound": "tensor", "ms": ms_stats_avg, "achieved": flops / (ms_stats_avg * 1e-3) / 1e12,
                            "peak": pk["tflops_sustained"], "unit": "TFLOP/s"},
        "fine_gather": {"bound": "hbm", "ms": stage_ms[1], "achieved": gather_bytes / (stage_ms[1] * 1e-3) / 1e9,
                        "peak": pk["hbm_gbs"], "unit": "GB/s"},
        "fine_match": {"bound": "hbm", "ms": stage_ms[2], "achieved": match_bytes / (stage_ms[2] * 1e-3) / 1e9,
                       "peak": pk["hbm_gbs"], "unit": "GB/s"},
    }
    # fine transformer: 4 encoder-layer applications per match (self x 2 sides, cross x 2), each
    # 49 rows x (q,k,v,merge 4 x 64x64 + mlp 128x128 + 64x128) MACs = 4.01 MFLOP; fp32-grade GEMMs
    # run as 3 bf16 tensor-core terms, so the ALGORITHMIC flop below is a third of what the pipe executes
    ft_flops = M * 4 * 49 * 2.0 * (4 * 64 * 64 + 128 * 128 + 64 * 128)
    kernels["fine_transformer"] = {"bound": "tensor", "ms": ft_ms, "achieved": ft_flops / (ft_ms * 1e-3) / 1e12,
                                   "peak": pk["tflops_sustained"], "unit": "TFLOP/s",
                                   "in_step": bool(args.with_fine_transformer), "traffic": load_traffic("fine_transformer"),
                                   "note": "fp32-equivalent flop; executed as 3-term bf16 splits, SIMT/latency-bound"}
    if ct_ms is not None:
        # 8 encoder layers x 2 sides x B x N rows.  Every intermediate of a half-pass makes one HBM round trip as
        # bf16 hi + lo planes (or fp32), 1 KB per row per 256 channels: KV 3 KB, K'^T V 2, Q 2, MSG 2, MERGE 2,
        # MLP1 4, MLP2 5 KB per row + 64 B of Z = the stage's algorithmic bytes; the reference's
        # GEMMs are 2 x (4 x 256^2 + 512^2 + 256 x 512) flop per row, executed as 3-term bf16 splits
        n_half = len(COARSE_CFG["layer_names"]) * 2
        ct_bytes = n_half * B * N * (20 * 1024 + 64.0)
        ct_flops = n_half * B * N * 2.0 * (4 * 256 * 256 + 512 * 512 + 256 * 512)
        kernels["coarse_transformer"] = {"bound": "hbm", "ms": ct_ms, "achieved": ct_bytes / (ct_ms * 1e-3) / 1e9,
                                         "peak": pk["hbm_gbs"], "unit": "GB/s", "in_step": False,
                                         "traffic": load_traffic("coarse_transformer"),
                                         "tflops_fp32_equivalent": ct_flops / (ct_ms * 1e-3) / 1e12,
                                         "tflops_bf16_issued": 3 * (ct_flops + n_half * B * N * 2.0 * 256 * 256) / (ct_ms * 1e-3) / 1e12,
                                         "note": "16 half-passes of 6 fused-epilogue CTA-pair GEMMs + the K'^T V reduction; "
                                                 "HBM-bound by the operand round trips between the GEMMs"}
    for kname, kv in kernels.items():
        kv["frac"] = kv["achieved"] / kv["peak"]
        kv["share_of_step"] = kv["ms"] / ms_step
    top = max((n for n in kernels if kernels[n].get("in_step", True)), key=lambda n: kernels[n]["ms"])
    traffic = load_traffic(top)
    roofline = dict(kernels[top], kernel=top, traffic=traffic, peak_source=pk["source"] +
                    (" (sustained bf16)" if kernels[top]["bound"] == "tensor" else " (copy)"))
    line = {
        "metric": METRIC if (args.width, args.height) == (640, 512) else
                  f"matched frame pairs/sec at {args.width}x{args.height}", "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"{B} planted {args.width}x{args.height} pairs/GPU/step ({CONFIGS[args.config]['name']}"
                               + (", consecutive frames of one sequence" if args.sequence else "") + "): "
                               "coarse match + fine gather + " + ("fine transformer + " if args.with_fine_transformer else "")
                               + "fine match + record pack"
                               + (" + NCCL all_gather of match records (side stream, overlapped with the next step)" if world > 1 else ""),
                   "N0": N, "N1": N, "C": 256, "fine_map": [64, hw_f[0], hw_f[1]], "window": 7,
                   "matches_per_step": M, "thr": 0.2, "border_rm": 2, "temperature": 0.1,
                   "l2": f"inputs ({in_bytes / 1e6:.0f} MB/step) exceed the 126 MB L2; no explicit flush",
                   "windows": "fp32 (what the reference hands to its fine transformer)"},
        "roofline": roofline, "kernels": kernels,
        "stage_ms": {"coarse_transformer": ct_ms, "coarse_total": stage_ms[0], "fine_gather": stage_ms[1],
                     "fine_transformer": ft_ms, "fine_match": stage_ms[2], "pack_and_gather": stage_ms[3]},
        "e2e": e2e, "gpu_launches": K * (LAUNCHES_PER_STEP + (4 if args.with_fine_transformer else 0)), "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args)
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
    else:
        run_b200(a)
