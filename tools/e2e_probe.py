#!/usr/bin/env python
"""Where does the end-to-end (host buffers) time go?  Times the raw H2D copies of one step's
features and MatchHead.match_host with several chunk counts.  GPU box only."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from featurematching_b200 import pipeline, synth  # noqa: E402

dev = torch.device("cuda:0")
hw_i, hw_c, hw_f, N, B = (512, 640), (64, 80), (256, 320), 5120, 32
cfg = {"fine_window_size": 7, "fine_concat_coarse_feat": True, "coarse": {"d_model": 256},
       "fine": {"d_model": 64}, "match_coarse": {"thr": 0.2, "border_rm": 2, "dsmax_temperature": 0.1}}
fc0, fc1, _ = synth.coarse_features(B, N, N, 256, 0.1, "planted", 1234, device=dev)
ff0, ff1 = synth.fine_features(B, 64, hw_f[0], hw_f[1], 4321, device=dev)
host = [t.cpu().pin_memory() for t in (fc0, fc1, ff0, ff1)]
print("pinned:", [t.is_pinned() for t in host], [t[8:16].is_pinned() for t in host])
nbytes = sum(t.numel() * t.element_size() for t in host)
dst = [torch.empty_like(t, device=dev) for t in host]


def timed(fn, n=5):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


def raw_copy():
    for d, t in zip(dst, host):
        d.copy_(t, non_blocking=True)


ms = timed(raw_copy)
print(f"raw H2D of one step: {ms:.2f} ms = {nbytes / ms / 1e6:.1f} GB/s")


def raw_copy_chunked():
    for k in range(4):
        for d, t in zip(dst, host):
            d[8 * k:8 * k + 8].copy_(t[8 * k:8 * k + 8], non_blocking=True)


ms = timed(raw_copy_chunked)
print(f"raw H2D in 4 chunks: {ms:.2f} ms = {nbytes / ms / 1e6:.1f} GB/s")

head = pipeline.MatchHead(cfg, dev)
head.load_params(synth.hot_path_params(seed=99))
for ch in (1, 2, 4, 8, 16):
    ms = timed(lambda: head.match_host(*host, hw_i, hw_c, hw_c, hw_f, chunks=ch))
    print(f"match_host chunks={ch}: {ms:.2f} ms/step = {B / ms * 1e3:.0f} pairs/s; enqueue/wait/unpack ms = "
          + "/".join(f"{x:.2f}" for x in head.last_timing))
# device-only time of the same head on resident inputs
r = head.run(fc0, fc1, ff0, ff1, hw_i, hw_c, hw_c, hw_f)
ms = timed(lambda: head.run(fc0, fc1, ff0, ff1, hw_i, hw_c, hw_c, hw_f), n=20)
print(f"head.run resident B=32: {ms:.2f} ms")
ms = timed(lambda: [head.run(fc0[8 * k:8 * k + 8], fc1[8 * k:8 * k + 8], ff0[8 * k:8 * k + 8], ff1[8 * k:8 * k + 8], hw_i, hw_c, hw_c, hw_f) for k in range(4)], n=20)
print(f"head.run resident 4 x B=8: {ms:.2f} ms")
