"""fm_pose_from_matches_static: the pose solver for hosts that cannot allocate device memory (a plug-in whose
process gave all of the card to a framework arena, a per-frame callback where allocation latency is not allowed, a bare
C / cgo / JNI binding with no allocator).  Its scratch is an array inside libfm_b200.so instead of a caller workspace.

Checked here, through the C ABI only:
  * CPU: a problem larger than FM_POSE_STATIC_SCRATCH_BYTES is refused with FM_ERR_WORKSPACE and a message that names
    the sizes (the check comes before any device work, so it runs without a GPU);
  * GPU: two calls in a row -- a large batch, then a SMALLER one that finds the large one's leftovers in the shared
    array -- each equal to what fm_pose_from_matches returns for the same inputs and seed.  Both entry points launch the
    same deterministic kernels, so "equal" is bit-equal, the bar tests/test_pose.py already sets for a repeated call;
    the free device memory does not move across the second call.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure here should not hide the
parity tests.  It passes alone in a fresh process (the first call on the device does the one-time address look-up) and
after the rest of the suite."""
import ctypes

import numpy as np
import pytest
import torch

from featurematching_b200 import _lib, build
from tests import pose_scenes as ps


@pytest.fixture(scope="module")
def lib():
    build.build_library()
    return _lib.load()


def test_static_pose_refuses_sizes_beyond_its_array(lib):
    buf = (ctypes.c_char * 4096)()
    p = ctypes.c_void_p((ctypes.addressof(buf) + 255) & ~255)
    B, cap, H = 32, 131072, 4096                       # the bench's pose problem: about 7 MB of scratch
    assert lib.fm_pose_workspace_bytes(B, cap, H) > _lib.POSE_STATIC_SCRATCH_BYTES
    rc = lib.fm_pose_from_matches_static(p, p, p, 3, None, cap, B, p, 1.0, H, 3, 0, p, p, p, p, p, p, None)
    assert rc == -2, rc                                # FM_ERR_WORKSPACE
    msg = lib.fm_last_error_string()
    assert b"fm_pose_from_matches_static" in msg and str(_lib.POSE_STATIC_SCRATCH_BYTES).encode() in msg
    rc = lib.fm_pose_from_matches_static(None, p, p, 3, None, 16, 1, p, 1.0, 64, 3, 0, p, p, p, p, p, p, None)
    assert rc == -1                                    # FM_ERR_BAD_ARG


def _problem(cases, dev, pad):
    """The head's layout: sorted b_ids, [M, 3] rows (x, y, std), capacity-padded, count on the device."""
    rows = [ps.scene(s, M, o) for s, M, o in cases]
    b = np.concatenate([np.full(len(r[3]), k, np.int64) for k, r in enumerate(rows)])
    p0 = np.concatenate([np.concatenate([r[3], np.full((len(r[3]), 1), 0.5, np.float32)], 1) for r in rows])
    p1 = np.concatenate([np.concatenate([r[4], np.full((len(r[4]), 1), 0.5, np.float32)], 1) for r in rows])
    M = len(b)
    padf = lambda a, v: np.concatenate([a, np.full((pad,) + a.shape[1:], v, a.dtype)])
    K = np.stack([r[0] for r in rows]).astype(np.float32).reshape(len(rows), 9)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    return {"B": len(rows), "M": M, "cap": M + pad, "b": t(padf(b, 0)), "p0": t(padf(p0, np.nan)), "p1": t(padf(p1, np.nan)),
            "K": t(K), "count": torch.tensor([M], dtype=torch.int32, device=dev)}


def _outputs(pr, dev):
    B, cap = pr["B"], pr["cap"]
    return {"R": torch.full((B, 9), -7.0, device=dev), "t": torch.full((B, 3), -7.0, device=dev),
            "E": torch.full((B, 9), -7.0, device=dev), "n": torch.full((B,), -7, dtype=torch.int32, device=dev),
            "st": torch.full((B,), -7, dtype=torch.int32, device=dev), "mask": torch.zeros(cap, dtype=torch.uint8, device=dev)}


def _call(lib, pr, o, H, seed, stream, workspace=None):
    head = (_lib.ptr(pr["b"]), _lib.ptr(pr["p0"]), _lib.ptr(pr["p1"]), 3, _lib.ptr(pr["count"]), pr["cap"], pr["B"],
            _lib.ptr(pr["K"]), 1.0, H, _lib.POSE_POLISH | _lib.POSE_IDENTITY_K, seed, _lib.ptr(o["R"]), _lib.ptr(o["t"]),
            _lib.ptr(o["E"]), _lib.ptr(o["n"]), _lib.ptr(o["st"]), _lib.ptr(o["mask"]))
    if workspace is None:
        _lib.check(lib.fm_pose_from_matches_static(*head, stream), "fm_pose_from_matches_static")
    else:
        _lib.check(lib.fm_pose_from_matches(*head, _lib.ptr(workspace), workspace.numel(), stream), "fm_pose_from_matches")


@pytest.mark.gpu
def test_static_pose_equals_workspace_pose_on_two_calls_in_a_row(lib):
    dev = torch.device("cuda:0")
    H = 1024
    # call 1 fills most of the array; call 2 is smaller in every dimension and lands on call 1's leftovers.
    # Pair 2 of the first batch has seven matches (status 1, R = I, t = 0), the others solve.
    first = _problem([(0, 2000, 0.1), (1, 1500, 0.3), (6, 7, 0.0), (3, 900, 0.4), (4, 300, 0.2)], dev, pad=37)
    second = _problem([(5, 60, 0.1), (2, 700, 0.25)], dev, pad=5)
    with torch.cuda.device(dev):
        stream = _lib.stream_ptr(dev)
        ws = {}
        for name, pr in (("first", first), ("second", second)):
            need = lib.fm_pose_workspace_bytes(pr["B"], pr["cap"], H)
            assert 0 < need <= _lib.POSE_STATIC_SCRATCH_BYTES
            ws[name] = torch.empty(need, dtype=torch.uint8, device=dev)
        results = {}
        for seed, (name, pr) in enumerate((("first", first), ("second", second)), start=3):
            got, want = _outputs(pr, dev), _outputs(pr, dev)   # every caller-side tensor exists before the call
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info(dev)[0]
            _call(lib, pr, got, H, seed, stream)
            torch.cuda.synchronize()
            if name == "second":                              # by now the module and its kernels are resident
                assert torch.cuda.mem_get_info(dev)[0] == free_before, "the static-scratch call changed free device memory"
            _call(lib, pr, want, H, seed, stream, workspace=ws[name])
            torch.cuda.synchronize()
            results[name] = (got, want)
    for name, pr in (("first", first), ("second", second)):
        got, want = results[name]
        assert want["st"].min().item() >= 0, "the workspace entry point left a status unwritten"
        for k in ("st", "n", "mask", "R", "t", "E"):
            assert torch.equal(got[k], want[k]), f"{name} call: {k} differs from fm_pose_from_matches"
        assert not got["mask"][pr["M"]:].any()                # padding untouched
    st = results["first"][0]["st"].cpu().tolist()
    assert st == [0, 0, 1, 0, 0], st
    assert results["second"][0]["st"].cpu().tolist() == [0, 0]
    # and the answers are poses, not two equal piles of garbage: rotation within 5e-3 rad of the planted one (noise-free)
    K, Rtrue, ttrue, x0, x1, inl = ps.scene(0, 2000, 0.1)
    R0 = results["first"][0]["R"][0].cpu().numpy().astype(np.float64).reshape(3, 3)
    assert ps.rot_angle(Rtrue, R0) <= 5e-3
