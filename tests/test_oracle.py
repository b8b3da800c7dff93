"""The oracle restatement is pinned against outputs of the reference itself
(tests/golden/*.npz, produced by oracle/make_golden.py from /root/reference)."""
import os

import numpy as np
import pytest
import torch

from oracle import compare, matcher_oracle as mo, reference_shims

COARSE_CASES = ["coarse_planted", "coarse_nonsquare", "coarse_iid", "coarse_dup",
                "coarse_noborder_lowthr", "coarse_hot"]


def load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


@pytest.mark.parametrize("name", COARSE_CASES)
def test_coarse_oracle_matches_reference_golden(golden_dir, name):
    torch.set_num_threads(1)
    g = load(golden_dir, name)
    f0, f1 = compare.from_bits(g["feat_c0"]), compare.from_bits(g["feat_c1"])
    out = mo.coarse_match(f0, f1, tuple(g["hw_i"]), tuple(g["hw0_c"]), tuple(g["hw1_c"]),
                          float(g["thr"]), int(g["border_rm"]), float(g["temperature"]))
    for k in ("b_ids", "i_ids", "j_ids"):
        assert np.array_equal(out[k].numpy(), g[k]), k
    assert out["b_ids"].dtype == torch.int64
    np.testing.assert_allclose(out["mconf"].numpy(), g["mconf"], rtol=1e-6, atol=0)
    assert np.array_equal(out["mkpts0_c"].numpy(), g["mkpts0_c"])
    assert np.array_equal(out["mkpts1_c"].numpy(), g["mkpts1_c"])
    assert np.array_equal(out["gt_mask"].numpy(), g["gt_mask"])
    assert out["mkpts0_c"].shape == (g["b_ids"].shape[0], 2)


def test_fp64_oracle_agrees_with_fp32(golden_dir):
    g = load(golden_dir, "coarse_nonsquare")
    f0, f1 = compare.from_bits(g["feat_c0"]), compare.from_bits(g["feat_c1"])
    args = (f0, f1, tuple(g["hw_i"]), tuple(g["hw0_c"]), tuple(g["hw1_c"]))
    a = mo.coarse_match(*args, dtype=torch.float32)
    b = mo.coarse_match(*args, dtype=torch.float64, return_conf=True)
    assert compare.check_coarse(a, b, b["conf_matrix"]) == 0


def test_fine_chain_oracle_matches_reference_golden(golden_dir):
    torch.set_num_threads(1)
    g = load(golden_dir, "fine_chain")
    t = {k: torch.from_numpy(v) for k, v in g.items() if v.dtype != np.uint16}
    fc0, fc1 = compare.from_bits(g["feat_c0"]), compare.from_bits(g["feat_c1"])
    ff0, ff1 = compare.from_bits(g["feat_f0"]), compare.from_bits(g["feat_f1"])
    hw_c, hw_f, hw_i = tuple(g["hw_c"]), tuple(g["hw_f"]), tuple(g["hw_i"])
    w0, w1 = mo.fine_preprocess(ff0, ff1, fc0, fc1, t["b_ids"], t["i_ids"], t["j_ids"],
                                hw_c, hw_c, hw_f, 7, t)
    compare.check_windows(w0, g["win0"], rtol=1e-5, atol=1e-5)
    compare.check_windows(w1, g["win1"], rtol=1e-5, atol=1e-5)
    # windows without the merge layer equal a plain unfold
    r0, _ = mo.fine_preprocess(ff0, ff1, fc0, fc1, t["b_ids"], t["i_ids"], t["j_ids"],
                               hw_c, hw_c, hw_f, 7, None)
    unf = torch.nn.functional.unfold(ff0.float(), (7, 7), stride=4, padding=2)
    unf = unf.view(1, 64, 49, -1).permute(0, 3, 2, 1)
    assert torch.equal(r0, unf[t["b_ids"], t["i_ids"]])
    m0, m1 = mo.fine_match(torch.from_numpy(g["win0"]), torch.from_numpy(g["win1"]),
                           t["mkpts0_c"], t["mkpts1_c"], t, hw_i, hw_f)
    compare.check_fine(m0, m1, g["mkpts0_f"], g["mkpts1_f"])
    assert (m0 - torch.from_numpy(g["mkpts0_f"])).abs().max() < 2e-5


def test_fine_match_special_cases(golden_dir):
    g = load(golden_dir, "fine_match_special")
    t = {k: torch.from_numpy(v) for k, v in g.items() if v.dtype != np.uint16}
    w0, w1 = compare.from_bits(g["win0"]), compare.from_bits(g["win1"])
    m0, m1 = mo.fine_match(w0, w1, t["mkpts0_c"], t["mkpts1_c"], t, tuple(g["hw_i"]),
                           tuple(g["hw_f"]))
    compare.check_fine(m0, m1, g["mkpts0_f"], g["mkpts1_f"])
    # uniform heat-map (rows 8..15): expectation 0 -> +3 px offset, std = 2 * sqrt(sum g^2 / 7)
    grid = mo.window_grid(7)
    std_uniform = 2 * float(torch.sqrt((grid[:, 0] ** 2).mean()))
    assert torch.allclose(m0[8:16, 2], torch.full((8,), std_uniform), atol=1e-5)
    assert torch.allclose(m0[8:16, :2], t["mkpts0_c"][8:16] + 3.0, atol=1e-5)
    # one-hot heat-map (rows 16..23): std hits the clamp 2 * sqrt(1e-10)
    assert torch.allclose(m0[16:24, 2], torch.full((8,), 2e-5), atol=1e-6)
    for m in range(16, 24):
        cell = (m * 11) % 49
        expect = t["mkpts0_c"][m] + grid[cell] * 3 * 2.0 + 3
        assert torch.allclose(m0[m, :2], expect, atol=1e-4)


def test_empty_match_list_shapes():
    f0 = torch.randn(1, 80, 256).bfloat16()
    f1 = torch.randn(1, 80, 256).bfloat16()
    out = mo.coarse_match(f0, f1, (64, 80), (8, 10), (8, 10))
    assert out["b_ids"].shape == (0,) and out["mkpts0_c"].shape == (0, 2)
    w0, w1 = mo.fine_preprocess(torch.zeros(1, 64, 32, 40), torch.zeros(1, 64, 32, 40), f0, f1,
                                out["b_ids"], out["i_ids"], out["j_ids"], (8, 10), (8, 10),
                                (32, 40))
    assert w0.shape == (0, 49, 64)
    m0, m1 = mo.fine_match(w0, w1, out["mkpts0_c"], out["mkpts1_c"], None, (64, 80), (32, 40))
    assert m0.shape == (0, 2)


@pytest.mark.skipif(not reference_shims.available(), reason="reference not present")
def test_oracle_against_live_reference_medium():
    """Build container only: the real reference vs the restatement at a medium size."""
    from featurematching_b200 import synth
    ref = reference_shims.load_reference()
    cfg = reference_shims.default_config()
    hw_c = (20, 24)
    N = hw_c[0] * hw_c[1]
    f0, f1, _ = synth.coarse_features(2, N, N, 256, 0.4, "dup", 77)
    data = {"hw0_i": (160, 192), "hw1_i": (160, 192), "hw0_c": hw_c, "hw1_c": hw_c}
    with torch.no_grad():
        ref.CoarseMatching(cfg["match_coarse"]).eval()(f0.float(), f1.float(), data)
    out = mo.coarse_match(f0, f1, (160, 192), hw_c, hw_c)
    assert data["b_ids"].numel() > 50
    for k in ("b_ids", "i_ids", "j_ids"):
        assert torch.equal(out[k], data[k])
    assert torch.allclose(out["mconf"], data["mconf"], rtol=1e-6, atol=0)
