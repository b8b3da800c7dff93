"""Parity rules between a matcher implementation and the oracle.  TEST INFRASTRUCTURE.

north_star's bar:
  * mutual-NN match indices bit-exact, except on ties within a stated epsilon of the score;
  * fine coordinates within 1e-3 px;  * confidences within 1e-3 relative.

TIE_EPS is that stated epsilon: a match present on one side only is tolerated iff, in the
fp64 oracle, its confidence is within TIE_EPS (relative) of the threshold or of the best
rival in its row or column -- i.e. fp32 rounding alone decides it.
"""
import numpy as np
import torch

TIE_EPS = 1e-4
COORD_TOL_PX = 1e-3
CONF_RTOL = 1e-3
STD_RTOL = 1e-3
STD_ATOL = 1e-5


def _triples(d):
    b = torch.as_tensor(d["b_ids"]).cpu().long()
    i = torch.as_tensor(d["i_ids"]).cpu().long()
    j = torch.as_tensor(d["j_ids"]).cpu().long()
    return torch.stack([b, i, j], 1)


def _ambiguous(conf64, b, i, j, thr):
    c = conf64[b, i, j].item()
    row = conf64[b, i].clone(); row[j] = -1
    col = conf64[b, :, j].clone(); col[i] = -1
    scale = max(c, thr)
    return (abs(c - thr) <= TIE_EPS * scale or abs(c - row.max().item()) <= TIE_EPS * scale
            or abs(c - col.max().item()) <= TIE_EPS * scale)


def check_coarse(ours, ref, conf64=None, thr=0.2):
    """Compares b/i/j ids (order included), mconf and coarse coordinates.  `conf64` (fp64
    confidence matrix from the oracle) enables the tie rule; without it sets must be equal.
    Returns the number of tolerated tie differences; raises AssertionError otherwise."""
    to, tr = _triples(ours), _triples(ref)
    so = {tuple(x) for x in to.tolist()}
    sr = {tuple(x) for x in tr.tolist()}
    diff = so ^ sr
    if diff:
        assert conf64 is not None, f"match sets differ in {len(diff)} entries: {sorted(diff)[:8]}"
        bad = [t for t in diff if not _ambiguous(conf64, *t, thr)]
        assert not bad, f"{len(bad)} non-tie differences, e.g. {sorted(bad)[:8]}"
    # order: lexicographic (b, i, j), as torch.where gives (coarse_matching_new.py:109)
    if to.shape[0] > 1:
        key = (to[:, 0] * (1 << 42)) + (to[:, 1] * (1 << 21)) + to[:, 2]
        assert bool((key[1:] > key[:-1]).all()), "match list is not sorted by (b, i, j)"
    common = sorted(so & sr)
    if common:
        io = {t: k for k, t in enumerate(map(tuple, to.tolist()))}
        ir = {t: k for k, t in enumerate(map(tuple, tr.tolist()))}
        ko = torch.tensor([io[t] for t in common])
        kr = torch.tensor([ir[t] for t in common])
        mo = torch.as_tensor(ours["mconf"]).cpu().double()[ko]
        mr = torch.as_tensor(ref["mconf"]).cpu().double()[kr]
        rel = ((mo - mr).abs() / mr.abs()).max().item()
        assert rel <= CONF_RTOL, f"mconf relative error {rel:.3e} > {CONF_RTOL}"
        for k in ("mkpts0_c", "mkpts1_c"):
            a = torch.as_tensor(ours[k]).cpu().double()[ko]
            b = torch.as_tensor(ref[k]).cpu().double()[kr]
            assert torch.equal(a, b), f"{k} differs"
    return len(diff)


def common_rows(ours, ref):
    """Row indices (into ours, into ref) of the matches both lists hold, keyed on (b, i, j): lets the fine columns be
    compared when a tolerated tie made the two lists differ in a few entries."""
    key = lambda d: list(zip(torch.as_tensor(d["b_ids"]).cpu().tolist(), torch.as_tensor(d["i_ids"]).cpu().tolist(),
                             torch.as_tensor(d["j_ids"]).cpu().tolist()))
    ko_map = {t: k for k, t in enumerate(key(ours))}
    kr_map = {t: k for k, t in enumerate(key(ref))}
    common = sorted(set(ko_map) & set(kr_map))
    return (torch.tensor([ko_map[t] for t in common], dtype=torch.long),
            torch.tensor([kr_map[t] for t in common], dtype=torch.long))


def check_fine(ours0, ours1, ref0, ref1, std_ref0=None, std_ref1=None):
    """[M, 3] = (x px, y px, std): coordinates within COORD_TOL_PX, std within STD_ATOL + STD_RTOL |std|.

    std_ref0 / std_ref1 (optional [M, 3], the oracle evaluated in fp64): the std column is then compared with THESE.
    The reference's std = sqrt(E[g^2] - E[g]^2) (fine_matching_new.py:69-73) cancels in fp32 when the heat map is
    peaked -- at std ~ 0.006 the fp32 oracle itself sits 1e-3 relative from the exact value of its own formula and a
    few matches per 18 000 land outside the bar through no fault of either side (tools/std_probe.py);
    fp64 is that exact value.  The coordinates are always compared with the fp32 oracle."""
    for o, r, r64 in ((ours0, ref0, std_ref0), (ours1, ref1, std_ref1)):
        o = torch.as_tensor(o).cpu().double()
        r = torch.as_tensor(r).cpu().double()
        assert o.shape == r.shape, f"shape {tuple(o.shape)} vs {tuple(r.shape)}"
        if o.numel() == 0:
            continue
        err = (o[:, :2] - r[:, :2]).abs().max().item()
        assert err <= COORD_TOL_PX, f"fine coordinate error {err:.3e} px > {COORD_TOL_PX}"
        if o.shape[1] > 2:
            s = r[:, 2] if r64 is None else torch.as_tensor(r64).cpu().double()[:, 2]
            serr = (o[:, 2] - s).abs()
            tol = STD_ATOL + STD_RTOL * s.abs()
            k = int(torch.argmax(serr / tol))
            assert bool((serr <= tol).all()), (f"std error {serr[k].item():.3e} at match {k} (ours {o[k, 2].item():.6e}, "
                                               f"oracle {s[k].item():.6e}{'' if r64 is None else ' fp64'}, fp32 oracle {r[k, 2].item():.6e})")


def check_windows(ours, ref, rtol=2e-5, atol=2e-5):
    o = torch.as_tensor(ours).cpu().double()
    r = torch.as_tensor(ref).cpu().double()
    assert o.shape == r.shape, f"shape {tuple(o.shape)} vs {tuple(r.shape)}"
    if o.numel():
        err = (o - r).abs()
        tol = atol + rtol * r.abs()
        assert bool((err <= tol).all()), f"window error max {err.max().item():.3e}"


def from_bits(a):
    """uint16 bf16 bit patterns (numpy) -> bf16 torch tensor."""
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int16)).view(torch.bfloat16)
