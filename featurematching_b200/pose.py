"""Relative pose from the matches, on the GPU -- host mirror of the reference's
OpenCV-on-the-CPU consumer of `mkpts*_f`:

    estimate_pose            network/utils/pose_pred.py:130-151 (and utils/metrics.py:79-109)
    Pose_Pred.get_opencv_pose network/utils/pose_pred.py:170-190
    compute_pose_errors      utils/metrics.py:124-159 (per batch element, mask = m_bids == bs)

The reference copies the match list to the host and calls cv2.findEssentialMat(..., RANSAC, 0.999, 1.0) +
cv2.recoverPose per pair; here every pair of a step is solved in one batched call of fm_pose_from_matches that
reads the head's device buffers where they lie (no device->host copy of the matches).  No CPU fallback.
"""
import torch

from . import _lib


class PoseFromMatches:
    """Batched RANSAC essential matrix + recoverPose (csrc/pose_ransac.cu).

    num_hypotheses  8-point samples per pair (deterministic in `seed`)
    threshold_px    RANSAC threshold in pixels (the reference: 1.0, pose_pred.py:136)
    polish          least-squares local optimisation of the best hypotheses (OpenCV returns its minimal-sample model)
    identity_k      rank the (R, +-t) candidates the way the reference's call does: recoverPose on PIXEL coordinates
                    with np.eye(3) (pose_pred.py:146); False = on normalised coordinates (cv2.recoverPose(E, p0, p1, K))
    """

    def __init__(self, num_hypotheses=4096, threshold_px=1.0, polish=True, identity_k=True, seed=0):
        self.num_hypotheses = int(num_hypotheses)
        self.threshold_px = float(threshold_px)
        self.flags = (_lib.POSE_POLISH if polish else 0) | (_lib.POSE_IDENTITY_K if identity_k else 0)
        self.seed = int(seed)
        self._ws = None

    def estimate(self, b_ids, mkpts0_f, mkpts1_f, K0, num_pairs, count=None, capacity=None, want_mask=True):
        """b_ids int64 [cap] sorted; mkpts*_f float32 [cap, 2 or 3]; K0 [num_pairs, 3, 3]; count = device int32
        match count for padded buffers.  Enqueues on the current stream, no host synchronisation.  Returns a dict of
        device tensors: R [B, 3, 3], t [B, 3], E [B, 3, 3], n_inliers [B], status [B] (0 ok, 1 = fewer than 8
        matches, 2 = no model, 3 = no inlier in front of both cameras, recoverPose's n = 0; R = I, t = 0 for 1-3, like
        pose_pred.py:177-179) and inlier_mask uint8 [cap].  The reference's 5-point solver also answers 5-7 matches;
        here those pairs are status 1."""
        lib = _lib.load()
        for t_ in (b_ids, mkpts0_f, mkpts1_f, K0):
            if not t_.is_cuda:
                raise _lib.FmError("featurematching_b200.pose runs on CUDA tensors only (no CPU fallback)")
        dev = mkpts0_f.device
        cap = int(b_ids.shape[0]) if capacity is None else int(capacity)
        B = int(num_pairs)
        b = b_ids.to(torch.int64).contiguous()
        p0 = mkpts0_f.to(torch.float32).contiguous()
        p1 = mkpts1_f.to(torch.float32).contiguous()
        if p0.dim() != 2 or p0.shape[1] < 2 or p1.shape != p0.shape:
            raise _lib.FmError(f"mkpts shapes {tuple(p0.shape)} / {tuple(p1.shape)}: need two equal [M, >= 2] tensors")
        K = K0.to(device=dev, dtype=torch.float32).reshape(-1, 9).contiguous()
        if K.shape[0] == 1 and B > 1:
            K = K.expand(B, 9).contiguous()     # get_opencv_pose uses data['K0'][0] for everything
        if K.shape[0] != B:
            raise _lib.FmError(f"K0 holds {K.shape[0]} matrices for {B} pairs")
        out = {"R": torch.empty(B, 3, 3, device=dev), "t": torch.empty(B, 3, device=dev),
               "E": torch.empty(B, 3, 3, device=dev), "n_inliers": torch.empty(B, dtype=torch.int32, device=dev),
               "status": torch.empty(B, dtype=torch.int32, device=dev),
               "inlier_mask": torch.zeros(max(cap, 1), dtype=torch.uint8, device=dev) if want_mask else None}
        nbytes = lib.fm_pose_workspace_bytes(B, cap, self.num_hypotheses)
        if nbytes == 0:
            raise _lib.FmError("bad pose problem size")
        if self._ws is None or self._ws.numel() < nbytes or self._ws.device != dev:
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            rc = lib.fm_pose_from_matches(_lib.ptr(b), _lib.ptr(p0), _lib.ptr(p1), int(p0.shape[1]), _lib.ptr(count), cap, B,
                                          _lib.ptr(K), self.threshold_px, self.num_hypotheses, self.flags, self.seed,
                                          _lib.ptr(out["R"]), _lib.ptr(out["t"]), _lib.ptr(out["E"]),
                                          _lib.ptr(out["n_inliers"]), _lib.ptr(out["status"]), _lib.ptr(out["inlier_mask"]),
                                          _lib.ptr(self._ws), self._ws.numel(), _lib.stream_ptr(dev))
        _lib.check(rc, "fm_pose_from_matches")
        self._keepalive = (b, p0, p1, K)   # the launch is asynchronous
        if want_mask:
            out["inlier_mask"] = out["inlier_mask"][:cap]
        return out

    def valid_hypotheses(self, num_pairs, capacity):
        """Debug: int32 [num_pairs] = how many of the samples of the LAST estimate() call (same num_pairs and capacity)
        gave a model.  Reads the workspace that call left behind."""
        lib = _lib.load()
        dev = self._ws.device
        out = torch.empty(int(num_pairs), dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            rc = lib.fm_debug_pose_valid_hypotheses(_lib.ptr(self._ws), int(num_pairs), int(capacity), self.num_hypotheses,
                                                    _lib.ptr(out), _lib.stream_ptr(dev))
        _lib.check(rc, "fm_debug_pose_valid_hypotheses")
        return out

    def estimate_from_data(self, data):
        """`data` as net.forward leaves it (m_bids, mkpts0_f, mkpts1_f) plus data['K0'] [bs, 3, 3]: one pose per
        batch element, the loop of utils/metrics.py:146-159 in one call.  Writes data['R_pred'], ['t_pred'],
        ['inliers'] (bool mask aligned with the matches) and returns the dict of `estimate`."""
        bs = int(data['bs']) if 'bs' in data else int(data['K0'].shape[0])
        out = self.estimate(data['m_bids'], data['mkpts0_f'], data['mkpts1_f'], data['K0'], bs)
        data.update({'R_pred': out['R'], 't_pred': out['t'], 'inliers': out['inlier_mask'].bool()})
        return out


def rotation_to_euler(R):
    """rotationMatrixToEulerAngles (pose_pred.py:110-127) on a [.., 3, 3] tensor, vectorised; returns [.., 3] (x, y, z)."""
    sy = torch.sqrt(R[..., 0, 0] ** 2 + R[..., 1, 0] ** 2)
    singular = sy < 1e-6
    x = torch.where(singular, torch.atan2(-R[..., 1, 2], R[..., 1, 1]), torch.atan2(R[..., 2, 1], R[..., 2, 2]))
    y = torch.atan2(-R[..., 2, 0], sy)
    z = torch.where(singular, torch.zeros_like(sy), torch.atan2(R[..., 1, 0], R[..., 0, 0]))
    return torch.stack([x, y, z], -1)


def get_opencv_pose(data, solver=None):
    """Drop-in for Pose_Pred.get_opencv_pose (pose_pred.py:170-190): the same four [1, 1, 3] tensors
    (axisangle, translation, axisangle_inv, translation_inv; Euler angles reversed to (z, y, x) as `[::-1]` does),
    from ALL matches with data['K0'][0], computed on the device."""
    solver = solver or PoseFromMatches()
    dev = data['mkpts0_f'].device
    M = data['mkpts0_f'].shape[0]
    if M < 5:   # pose_pred.py:131-132 (5 <= M < 8 also ends as R = I, t = 0: status 1 of the 8-point solver)
        R = torch.eye(3, device=dev)[None]
        t = torch.zeros(1, 3, device=dev)
    else:
        out = solver.estimate(torch.zeros(M, dtype=torch.int64, device=dev), data['mkpts0_f'], data['mkpts1_f'],
                              data['K0'][0:1], 1, want_mask=False)
        R, t = out['R'], out['t']
    R_inv = R.transpose(1, 2)   # np.linalg.inv of a rotation
    t_inv = -t
    flip = lambda e: e.flip(-1)[None]
    return flip(rotation_to_euler(R)), t[None], flip(rotation_to_euler(R_inv)), t_inv[None]
