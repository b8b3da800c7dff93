"""Synthetic two-view geometry with PLANTED inliers for the pose tests (numpy-seeded; TEST INFRASTRUCTURE)."""
import cv2
import numpy as np


def scene(seed, M=2000, outlier=0.2, f=520.0, W=640, H=512, noise=0.0, clean_margin_px=4.0):
    """K (fx = fy, so the reference's focal = (fx + fy) / 2 model is exact), R, unit t, pixel matches x0 / x1
    [M, 2] float32 and the planted-inlier mask.  Outliers are uniform in image 1 but re-drawn until they lie more
    than `clean_margin_px` from the true epipolar line, so no outlier is an accidental inlier of the true model."""
    rs = np.random.RandomState(seed)
    K = np.array([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1.0]])
    R, _ = cv2.Rodrigues(rs.uniform(-0.1, 0.1, 3))
    t = rs.uniform(-1, 1, 3)
    t /= np.linalg.norm(t)
    X = np.stack([rs.uniform(-2, 2, M * 4), rs.uniform(-1.6, 1.6, M * 4), rs.uniform(3, 9, M * 4)], 1)
    x0 = (K @ X.T).T
    x0 = x0[:, :2] / x0[:, 2:]
    X1 = (R @ X.T).T + t * 0.5
    x1 = (K @ X1.T).T
    x1 = x1[:, :2] / x1[:, 2:]
    ok = ((x0[:, 0] > 0) & (x0[:, 0] < W) & (x0[:, 1] > 0) & (x0[:, 1] < H) & (x1[:, 0] > 0) & (x1[:, 0] < W)
          & (x1[:, 1] > 0) & (x1[:, 1] < H) & (X1[:, 2] > 0.5))
    x0, x1 = x0[ok][:M], x1[ok][:M].copy()
    n = len(x0)
    x1 += noise * rs.standard_normal(x1.shape)
    inl = np.ones(n, bool)
    idx = rs.choice(n, int(outlier * n), replace=False)
    inl[idx] = False
    tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
    F = np.linalg.inv(K).T @ (tx @ R) @ np.linalg.inv(K)
    for k in idx:
        line = F @ np.array([x0[k, 0], x0[k, 1], 1.0])
        while True:
            cand = np.array([rs.uniform(0, W), rs.uniform(0, H)])
            if abs(line[0] * cand[0] + line[1] * cand[1] + line[2]) / np.hypot(line[0], line[1]) > clean_margin_px + 3 * noise:
                x1[k] = cand
                break
    return K, R, t, np.ascontiguousarray(x0.astype(np.float32)), np.ascontiguousarray(x1.astype(np.float32)), inl


def reference_estimate_pose(kpts0, kpts1, K0, identity_k=True):
    """network/utils/pose_pred.py:130-151 verbatim in behaviour (cv2 is the oracle of this stage); identity_k = False
    hands recoverPose the camera matrix instead of np.eye(3)."""
    if len(kpts0) < 5:
        return None
    focal = 0.5 * (K0[0, 0] + K0[1, 1])
    pp = (K0[0, 2], K0[1, 2])
    E, mask = cv2.findEssentialMat(kpts0, kpts1, focal, pp, cv2.RANSAC, 0.999, 1.0)
    if E is None:
        return None
    Kr = np.eye(3) if identity_k else np.array([[focal, 0, pp[0]], [0, focal, pp[1]], [0, 0, 1.0]])
    best, ret = 0, None
    for _E in np.split(E, len(E) / 3):
        n, R, t, _ = cv2.recoverPose(_E, kpts0, kpts1, Kr, 1e9, mask=mask.copy())
        if n > best:
            ret = (R, t[:, 0], mask.ravel() > 0)
            best = n
    return ret


def rot_angle(R1, R2):
    return float(np.arccos(np.clip((np.trace(R1.T @ R2) - 1) / 2, -1, 1)))


def vec_angle(a, b):
    return float(np.arccos(np.clip((a @ b) / np.linalg.norm(a) / np.linalg.norm(b), -1, 1)))
