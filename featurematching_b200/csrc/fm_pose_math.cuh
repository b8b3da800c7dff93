// Two-view geometry for the pose-from-matches stage: everything the kernels of
// pose_ransac.cu evaluate per hypothesis / per match / per pair, as plain functions that compile for the device
// AND for the host -- tests/pose_hostcheck.cpp builds them with g++ and checks the arithmetic against OpenCV on
// the CPU before any GPU time is spent.  (The host build is a test harness; the product path is the CUDA one.)
//
// What the reference does (network/utils/pose_pred.py:130-151, utils/metrics.py:79-109), through OpenCV:
//   E, mask = cv2.findEssentialMat(kpts0, kpts1, focal = (fx + fy) / 2, pp = (cx, cy), RANSAC, 0.999, 1.0)
//       points normalised as ((x - cx) / focal, (y - cy) / focal); a match is an inlier of a model E iff its
//       Sampson error  (x1' E x0)^2 / ((E x0)_0^2 + (E x0)_1^2 + (E' x1)_0^2 + (E' x1)_1^2)  <=  (1.0 / focal)^2
//   n, R, t, _ = cv2.recoverPose(E, kpts0, kpts1, np.eye(3), 1e9, mask)
//       E = U diag(1,1,0) V' -> R1 = U W V', R2 = U W' V', t = u3; the four (R, +-t) candidates are ranked by the
//       number of inliers whose DLT triangulation has positive, finite depth in both cameras.  NOTE the call
//       passes the PIXEL coordinates with an identity camera matrix; kIdentityK reproduces exactly that.
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define FM_HD __host__ __device__ __forceinline__
#else
#define FM_HD inline
#endif

namespace fm {
namespace pose {

// ---------------------------------------------------------------- small dense linear algebra
// Cyclic Jacobi eigen-decomposition of a symmetric N x N matrix (row-major, destroyed): eigenvalues on the
// diagonal of `a`, eigenvectors in the COLUMNS of `v`.
template <typename T, int N>
FM_HD void jacobi_eigen(T* a, T* v, int sweeps) {
    for (int i = 0; i < N; ++i)
        for (int j = 0; j < N; ++j) v[i * N + j] = i == j ? T(1) : T(0);
    for (int s = 0; s < sweeps; ++s) {
        T off = 0, diag = 0;
        for (int p = 0; p < N; ++p) {
            diag += a[p * N + p] * a[p * N + p];
            for (int q = p + 1; q < N; ++q) off += a[p * N + q] * a[p * N + q];
        }
        // converged: the off-diagonal mass is below the rounding noise of the diagonal (cyclic Jacobi converges
        // quadratically: 5-7 sweeps at these sizes; the fixed sweep counts are only the upper bound)
        if (off <= diag * (sizeof(T) == 8 ? T(1e-30) : T(1e-13))) break;
        for (int p = 0; p < N; ++p)
            for (int q = p + 1; q < N; ++q) {
                const T apq = a[p * N + q];
                if (apq == T(0)) continue;
                const T theta = (a[q * N + q] - a[p * N + p]) / (T(2) * apq);
                const T tt = (theta >= T(0) ? T(1) : T(-1)) / (std::fabs(theta) + std::sqrt(theta * theta + T(1)));
                const T c = T(1) / std::sqrt(tt * tt + T(1)), sn = tt * c;
                for (int k = 0; k < N; ++k) {   // A <- A J (columns p, q)
                    const T akp = a[k * N + p], akq = a[k * N + q];
                    a[k * N + p] = c * akp - sn * akq;
                    a[k * N + q] = sn * akp + c * akq;
                }
                for (int k = 0; k < N; ++k) {   // A <- J' A (rows p, q)
                    const T apk = a[p * N + k], aqk = a[q * N + k];
                    a[p * N + k] = c * apk - sn * aqk;
                    a[q * N + k] = sn * apk + c * aqk;
                }
                for (int k = 0; k < N; ++k) {
                    const T vkp = v[k * N + p], vkq = v[k * N + q];
                    v[k * N + p] = c * vkp - sn * vkq;
                    v[k * N + q] = sn * vkp + c * vkq;
                }
            }
    }
}

template <typename T>
FM_HD void cross3(const T* a, const T* b, T* c) {
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}

// E (row-major 3 x 3) = U diag(s) V' with s sorted descending; U, V proper or improper orthogonal (det +-1),
// columns in `u`, `v` (row-major storage: u[r * 3 + c]).  Built from the eigen-decomposition of E' E; the left
// vector of a vanishing singular value is completed by a cross product.
template <typename T>
FM_HD void svd3(const T* E, T* u, T* s, T* v) {
    T a[9], w[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) a[i * 3 + j] = E[0 * 3 + i] * E[0 * 3 + j] + E[1 * 3 + i] * E[1 * 3 + j] + E[2 * 3 + i] * E[2 * 3 + j];
    jacobi_eigen<T, 3>(a, w, 12);
    int ord[3] = {0, 1, 2};
    for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 2 - i; ++j)
            if (a[ord[j] * 4] < a[ord[j + 1] * 4]) { const int t = ord[j]; ord[j] = ord[j + 1]; ord[j + 1] = t; }
    for (int c = 0; c < 3; ++c) {
        const T ev = a[ord[c] * 4];
        s[c] = ev > T(0) ? std::sqrt(ev) : T(0);
        for (int r = 0; r < 3; ++r) v[r * 3 + c] = w[r * 3 + ord[c]];
    }
    for (int c = 0; c < 2; ++c) {   // u_c = E v_c / s_c
        T n = 0, uc[3];
        for (int r = 0; r < 3; ++r) {
            uc[r] = E[r * 3 + 0] * v[0 * 3 + c] + E[r * 3 + 1] * v[1 * 3 + c] + E[r * 3 + 2] * v[2 * 3 + c];
            n += uc[r] * uc[r];
        }
        n = n > T(0) ? T(1) / std::sqrt(n) : T(0);
        for (int r = 0; r < 3; ++r) u[r * 3 + c] = uc[r] * n;
    }
    {   // second left vector exactly orthogonal to the first, third by the cross product
        T d = u[0] * u[1] + u[3] * u[4] + u[6] * u[7], n = 0;
        for (int r = 0; r < 3; ++r) { u[r * 3 + 1] -= d * u[r * 3 + 0]; n += u[r * 3 + 1] * u[r * 3 + 1]; }
        n = n > T(0) ? T(1) / std::sqrt(n) : T(0);
        for (int r = 0; r < 3; ++r) u[r * 3 + 1] *= n;
        const T u0[3] = {u[0], u[3], u[6]}, u1[3] = {u[1], u[4], u[7]};
        T u2[3];
        cross3(u0, u1, u2);
        u[2] = u2[0]; u[5] = u2[1]; u[8] = u2[2];
    }
}

// Nearest essential matrix (singular values 1, 1, 0), Frobenius sense: E <- u1 v1' + u2 v2'.
template <typename T>
FM_HD void project_essential(T* E) {
    T u[9], s[3], v[9];
    svd3(E, u, s, v);
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) E[r * 3 + c] = u[r * 3 + 0] * v[c * 3 + 0] + u[r * 3 + 1] * v[c * 3 + 1];
}

// ---------------------------------------------------------------- hypotheses: the 8-point equations
// One epipolar constraint x1' E x0 = 0 as a row of the 9-column design matrix (E row-major).
template <typename T>
FM_HD void epipolar_row(T x0, T y0, T x1, T y1, T* row) {
    row[0] = x1 * x0; row[1] = x1 * y0; row[2] = x1;
    row[3] = y1 * x0; row[4] = y1 * y0; row[5] = y1;
    row[6] = x0;      row[7] = y0;      row[8] = T(1);
}

// Null vector of an 8 x 9 system by Gauss-Jordan elimination with full pivoting (`A` row-major, destroyed).
// Returns false if the system is rank deficient beyond the expected one-dimensional null space.
template <typename T>
FM_HD bool null_vector_8x9(T* A, T* e) {
    int colperm[9];
    for (int c = 0; c < 9; ++c) colperm[c] = c;
    for (int k = 0; k < 8; ++k) {
        int pr = k, pc = k;
        T best = 0;
        for (int r = k; r < 8; ++r)
            for (int c = k; c < 9; ++c) {
                const T m = std::fabs(A[r * 9 + c]);
                if (m > best) { best = m; pr = r; pc = c; }
            }
        if (!(best > T(1e-12))) return false;
        if (pr != k) for (int c = 0; c < 9; ++c) { const T t = A[k * 9 + c]; A[k * 9 + c] = A[pr * 9 + c]; A[pr * 9 + c] = t; }
        if (pc != k) {
            for (int r = 0; r < 8; ++r) { const T t = A[r * 9 + k]; A[r * 9 + k] = A[r * 9 + pc]; A[r * 9 + pc] = t; }
            const int t = colperm[k]; colperm[k] = colperm[pc]; colperm[pc] = t;
        }
        const T inv = T(1) / A[k * 9 + k];
        for (int c = k; c < 9; ++c) A[k * 9 + c] *= inv;
        for (int r = 0; r < 8; ++r) {
            if (r == k) continue;
            const T f = A[r * 9 + k];
            if (f == T(0)) continue;
            for (int c = k; c < 9; ++c) A[r * 9 + c] -= f * A[k * 9 + c];
        }
    }
    // reduced system: x_k = -A[k][8] x_8 for the permuted unknowns
    T n = T(1);
    for (int k = 0; k < 8; ++k) n += A[k * 9 + 8] * A[k * 9 + 8];
    n = T(1) / std::sqrt(n);
    for (int k = 0; k < 8; ++k) e[colperm[k]] = -A[k * 9 + 8] * n;
    e[colperm[8]] = n;
    return true;
}

// Sampson error of a normalised match under E, as OpenCV's EMEstimatorCallback::computeError defines it.
template <typename T>
FM_HD T sampson_error(const T* E, T x0, T y0, T x1, T y1) {
    const T ex0 = E[0] * x0 + E[1] * y0 + E[2], ex1 = E[3] * x0 + E[4] * y0 + E[5], ex2 = E[6] * x0 + E[7] * y0 + E[8];
    const T et0 = E[0] * x1 + E[3] * y1 + E[6], et1 = E[1] * x1 + E[4] * y1 + E[7];
    const T r = x1 * ex0 + y1 * ex1 + ex2;
    return r * r / (ex0 * ex0 + ex1 * ex1 + et0 * et0 + et1 * et1);
}

// ---------------------------------------------------------------- recoverPose
// cv::decomposeEssentialMat: R1 = U W V', R2 = U W' V', t = u3, with U, V' made proper rotations first.
FM_HD void decompose_essential(const double* E, double* R1, double* R2, double* t) {
    double u[9], s[3], v[9];
    svd3(E, u, s, v);
    auto det3 = [](const double* m) {
        return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
    };
    if (det3(u) < 0) for (int i = 0; i < 9; ++i) u[i] = -u[i];
    if (det3(v) < 0) for (int i = 0; i < 9; ++i) v[i] = -v[i];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
            // W = [[0, 1, 0], [-1, 0, 0], [0, 0, 1]]:  (U W V')[r][c] = sum_k (U W)[r][k] V[c][k]
            const double uw0 = -u[r * 3 + 1], uw1 = u[r * 3 + 0], uw2 = u[r * 3 + 2];
            R1[r * 3 + c] = uw0 * v[c * 3 + 0] + uw1 * v[c * 3 + 1] + uw2 * v[c * 3 + 2];
            const double ut0 = u[r * 3 + 1], ut1 = -u[r * 3 + 0];   // U W'
            R2[r * 3 + c] = ut0 * v[c * 3 + 0] + ut1 * v[c * 3 + 1] + uw2 * v[c * 3 + 2];
        }
    t[0] = u[2]; t[1] = u[5]; t[2] = u[8];
}

// cv::triangulatePoints for one correspondence between P0 = [I | 0] and P1 = [R | t]: the right singular vector
// of the 4 x 4 DLT matrix for the smallest singular value (homogeneous X), then recoverPose's cheirality test.
FM_HD bool in_front_of_both(const double* R, const double* t, double x0, double y0, double x1, double y1, double dist) {
    double A[16], ata[16], vec[16];
    // rows: x P[2] - P[0], y P[2] - P[1] for both cameras
    A[0] = -1; A[1] = 0; A[2] = x0; A[3] = 0;
    A[4] = 0; A[5] = -1; A[6] = y0; A[7] = 0;
    for (int c = 0; c < 3; ++c) {
        A[8 + c] = x1 * R[6 + c] - R[0 + c];
        A[12 + c] = y1 * R[6 + c] - R[3 + c];
    }
    A[11] = x1 * t[2] - t[0];
    A[15] = y1 * t[2] - t[1];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) ata[i * 4 + j] = A[i] * A[j] + A[4 + i] * A[4 + j] + A[8 + i] * A[8 + j] + A[12 + i] * A[12 + j];
    jacobi_eigen<double, 4>(ata, vec, 16);
    int k = 0;
    for (int i = 1; i < 4; ++i) if (ata[i * 5] < ata[k * 5]) k = i;
    const double X = vec[0 * 4 + k], Y = vec[1 * 4 + k], Z = vec[2 * 4 + k], W = vec[3 * 4 + k];
    bool ok = Z * W > 0;                       // mask = Q.row(2).mul(Q.row(3)) > 0
    const double xn = X / W, yn = Y / W, zn = Z / W;
    ok = ok && zn < dist;
    const double z1 = R[6] * xn + R[7] * yn + R[8] * zn + t[2];   // depth in the second camera
    return ok && z1 > 0 && z1 < dist;
}

// The order in which cv::recoverPose ranks its candidates: (R1, t), (R2, t), (R1, -t), (R2, -t); first maximum wins.
FM_HD int best_candidate(const int* good) {
    if (good[0] >= good[1] && good[0] >= good[2] && good[0] >= good[3]) return 0;
    if (good[1] >= good[0] && good[1] >= good[2] && good[1] >= good[3]) return 1;
    if (good[2] >= good[0] && good[2] >= good[1] && good[2] >= good[3]) return 2;
    return 3;
}

// ---------------------------------------------------------------- sampling
FM_HD uint32_t hash32(uint32_t x) {   // lowbias32
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
// 8 distinct indices in [0, n) for hypothesis `h` of pair `b` (n >= 8): counter-based, no state
FM_HD void sample8(uint32_t seed, int b, int h, int n, int* idx) {
    uint32_t ctr = hash32(seed ^ hash32(static_cast<uint32_t>(b) * 0x9e3779b9u + static_cast<uint32_t>(h)));
    for (int k = 0; k < 8; ++k) {
        for (;;) {
            ctr = hash32(ctr + 0x6d2b79f5u);
            const int c = static_cast<int>((static_cast<uint64_t>(ctr) * static_cast<uint64_t>(n)) >> 32);
            bool dup = false;
            for (int j = 0; j < k; ++j) dup = dup || idx[j] == c;
            if (!dup) { idx[k] = c; break; }
        }
    }
}

}  // namespace pose
}  // namespace fm
