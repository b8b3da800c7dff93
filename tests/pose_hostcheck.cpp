// TEST HARNESS (host build of csrc/fm_pose_math.cuh): the pose-from-matches pipeline of csrc/pose_ransac.cu run
// sequentially on the CPU with the SAME arithmetic functions, so tests/test_pose.py can check them against OpenCV
// without a GPU.  Compiled on the fly by the test with g++; never linked into libfm_b200.so, never a fallback.
#include <cstdint>
#include <cstring>
#include <vector>

#include "../featurematching_b200/csrc/fm_pose_math.cuh"

using namespace fm::pose;

extern "C" int pose_hostcheck(const float* k0, const float* k1, int M, const float* K, float thr_px, int H, int flags,
                              uint32_t seed, float* R_out, float* t_out, float* E_out, int* n_inliers, uint8_t* mask) {
    const float focal = 0.5f * (K[0] + K[4]), cx = K[2], cy = K[5];
    if (M < 8) return 1;
    std::vector<float> x0(M), y0(M), x1(M), y1(M);
    for (int m = 0; m < M; ++m) {
        x0[m] = (k0[2 * m] - cx) / focal; y0[m] = (k0[2 * m + 1] - cy) / focal;
        x1[m] = (k1[2 * m] - cx) / focal; y1[m] = (k1[2 * m + 1] - cy) / focal;
    }
    const float thr2 = (thr_px / focal) * (thr_px / focal);
    const float trunc2 = 4.f * thr2;   // MSAC: sum of min(err, (2 thr)^2)
    // ---- hypotheses (pose_hypotheses_kernel + pose_score_kernel)
    std::vector<float> hypE(9 * H), cost(H, 3.4e38f);
    for (int h = 0; h < H; ++h) {
        int idx[8];
        sample8(seed, 0, h, M, idx);
        float A[72], E[9];
        for (int k = 0; k < 8; ++k) epipolar_row(x0[idx[k]], y0[idx[k]], x1[idx[k]], y1[idx[k]], A + 9 * k);
        if (!null_vector_8x9(A, E)) continue;
        project_essential(E);
        float c = 0.f;
        for (int m = 0; m < M; ++m) {
            const float e = sampson_error(E, x0[m], y0[m], x1[m], y1[m]);
            c += e < trunc2 ? e : trunc2;
        }
        cost[h] = c;
        std::memcpy(&hypE[9 * h], E, sizeof(E));
    }
    // ---- the kStarts best hypotheses, each polished by local optimisation (pose_polish_kernel)
    const int kStarts = (flags & 1) ? 16 : 1;
    std::vector<int> order(H);
    for (int h = 0; h < H; ++h) order[h] = h;
    for (int a = 0; a < kStarts && a < H; ++a)
        for (int b = a + 1; b < H; ++b)
            if (cost[order[b]] < cost[order[a]] || (cost[order[b]] == cost[order[a]] && order[b] < order[a])) { int t = order[a]; order[a] = order[b]; order[b] = t; }
    if (cost[order[0]] >= 3.4e38f) return 2;
    // consensus of a model: MSAC cost sum min(err, thr^2) (lower is better) and its inlier set err <= thr^2
    auto consensus = [&](const float* Em, std::vector<uint8_t>& is, int& cnt) {
        float c = 0.f;
        cnt = 0;
        for (int m = 0; m < M; ++m) {
            const float e = sampson_error(Em, x0[m], y0[m], x1[m], y1[m]);
            is[m] = e <= thr2;
            cnt += is[m];
            c += e <= thr2 ? e : thr2;
        }
        return c;
    };
    int n = -1;
    float best_cost = 3.4e38f;
    double E[9];
    std::vector<uint8_t> inl(M);
    for (int s = 0; s < kStarts && s < H; ++s) {
        if (cost[order[s]] >= 3.4e38f) break;
        float Ecur[9];
        std::memcpy(Ecur, &hypE[9 * order[s]], sizeof(Ecur));
        double Es[9];
        for (int i = 0; i < 9; ++i) Es[i] = Ecur[i];
        std::vector<uint8_t> is(M);
        int ns = 0;
        float cs = consensus(Ecur, is, ns);
        if (flags & 1) {
            // least-squares refit on the matches within kappa x threshold of the current model, kappa shrinking to 1
            // (a minimal 8-point sample is a poor estimator under pixel noise; a wide first consensus set pulls it
            // in), then at kappa = 1 while the MSAC cost keeps falling
            const float kappa[8] = {3.f, 2.5f, 2.f, 1.5f, 1.f, 1.f, 1.f, 1.f};
            for (int it = 0; it < 8; ++it) {
                const float t2 = thr2 * kappa[it] * kappa[it];
                double ata[81] = {0}, vec[81];
                for (int m = 0; m < M; ++m) {
                    if (!(sampson_error(Ecur, x0[m], y0[m], x1[m], y1[m]) <= t2)) continue;
                    double row[9];
                    epipolar_row<double>(x0[m], y0[m], x1[m], y1[m], row);
                    for (int i = 0; i < 9; ++i)
                        for (int j = 0; j < 9; ++j) ata[i * 9 + j] += row[i] * row[j];
                }
                jacobi_eigen<double, 9>(ata, vec, 30);
                int k = 0;
                for (int i = 1; i < 9; ++i) if (ata[i * 10] < ata[k * 10]) k = i;
                double Er[9];
                for (int i = 0; i < 9; ++i) Er[i] = vec[i * 9 + k];
                project_essential(Er);
                float Ef[9];
                for (int i = 0; i < 9; ++i) Ef[i] = static_cast<float>(Er[i]);
                int n2 = 0;
                std::vector<uint8_t> inl2(M);
                const float c2 = consensus(Ef, inl2, n2);
                if (kappa[it] == 1.f && !(c2 < cs)) break;
                for (int i = 0; i < 9; ++i) Ecur[i] = Ef[i];
                if (c2 < cs) { cs = c2; ns = n2; is = inl2; for (int i = 0; i < 9; ++i) Es[i] = Er[i]; }
            }
        }
        if (cs < best_cost) { best_cost = cs; n = ns; inl = is; for (int i = 0; i < 9; ++i) E[i] = Es[i]; }
    }
    double R1[9], R2[9], t[3], tn[3];
    decompose_essential(E, R1, R2, t);
    for (int i = 0; i < 3; ++i) tn[i] = -t[i];
    int good[4] = {0, 0, 0, 0};
    const bool identity_k = (flags & 2) != 0;
    for (int m = 0; m < M; ++m) {
        if (!inl[m]) continue;
        const double a0 = identity_k ? k0[2 * m] : x0[m], b0 = identity_k ? k0[2 * m + 1] : y0[m];
        const double a1 = identity_k ? k1[2 * m] : x1[m], b1 = identity_k ? k1[2 * m + 1] : y1[m];
        good[0] += in_front_of_both(R1, t, a0, b0, a1, b1, 1e9);
        good[1] += in_front_of_both(R2, t, a0, b0, a1, b1, 1e9);
        good[2] += in_front_of_both(R1, tn, a0, b0, a1, b1, 1e9);
        good[3] += in_front_of_both(R2, tn, a0, b0, a1, b1, 1e9);
    }
    const int c = best_candidate(good);
    if (good[c] == 0) {   // recoverPose's n = 0: the reference returns None (pose_pred.py:143-151)
        for (int i = 0; i < 9; ++i) { R_out[i] = (i % 4 == 0) ? 1.f : 0.f; E_out[i] = static_cast<float>(E[i]); }
        for (int i = 0; i < 3; ++i) t_out[i] = 0.f;
        *n_inliers = n;
        std::memcpy(mask, inl.data(), M);
        return 3;
    }
    const double* R = (c & 1) ? R2 : R1;
    const double* tt = (c & 2) ? tn : t;
    for (int i = 0; i < 9; ++i) { R_out[i] = static_cast<float>(R[i]); E_out[i] = static_cast<float>(E[i]); }
    for (int i = 0; i < 3; ++i) t_out[i] = static_cast<float>(tt[i]);
    *n_inliers = n;
    std::memcpy(mask, inl.data(), M);
    return 0;
}
