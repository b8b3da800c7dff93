// fm_pose_from_matches_static: the pose solver for callers that have no way to hand over scratch memory.
//
// fm_pose_from_matches (fm_api.cu) wants a workspace the CALLER allocated on the device.  Some hosts of this library
// cannot produce one at the point where they need a pose:
//   * a plug-in inside a serving process whose device memory was handed out once, at start-up, to the framework's own
//     arena -- the plug-in sees the tensors it is given and nothing else, and a cudaMalloc of its own would both
//     serialise the whole device behind it and may simply fail because the arena already owns the card;
//   * a per-frame callback of a live endoscopy tracker (the reference's domain): allocation is forbidden in the frame
//     loop because its latency is unbounded, and there is no "before the loop" the callback could use;
//   * a thin foreign binding (cgo, JNI, a C test harness) that links libfm_b200.so alone and has no CUDA runtime
//     handle, no torch, and therefore no allocator to ask.
// For them the scratch lives in THIS library's image: one file-scope __device__ array, of a fixed and deliberately
// small size, that every GPU which loads the library carries until the process exits.  The kernels are the ones
// fm_pose_from_matches launches (pose_ransac.cu, unchanged); only where their intermediate buffers live differs.
//
// What a user has to know
//   * the limit: fm_pose_workspace_bytes(B, capacity, num_hypotheses) must not exceed FM_POSE_STATIC_SCRATCH_BYTES
//     (2 MiB -- e.g. 8 pairs x 4096 hypotheses x 32768 matches); larger problems are refused with FM_ERR_WORKSPACE and
//     belong to fm_pose_from_matches;
//   * the first call on each device looks up the array's device address (cudaGetSymbolAddress, below).  That look-up is
//     not stream work -- it may load this file's module onto the device -- so make that first call during start-up,
//     outside any stream capture; later calls on the same device only enqueue kernels;
//   * there is ONE array per device, shared by every call from every thread and stream.  Two calls whose kernels can
//     overlap on the GPU (different streams without an event between them) overwrite each other's hypotheses and give
//     garbage: issue them on one stream or order them yourself;
//   * the array keeps whatever the previous call left in it.  Nothing here depends on that: every word the six kernels
//     read (pair ranges, normalised points, hypotheses, costs, polish starts and results) is written earlier in the same
//     call, so there is nothing to clear between calls -- tests/test_zz_no_alloc.py runs a small problem after a large
//     one to hold that property in place;
//   * no cudaMalloc / cudaFree / allocating library routine is called, on the first call or any other.
#include <mutex>

#include "fm_common.cuh"

namespace fm {
// pose_ransac.cu
size_t pose_workspace_bytes(int B, int capacity, int H);
cudaError_t launch_pose_from_matches(const int64_t* b_ids, const float* mk0, const float* mk1, int stride, const int* count,
                                     int capacity, int B, const float* K, float thr_px, int H, int flags, uint32_t seed,
                                     float* R, float* t, float* E, int* n_inliers, int* status, uint8_t* mask,
                                     void* workspace, cudaStream_t st);
}  // namespace fm

namespace {

// The scratch itself.  256-byte alignment is what pose_ransac.cu's carve() assumes of its base pointer.
static __device__ __align__(256) unsigned char g_pose_scratch[FM_POSE_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
void* g_scratch_addr[kMaxDevices] = {};   // device address of g_pose_scratch on device d, nullptr = not looked up yet

// `g_pose_scratch` in HOST code is only a handle the runtime knows the array by; it is not a device address and must
// never reach a kernel.  The address on the CURRENT device comes from cudaGetSymbolAddress, once per device.
int scratch_for_current_device(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        fm::set_error("no CUDA device available (this library has no CPU path): %s", cudaGetErrorString(e));
        return FM_ERR_CUDA;
    }
    if (dev < 0 || dev >= kMaxDevices) {
        fm::set_error("fm_pose_from_matches_static: device ordinal %d outside the %d this table covers", dev, kMaxDevices);
        return FM_ERR_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lk(g_setup_mu);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_pose_scratch);
        if (e != cudaSuccess || !p) {
            fm::set_error("cudaGetSymbolAddress(pose scratch) on device %d: %s", dev, cudaGetErrorString(e));
            return FM_ERR_CUDA;
        }
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return FM_OK;
}

}  // namespace

extern "C" int fm_pose_from_matches_static(const int64_t* b_ids, const float* mkpts0_f, const float* mkpts1_f,
                                           int pts_stride, const int32_t* count, int capacity, int B, const float* K0,
                                           float threshold_px, int num_hypotheses, int flags, uint32_t seed, float* R,
                                           float* t, float* E, int32_t* n_inliers, int32_t* pose_status,
                                           uint8_t* inlier_mask, void* stream) {
    if (!b_ids || !mkpts0_f || !mkpts1_f || !K0 || !R || !t || !n_inliers || !pose_status) {
        fm::set_error("null pointer argument");
        return FM_ERR_BAD_ARG;
    }
    if (B < 1 || capacity < 0 || pts_stride < 2 || num_hypotheses < 1 || num_hypotheses > (1 << 20) || !(threshold_px > 0.f)) {
        fm::set_error("bad argument to fm_pose_from_matches_static (B >= 1, capacity >= 0, pts_stride >= 2, 1 <= hypotheses <= 2^20, threshold > 0)");
        return FM_ERR_BAD_ARG;
    }
    const size_t need = fm::pose_workspace_bytes(B, capacity, num_hypotheses);
    if (need > static_cast<size_t>(FM_POSE_STATIC_SCRATCH_BYTES)) {
        fm::set_error("fm_pose_from_matches_static: %d pairs x %d hypotheses x %d matches need %zu bytes of scratch, the built-in "
                      "array holds %d; use fm_pose_from_matches with a workspace of your own for this size",
                      B, num_hypotheses, capacity, need, FM_POSE_STATIC_SCRATCH_BYTES);
        return FM_ERR_WORKSPACE;
    }
    void* scratch = nullptr;
    const int rc = scratch_for_current_device(&scratch);
    if (rc != FM_OK) return rc;
    const cudaError_t e = fm::launch_pose_from_matches(b_ids, mkpts0_f, mkpts1_f, pts_stride, count, capacity, B, K0,
                                                       threshold_px, num_hypotheses, flags, seed, R, t, E, n_inliers,
                                                       pose_status, inlier_mask, scratch, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) {
        fm::set_error("pose_from_matches (static scratch) launch: %s", cudaGetErrorString(e));
        return FM_ERR_CUDA;
    }
    return FM_OK;
}
