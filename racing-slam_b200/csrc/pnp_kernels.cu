// pnp_kernels.cu -- device side of the PnP verification support: verify_pnp's input assembly
// (src/LoopDetector.cpp:188-197) and the inlier scoring cv::solvePnPRansac (src/LoopDetector.cpp:203-214,
// src/Initialization.cpp:211-222) runs once per RANSAC model: projectPoints without distortion, squared
// reprojection error in float, err <= (float)(thr*thr).
//
// Arithmetic contract (checked bit for bit against cv2 4.13.0's projectPoints, tests/golden/pnp_*.npz):
// doubles, no contraction, the rotation row summed left to right then + t, z = z ? 1./z : 1, u = x*fx + cx
// rounded to float; dx, dy, dx*dx + dy*dy in float.  Every operation below is an explicit _rn intrinsic (the
// library is built with --fmad=false as well); the division is the IEEE one.
//
// Shapes: a PROBLEM is one (query, candidate) key-frame pair = a run of correspondences
// [corr_begin, corr_begin + corr_count) of the object / image arrays.  A MODEL is one pose hypothesis
// (R row-major 9 doubles, t 3 doubles) tied to a problem.  K_score: one warp per model, lanes over the
// correspondences (L1/L2 keep a problem's 20 B/correspondence resident while its models are scored; the
// kernel is FP64-pipe bound: 23 DMUL/DADD + one division per correspondence and model).  K_mask: one CTA
// per problem writes the mask of the model RANSAC kept.  K_assemble: one CTA per pair compacts the
// matched query rows in ascending order (ballot prefix) and gathers their 3-D / 2-D points.
#include "slamgpu_internal.cuh"

namespace slamgpu {

namespace {

constexpr int kPnpThreads = 256;

struct PnpCam {
    double fx, fy, cx, cy;
    float thr2;
};

__device__ __forceinline__ float pnp_error(const double* __restrict__ rt, const PnpCam& cam, float Xf, float Yf,
                                           float Zf, float2 ip)
{
    const double X = Xf, Y = Yf, Z = Zf;
    double x = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(rt[0], X), __dmul_rn(rt[1], Y)), __dmul_rn(rt[2], Z)), rt[9]);
    double y = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(rt[3], X), __dmul_rn(rt[4], Y)), __dmul_rn(rt[5], Z)), rt[10]);
    double z = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(rt[6], X), __dmul_rn(rt[7], Y)), __dmul_rn(rt[8], Z)), rt[11]);
    z = (z != 0.0) ? __ddiv_rn(1.0, z) : 1.0;
    x = __dmul_rn(x, z);
    y = __dmul_rn(y, z);
    const float u = __double2float_rn(__dadd_rn(__dmul_rn(x, cam.fx), cam.cx));
    const float v = __double2float_rn(__dadd_rn(__dmul_rn(y, cam.fy), cam.cy));
    const float dx = __fsub_rn(ip.x, u), dy = __fsub_rn(ip.y, v);
    return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
}

// one warp per model
__global__ void __launch_bounds__(kPnpThreads)
pnp_score_kernel(const float* __restrict__ object, const float2* __restrict__ image,
                 const int32_t* __restrict__ corr_begin, const int32_t* __restrict__ corr_count,
                 const double* __restrict__ hyp_rt, const int32_t* __restrict__ hyp_problem, int n_hyp, PnpCam cam,
                 int32_t* __restrict__ out_count)
{
    const int lane = threadIdx.x & 31;
    const int h = blockIdx.x * (kPnpThreads / 32) + (threadIdx.x >> 5);
    if (h >= n_hyp) return;
    double rt[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) rt[k] = __ldg(hyp_rt + 12 * static_cast<size_t>(h) + k);
    const int p = __ldg(hyp_problem + h);
    const int begin = __ldg(corr_begin + p), n = __ldg(corr_count + p);
    const float* obj = object + 3 * static_cast<size_t>(begin);
    const float2* img = image + begin;
    int good = 0;
    int i = lane;
    for (; i + 32 < n; i += 64) {     // two correspondences in flight per lane
        const float X0 = __ldg(obj + 3 * i), Y0 = __ldg(obj + 3 * i + 1), Z0 = __ldg(obj + 3 * i + 2);
        const float X1 = __ldg(obj + 3 * (i + 32)), Y1 = __ldg(obj + 3 * (i + 32) + 1), Z1 = __ldg(obj + 3 * (i + 32) + 2);
        const float2 p0 = __ldg(img + i), p1 = __ldg(img + i + 32);
        good += pnp_error(rt, cam, X0, Y0, Z0, p0) <= cam.thr2;
        good += pnp_error(rt, cam, X1, Y1, Z1, p1) <= cam.thr2;
    }
    if (i < n) {
        const float X0 = __ldg(obj + 3 * i), Y0 = __ldg(obj + 3 * i + 1), Z0 = __ldg(obj + 3 * i + 2);
        good += pnp_error(rt, cam, X0, Y0, Z0, __ldg(img + i)) <= cam.thr2;
    }
    good = __reduce_add_sync(0xFFFFFFFFu, good);
    if (lane == 0) out_count[h] = good;
}

// one CTA per problem: mask (and errors, when asked) of the model sel_hyp[p]; -1 = no model: zeros
__global__ void __launch_bounds__(kPnpThreads)
pnp_mask_kernel(const float* __restrict__ object, const float2* __restrict__ image,
                const int32_t* __restrict__ corr_begin, const int32_t* __restrict__ corr_count,
                const double* __restrict__ hyp_rt, const int32_t* __restrict__ sel_hyp, PnpCam cam,
                uint8_t* __restrict__ out_mask, float* __restrict__ out_err, int32_t* __restrict__ out_inliers)
{
    __shared__ double s_rt[12];
    __shared__ int s_good;
    const int p = blockIdx.x;
    const int h = sel_hyp[p];
    const int begin = corr_begin[p], n = corr_count[p];
    if (threadIdx.x < 12 && h >= 0) s_rt[threadIdx.x] = hyp_rt[12 * static_cast<size_t>(h) + threadIdx.x];
    if (threadIdx.x == 0) s_good = 0;
    __syncthreads();
    int good = 0;
    for (int i = threadIdx.x; i < n; i += kPnpThreads) {
        const size_t g = static_cast<size_t>(begin) + i;
        int f = 0;
        float e = 0.f;
        if (h >= 0) {
            e = pnp_error(s_rt, cam, object[3 * g], object[3 * g + 1], object[3 * g + 2], image[g]);
            f = e <= cam.thr2;
        }
        out_mask[g] = static_cast<uint8_t>(f);
        if (out_err) out_err[g] = e;
        good += f;
    }
    good = __reduce_add_sync(0xFFFFFFFFu, good);
    if ((threadIdx.x & 31) == 0 && good) atomicAdd(&s_good, good);
    __syncthreads();
    if (threadIdx.x == 0 && out_inliers) out_inliers[p] = s_good;
}

// one CTA per pair: correspondences = the query rows with a surviving match, ascending (the order
// match_descriptors emits them, src/MapMatcher.cpp:150-161); object = position of the train row's map point,
// image = the query key point (src/LoopDetector.cpp:192-196).
__global__ void __launch_bounds__(kPnpThreads)
pnp_assemble_kernel(const int32_t* __restrict__ match_train, const int32_t* __restrict__ frame_offset,
                    const int32_t* __restrict__ pairs, int max_rows, const float* __restrict__ row_pos,
                    const float2* __restrict__ row_xy, float* __restrict__ out_object, float2* __restrict__ out_image,
                    int32_t* __restrict__ out_query, int32_t* __restrict__ out_count)
{
    __shared__ int s_warp[kPnpThreads / 32];
    __shared__ int s_base;
    const int p = blockIdx.x;
    const int qf = pairs[2 * p], tf = pairs[2 * p + 1];
    const int q0 = frame_offset[qf], nq = frame_offset[qf + 1] - q0, t0 = frame_offset[tf];
    const int32_t* mt = match_train + static_cast<size_t>(p) * max_rows;
    const size_t out0 = static_cast<size_t>(p) * max_rows;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    for (int c = 0; c < nq; c += kPnpThreads) {
        const int q = c + threadIdx.x;
        const int t = q < nq ? mt[q] : -1;
        const unsigned b = __ballot_sync(0xFFFFFFFFu, t >= 0);
        if (lane == 0) s_warp[warp] = __popc(b);
        __syncthreads();
        int before = s_base;
        for (int w = 0; w < warp; ++w) before += s_warp[w];
        if (t >= 0) {
            const size_t o = out0 + before + __popc(b & ((1u << lane) - 1u));
            const size_t tr = static_cast<size_t>(t0) + t;
            out_object[3 * o] = row_pos[3 * tr];
            out_object[3 * o + 1] = row_pos[3 * tr + 1];
            out_object[3 * o + 2] = row_pos[3 * tr + 2];
            out_image[o] = row_xy[static_cast<size_t>(q0) + q];
            if (out_query) out_query[o] = q;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0;
            for (int w = 0; w < kPnpThreads / 32; ++w) tot += s_warp[w];
            s_base += tot;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) out_count[p] = s_base;
}

PnpCam make_cam(const double* K4, double reproj_error)
{
    PnpCam c;
    c.fx = K4[0];
    c.fy = K4[1];
    c.cx = K4[2];
    c.cy = K4[3];
    c.thr2 = static_cast<float>(reproj_error * reproj_error);   // ptsetreg.cpp findInliers
    return c;
}

} // namespace

int pnp_launch_score(slamgpu_ctx* ctx, const float* d_object, const float* d_image, const int32_t* d_corr_begin,
                     const int32_t* d_corr_count, const double* d_hyp_rt, const int32_t* d_hyp_problem, int n_hyp,
                     const double* K4, double reproj_error, int32_t* d_out_count)
{
    if (n_hyp <= 0) return SLAMGPU_OK;
    ProfScope prof(ctx, SLAMGPU_KERNEL_PNP_SCORE);
    const int per_cta = kPnpThreads / 32;
    pnp_score_kernel<<<(n_hyp + per_cta - 1) / per_cta, kPnpThreads, 0, ctx->stream>>>(
        d_object, reinterpret_cast<const float2*>(d_image), d_corr_begin, d_corr_count, d_hyp_rt, d_hyp_problem, n_hyp,
        make_cam(K4, reproj_error), d_out_count);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int pnp_launch_mask(slamgpu_ctx* ctx, const float* d_object, const float* d_image, const int32_t* d_corr_begin,
                    const int32_t* d_corr_count, int n_problems, const double* d_hyp_rt, const int32_t* d_sel_hyp,
                    const double* K4, double reproj_error, uint8_t* d_out_mask, float* d_out_err,
                    int32_t* d_out_inliers)
{
    if (n_problems <= 0) return SLAMGPU_OK;
    pnp_mask_kernel<<<n_problems, kPnpThreads, 0, ctx->stream>>>(
        d_object, reinterpret_cast<const float2*>(d_image), d_corr_begin, d_corr_count, d_hyp_rt, d_sel_hyp,
        make_cam(K4, reproj_error), d_out_mask, d_out_err, d_out_inliers);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int pnp_launch_assemble(slamgpu_ctx* ctx, const int32_t* d_match_train, const int32_t* d_frame_offset,
                        const int32_t* d_pairs, int n_pairs, int max_rows, const float* d_row_pos,
                        const float* d_row_xy, float* d_out_object, float* d_out_image, int32_t* d_out_query,
                        int32_t* d_out_count)
{
    if (n_pairs <= 0) return SLAMGPU_OK;
    pnp_assemble_kernel<<<n_pairs, kPnpThreads, 0, ctx->stream>>>(
        d_match_train, d_frame_offset, d_pairs, max_rows, d_row_pos, reinterpret_cast<const float2*>(d_row_xy),
        d_out_object, reinterpret_cast<float2*>(d_out_image), d_out_query, d_out_count);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

} // namespace slamgpu
