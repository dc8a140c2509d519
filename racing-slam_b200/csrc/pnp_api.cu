// pnp_api.cu -- PnP verification entry points of the C ABI (include/slamgpu.h): input assembly of
// verify_pnp (src/LoopDetector.cpp:188-197), inlier counts of every RANSAC model, RANSAC's model choice and the
// kept model's mask, as cv::solvePnPRansac (src/LoopDetector.cpp:203-214) computes them.  Kernels are in
// pnp_kernels.cu.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>

#include "slamgpu_internal.cuh"

using namespace slamgpu;

namespace {

#define SG_CTX(ctx)                                 \
    do {                                            \
        SG_CHECK((ctx) != nullptr, "null context"); \
        SG_CUDA(cudaSetDevice((ctx)->device));      \
    } while (0)

// OpenCV ptsetreg.cpp RANSACUpdateNumIters (host doubles; std::pow / std::log are what OpenCV calls).
int update_num_iters(double p, double ep, int model_points, int max_iters)
{
    p = std::min(std::max(p, 0.), 1.);
    ep = std::min(std::max(ep, 0.), 1.);
    double num = std::max(1. - p, DBL_MIN);
    double denom = 1. - std::pow(1. - ep, model_points);
    if (denom < DBL_MIN) return 0;
    num = std::log(num);
    denom = std::log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : static_cast<int>(std::lrint(num / denom));   // cvRound
}

int select_model(const int32_t* good_count, int n_models, int n_points, int model_points, double confidence,
                 int max_iters, int* out_iters)
{
    int niters = std::max(max_iters, 1), best = -1, max_good = 0, iter = 0;
    for (; iter < niters && iter < n_models; ++iter) {
        const int good = good_count[iter];
        if (good > std::max(max_good, model_points - 1)) {
            best = iter;
            max_good = good;
            niters = update_num_iters(confidence, static_cast<double>(n_points - good) / n_points, model_points, niters);
        }
    }
    if (out_iters) *out_iters = iter;
    return best;
}

bool finite4(const double* K4) { return std::isfinite(K4[0]) && std::isfinite(K4[1]) && std::isfinite(K4[2]) && std::isfinite(K4[3]); }

} // namespace

extern "C" {

int slamgpu_pnp_assemble_batch_dev(slamgpu_ctx* ctx, const int32_t* d_match_train, const int32_t* d_frame_offset,
                                   int n_frames, const int32_t* d_pairs, int n_pairs, int max_rows,
                                   const float* d_row_pos, const float* d_row_xy, float* d_out_object,
                                   float* d_out_image, int32_t* d_out_query, int32_t* d_out_count)
{
    SG_CTX(ctx);
    SG_CHECK(n_frames >= 0 && n_pairs >= 0 && max_rows >= 0, "negative size");
    if (n_pairs == 0) return SLAMGPU_OK;
    SG_CHECK(d_match_train && d_frame_offset && d_pairs && d_row_pos && d_row_xy && d_out_object && d_out_image && d_out_count,
             "null device pointer");
    SG_CHECK((reinterpret_cast<uintptr_t>(d_row_xy) & 7) == 0 && (reinterpret_cast<uintptr_t>(d_out_image) & 7) == 0,
             "d_row_xy / d_out_image must be 8-byte aligned");
    return pnp_launch_assemble(ctx, d_match_train, d_frame_offset, d_pairs, n_pairs, max_rows, d_row_pos, d_row_xy,
                               d_out_object, d_out_image, d_out_query, d_out_count);
}

int slamgpu_pnp_score_dev(slamgpu_ctx* ctx, const float* d_object, const float* d_image, const int32_t* d_corr_begin,
                          const int32_t* d_corr_count, int n_problems, const double* d_hyp_rt,
                          const int32_t* d_hyp_problem, int n_hyp, const double* K4, double reproj_error,
                          int32_t* d_out_count)
{
    SG_CTX(ctx);
    SG_CHECK(n_problems >= 0 && n_hyp >= 0, "negative size");
    SG_CHECK(K4 != nullptr && finite4(K4), "K4 = fx, fy, cx, cy must be finite");
    SG_CHECK(reproj_error >= 0.0, "negative reprojection error");
    if (n_hyp == 0) return SLAMGPU_OK;
    SG_CHECK(n_problems > 0, "models without a problem");
    SG_CHECK(d_object && d_image && d_corr_begin && d_corr_count && d_hyp_rt && d_hyp_problem && d_out_count,
             "null device pointer");
    SG_CHECK((reinterpret_cast<uintptr_t>(d_image) & 7) == 0, "d_image must be 8-byte aligned");
    return pnp_launch_score(ctx, d_object, d_image, d_corr_begin, d_corr_count, d_hyp_rt, d_hyp_problem, n_hyp, K4,
                            reproj_error, d_out_count);
}

int slamgpu_pnp_mask_dev(slamgpu_ctx* ctx, const float* d_object, const float* d_image, const int32_t* d_corr_begin,
                         const int32_t* d_corr_count, int n_problems, const double* d_hyp_rt, const int32_t* d_sel_hyp,
                         const double* K4, double reproj_error, uint8_t* d_out_mask, float* d_out_err,
                         int32_t* d_out_inliers)
{
    SG_CTX(ctx);
    SG_CHECK(n_problems >= 0, "negative size");
    SG_CHECK(K4 != nullptr && finite4(K4), "K4 = fx, fy, cx, cy must be finite");
    SG_CHECK(reproj_error >= 0.0, "negative reprojection error");
    if (n_problems == 0) return SLAMGPU_OK;
    SG_CHECK(d_object && d_image && d_corr_begin && d_corr_count && d_hyp_rt && d_sel_hyp && d_out_mask,
             "null device pointer");
    SG_CHECK((reinterpret_cast<uintptr_t>(d_image) & 7) == 0, "d_image must be 8-byte aligned");
    return pnp_launch_mask(ctx, d_object, d_image, d_corr_begin, d_corr_count, n_problems, d_hyp_rt, d_sel_hyp, K4,
                           reproj_error, d_out_mask, d_out_err, d_out_inliers);
}

int slamgpu_ransac_select(const int32_t* good_count, int n_models, int n_points, int model_points, double confidence,
                          int max_iters, int* out_best, int* out_iters)
{
    SG_CHECK(n_models >= 0 && n_points >= 0 && model_points >= 1, "bad sizes");
    SG_CHECK(out_best != nullptr, "null out_best");
    SG_CHECK(n_models == 0 || good_count != nullptr, "null good_count");
    if (n_points == 0) {     // no correspondences: nothing can be kept
        *out_best = -1;
        if (out_iters) *out_iters = 0;
        return SLAMGPU_OK;
    }
    *out_best = select_model(good_count, n_models, n_points, model_points, confidence, max_iters, out_iters);
    return SLAMGPU_OK;
}

int slamgpu_pnp_ransac_score(slamgpu_ctx* ctx, const float* object, const float* image, const int32_t* corr_offset,
                             int n_problems, const double* hyp_rt, const int32_t* hyp_offset, const double* K4,
                             double reproj_error, double confidence, int model_points, int max_iters,
                             int32_t* out_count, int32_t* out_best, int32_t* out_inliers, uint8_t* out_mask)
{
    SG_CTX(ctx);
    SG_CHECK(n_problems >= 0, "negative problem count");
    SG_CHECK(model_points >= 1, "model_points must be positive");
    SG_CHECK(K4 != nullptr && finite4(K4), "K4 = fx, fy, cx, cy must be finite");
    SG_CHECK(reproj_error >= 0.0, "negative reprojection error");
    if (n_problems == 0) return SLAMGPU_OK;
    SG_CHECK(corr_offset && hyp_offset && out_best && out_inliers, "null pointer");
    SG_CHECK(corr_offset[0] == 0 && hyp_offset[0] == 0, "offsets must start at 0");
    for (int p = 0; p < n_problems; ++p)
        SG_CHECK(corr_offset[p + 1] >= corr_offset[p] && hyp_offset[p + 1] >= hyp_offset[p], "offsets must be non-decreasing");
    const size_t NC = static_cast<size_t>(corr_offset[n_problems]), NH = static_cast<size_t>(hyp_offset[n_problems]);
    const size_t NP = static_cast<size_t>(n_problems);
    SG_CHECK(NC == 0 || (object && image), "null correspondences");
    SG_CHECK(NH == 0 || (hyp_rt && out_count), "null models");

    // device: [object][image][begin][count][hyp_rt][hyp_problem][sel][out_count][inliers][mask]
    const size_t b_obj = Carver::padded(NC * 12), b_img = Carver::padded(NC * 8), b_p = Carver::padded(NP * 4);
    const size_t b_rt = Carver::padded(NH * 96), b_h = Carver::padded(NH * 4), b_mask = Carver::padded(NC);
    SG_TRY(ensure_device_io(ctx, b_obj + b_img + 4 * b_p + b_rt + 2 * b_h + b_mask + 256));
    const size_t up_bytes = b_obj + b_img + 2 * b_p + b_rt + b_h;
    SG_TRY(ensure_host_stage(ctx, std::max(up_bytes, 2 * b_p + b_mask) + 256));
    Carver dio(ctx->d_io.ptr);
    float* d_obj = dio.take<float>(NC * 3);
    float* d_img = dio.take<float>(NC * 2);
    int32_t* d_begin = dio.take<int32_t>(NP);
    int32_t* d_cnt = dio.take<int32_t>(NP);
    double* d_rt = dio.take<double>(NH * 12);
    int32_t* d_hp = dio.take<int32_t>(NH);
    int32_t* d_sel = dio.take<int32_t>(NP);
    int32_t* d_good = dio.take<int32_t>(NH);
    int32_t* d_inl = dio.take<int32_t>(NP);
    uint8_t* d_mask = dio.take<uint8_t>(NC);

    // one staged upload: the device layout up to hyp_problem is mirrored in the pinned buffer
    Carver hs(ctx->h_stage.ptr);
    float* h_obj = hs.take<float>(NC * 3);
    float* h_img = hs.take<float>(NC * 2);
    int32_t* h_begin = hs.take<int32_t>(NP);
    int32_t* h_cnt = hs.take<int32_t>(NP);
    double* h_rt = hs.take<double>(NH * 12);
    int32_t* h_hp = hs.take<int32_t>(NH);
    if (NC) {
        std::memcpy(h_obj, object, NC * 12);
        std::memcpy(h_img, image, NC * 8);
    }
    if (NH) std::memcpy(h_rt, hyp_rt, NH * 96);
    for (int p = 0; p < n_problems; ++p) {
        h_begin[p] = corr_offset[p];
        h_cnt[p] = corr_offset[p + 1] - corr_offset[p];
        for (int h = hyp_offset[p]; h < hyp_offset[p + 1]; ++h) h_hp[h] = p;
    }
    SG_CUDA(cudaMemcpyAsync(d_obj, ctx->h_stage.ptr, up_bytes, cudaMemcpyHostToDevice, ctx->stream));
    SG_TRY(pnp_launch_score(ctx, d_obj, d_img, d_begin, d_cnt, d_rt, d_hp, static_cast<int>(NH), K4, reproj_error, d_good));
    if (NH) {
        SG_CUDA(cudaMemcpyAsync(out_count, d_good, NH * 4, cudaMemcpyDeviceToHost, ctx->stream));
    }
    SG_CUDA(cudaStreamSynchronize(ctx->stream));

    // RANSAC's choice per problem (host doubles, as OpenCV), then the kept models' masks
    int32_t* h_sel = static_cast<int32_t*>(ctx->h_stage.ptr);
    for (int p = 0; p < n_problems; ++p) {
        const int h0 = hyp_offset[p], nh = hyp_offset[p + 1] - h0, nc = corr_offset[p + 1] - corr_offset[p];
        const int best = nc > 0 ? select_model(out_count + h0, nh, nc, model_points, confidence, max_iters, nullptr) : -1;
        out_best[p] = best;
        h_sel[p] = best < 0 ? -1 : h0 + best;
    }
    SG_CUDA(cudaMemcpyAsync(d_sel, h_sel, NP * 4, cudaMemcpyHostToDevice, ctx->stream));
    SG_TRY(pnp_launch_mask(ctx, d_obj, d_img, d_begin, d_cnt, n_problems, d_rt, d_sel, K4, reproj_error, d_mask, nullptr, d_inl));
    uint8_t* h_back = static_cast<uint8_t*>(ctx->h_stage.ptr) + Carver::padded(NP * 4);
    SG_CUDA(cudaMemcpyAsync(h_back, d_inl, NP * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_mask && NC) SG_CUDA(cudaMemcpyAsync(h_back + b_p, d_mask, NC, cudaMemcpyDeviceToHost, ctx->stream));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    std::memcpy(out_inliers, h_back, NP * 4);
    if (out_mask && NC) std::memcpy(out_mask, h_back + b_p, NC);
    return SLAMGPU_OK;
}

} // extern "C"
