// bf_api.cu -- brute-force entry points of the C ABI (include/slamgpu.h).
// Host-buffer forms stage through the context's pinned buffer, copy H2D, run the kernels of
// bf_kernels.cu on the context's stream, copy D2H and return when the caller's buffers are filled.
#include <algorithm>
#include <cstring>

#include "slamgpu_internal.cuh"

using namespace slamgpu;

namespace {

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

#define SG_CTX(ctx)                                 \
    do {                                            \
        SG_CHECK((ctx) != nullptr, "null context"); \
        SG_CUDA(cudaSetDevice((ctx)->device));      \
    } while (0)

int check_sets(const void* q, int nq, const void* t, int nt)
{
    SG_CHECK(nq >= 0 && nt >= 0, "negative descriptor count");
    SG_CHECK(nt <= SLAMGPU_MAX_TRAIN && nq <= SLAMGPU_MAX_TRAIN, "descriptor set larger than SLAMGPU_MAX_TRAIN");
    SG_CHECK(nq == 0 || q != nullptr, "null query descriptors");
    SG_CHECK(nt == 0 || t != nullptr, "null train descriptors");
    return SLAMGPU_OK;
}

int check_dev_sets(const void* q, int nq, const void* t, int nt)
{
    SG_TRY(check_sets(q, nq, t, nt));
    SG_CHECK(aligned16(q) && aligned16(t), "device descriptor pointers must be 16-byte aligned");
    return SLAMGPU_OK;
}

// kNN-2 over device buffers; scratch from d_ws.
int knn2_core(slamgpu_ctx* ctx, const uint8_t* dq, int nq, const uint8_t* dt, int nt, int32_t* d_idx, int32_t* d_dist)
{
    if (nq == 0) return SLAMGPU_OK;
    if (nt == 0) {   // every neighbour missing
        SG_CUDA(cudaMemsetAsync(d_idx, 0xFF, sizeof(int32_t) * 2 * nq, ctx->stream));
        SG_CUDA(cudaMemsetAsync(d_dist, 0xFF, sizeof(int32_t) * 2 * nq, ctx->stream));
        return SLAMGPU_OK;
    }
    const BfPlan plan = bf_make_plan(ctx->n_sm, nq, nt, 1);
    SG_TRY(ensure_device_ws(ctx, Carver::padded(bf_partial_elems(plan, 1, nq) * sizeof(uint2))));
    Carver ws(ctx->d_ws.ptr);
    uint2* keys = ws.take<uint2>(bf_partial_elems(plan, 1, nq));
    SG_TRY(bf_launch_scan(ctx, plan, dq, nq, dt, nt, nullptr, nullptr, 1, nq, false, keys));
    return bf_launch_finish_raw(ctx, plan, keys, nq, nt, nq, d_idx, d_dist);
}

// cross-check over device buffers.  strict_max_dist < 0: raw; otherwise also distance < strict_max_dist.
int crosscheck_core(slamgpu_ctx* ctx, const uint8_t* dq, int nq, const uint8_t* dt, int nt, float strict_max_dist,
                    int32_t* d_idx, int32_t* d_dist)
{
    if (nq == 0) return SLAMGPU_OK;
    if (nt == 0) {
        SG_CUDA(cudaMemsetAsync(d_idx, 0xFF, sizeof(int32_t) * nq, ctx->stream));
        SG_CUDA(cudaMemsetAsync(d_dist, 0xFF, sizeof(int32_t) * nq, ctx->stream));
        return SLAMGPU_OK;
    }
    const BfPlan row = bf_make_plan(ctx->n_sm, nq, nt, 1);
    const BfPlan col = bf_make_plan(ctx->n_sm, nt, nq, 1);
    const size_t n_row = bf_partial_elems(row, 1, nq), n_col = bf_partial_elems(col, 1, nt);
    SG_TRY(ensure_device_ws(ctx, Carver::padded(n_row * sizeof(uint2)) + Carver::padded(n_col * sizeof(uint2))));
    Carver ws(ctx->d_ws.ptr);
    uint2* row_keys = ws.take<uint2>(n_row);
    uint2* col_keys = ws.take<uint2>(n_col);
    SG_TRY(bf_launch_scan(ctx, row, dq, nq, dt, nt, nullptr, nullptr, 1, nq, false, row_keys));
    SG_TRY(bf_launch_scan(ctx, col, dt, nt, dq, nq, nullptr, nullptr, 1, nt, false, col_keys));
    return bf_launch_mutual(ctx, row, row_keys, nq, col, col_keys, nt, nq, nt, strict_max_dist, d_idx, d_dist);
}

int match_descriptors_core(slamgpu_ctx* ctx, const uint8_t* dq, int nq, const uint8_t* dt, int nt, float max_dist,
                           float ratio, int32_t* d_out_train, int32_t* d_out_count)
{
    SG_CUDA(cudaMemsetAsync(d_out_count, 0, sizeof(int32_t), ctx->stream));
    if (nq == 0) return SLAMGPU_OK;
    if (nt == 0) {   // src/MapMatcher.cpp:139-141: no matched key-frame keypoints -> no matches
        SG_CUDA(cudaMemsetAsync(d_out_train, 0xFF, sizeof(int32_t) * nq, ctx->stream));
        return SLAMGPU_OK;
    }
    const BfPlan plan = bf_make_plan(ctx->n_sm, nq, nt, 1);
    SG_TRY(ensure_device_ws(ctx, Carver::padded(bf_partial_elems(plan, 1, nq) * sizeof(uint2))));
    Carver ws(ctx->d_ws.ptr);
    uint2* keys = ws.take<uint2>(bf_partial_elems(plan, 1, nq));
    SG_TRY(bf_launch_scan(ctx, plan, dq, nq, dt, nt, nullptr, nullptr, 1, nq, false, keys));
    return bf_launch_finish_filter(ctx, plan, keys, nq, nt, nullptr, nullptr, 1, nq, max_dist, ratio, d_out_train,
                                   d_out_count);
}

// Pairs are processed in slices so that the partial-key scratch stays around L2 size.
constexpr size_t kBatchScratchBytes = size_t(64) << 20;

int batch_core(slamgpu_ctx* ctx, const uint8_t* d_desc, const int32_t* d_frame_offset, const int32_t* d_pairs,
               int n_pairs, int max_rows, float max_dist, float ratio, int32_t* d_out_train, int32_t* d_out_count,
               // optional ordered compaction (host-buffer form): per-slice out_train is then a scratch
               // of slice_pairs x max_rows and the lists are filled slice by slice
               int32_t* d_offset, int32_t* d_query_list, int32_t* d_train_list, int64_t capacity, int slice_pairs)
{
    if (n_pairs == 0) return SLAMGPU_OK;
    if (max_rows == 0) {
        SG_CUDA(cudaMemsetAsync(d_out_count, 0, sizeof(int32_t) * n_pairs, ctx->stream));
        if (d_offset) SG_CUDA(cudaMemsetAsync(d_offset, 0, sizeof(int32_t) * (n_pairs + 1), ctx->stream));
        return SLAMGPU_OK;
    }
    const BfPlan plan = bf_make_plan(ctx->n_sm, max_rows, max_rows, std::min(n_pairs, slice_pairs));
    const size_t per_pair = bf_partial_elems(plan, 1, max_rows);
    SG_TRY(ensure_device_ws(ctx, Carver::padded(per_pair * sizeof(uint2) * std::min(n_pairs, slice_pairs))));
    Carver ws(ctx->d_ws.ptr);
    uint2* keys = ws.take<uint2>(per_pair * std::min(n_pairs, slice_pairs));
    if (d_offset) SG_CUDA(cudaMemsetAsync(d_offset, 0, sizeof(int32_t), ctx->stream));
    for (int p0 = 0; p0 < n_pairs; p0 += slice_pairs) {
        const int np = std::min(slice_pairs, n_pairs - p0);
        const int32_t* pairs = d_pairs + 2 * static_cast<size_t>(p0);
        int32_t* out_train = d_offset ? d_out_train : d_out_train + static_cast<size_t>(p0) * max_rows;
        SG_TRY(bf_launch_scan(ctx, plan, d_desc, 0, d_desc, 0, d_frame_offset, pairs, np, max_rows, false, keys));
        SG_TRY(bf_launch_finish_filter(ctx, plan, keys, 0, 0, d_frame_offset, pairs, np, max_rows, max_dist, ratio,
                                       out_train, d_out_count + p0));
        if (d_offset) {
            SG_TRY(bf_launch_scan_counts(ctx, d_out_count + p0, np, d_offset + p0));
            SG_TRY(bf_launch_compact(ctx, out_train, d_frame_offset, pairs, 0, np, max_rows, d_offset + p0,
                                     d_query_list, d_train_list, capacity));
        }
    }
    return SLAMGPU_OK;
}

int slice_for(int n_sm, int max_rows)
{
    // one slice = enough pairs to fill the GPU a few times over, bounded by the scratch budget
    const size_t per_pair = static_cast<size_t>(max_rows) * sizeof(uint2);
    size_t s = per_pair ? kBatchScratchBytes / per_pair : 1;
    s = std::max<size_t>(s, static_cast<size_t>(n_sm) * 2);
    return static_cast<int>(std::min<size_t>(s, 1 << 20));
}

} // namespace

extern "C" {

int slamgpu_knn2_dev(slamgpu_ctx* ctx, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                     int32_t* d_out_idx, int32_t* d_out_dist)
{
    SG_CTX(ctx);
    SG_TRY(check_dev_sets(d_query, nq, d_train, nt));
    SG_CHECK(nq == 0 || (d_out_idx && d_out_dist), "null output");
    SG_CHECK((reinterpret_cast<uintptr_t>(d_out_idx) & 7u) == 0 && (reinterpret_cast<uintptr_t>(d_out_dist) & 7u) == 0,
             "kNN outputs must be 8-byte aligned");
    return knn2_core(ctx, d_query, nq, d_train, nt, d_out_idx, d_out_dist);
}

int slamgpu_crosscheck_dev(slamgpu_ctx* ctx, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                           int32_t* d_out_idx, int32_t* d_out_dist)
{
    SG_CTX(ctx);
    SG_TRY(check_dev_sets(d_query, nq, d_train, nt));
    SG_CHECK(nq == 0 || (d_out_idx && d_out_dist), "null output");
    return crosscheck_core(ctx, d_query, nq, d_train, nt, -1.0f, d_out_idx, d_out_dist);
}

int slamgpu_match_descriptors_dev(slamgpu_ctx* ctx, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                                  float max_dist, float ratio, int32_t* d_out_train, int32_t* d_out_count)
{
    SG_CTX(ctx);
    SG_TRY(check_dev_sets(d_query, nq, d_train, nt));
    SG_CHECK(d_out_count != nullptr && (nq == 0 || d_out_train != nullptr), "null output");
    return match_descriptors_core(ctx, d_query, nq, d_train, nt, max_dist, ratio, d_out_train, d_out_count);
}

int slamgpu_match_descriptors_batch_dev(slamgpu_ctx* ctx, const uint8_t* d_desc, const int32_t* d_frame_offset,
                                        int n_frames, const int32_t* d_pairs, int n_pairs, int max_rows,
                                        float max_dist, float ratio, int32_t* d_out_train, int32_t* d_out_count)
{
    SG_CTX(ctx);
    SG_CHECK(n_frames >= 0 && n_pairs >= 0 && max_rows >= 0, "negative size");
    SG_CHECK(max_rows <= SLAMGPU_MAX_TRAIN, "frame larger than SLAMGPU_MAX_TRAIN");
    if (n_pairs == 0) return SLAMGPU_OK;
    SG_CHECK(d_desc && d_frame_offset && d_pairs && d_out_train && d_out_count, "null pointer");
    SG_CHECK(aligned16(d_desc), "device descriptor pointer must be 16-byte aligned");
    return batch_core(ctx, d_desc, d_frame_offset, d_pairs, n_pairs, max_rows, max_dist, ratio, d_out_train,
                      d_out_count, nullptr, nullptr, nullptr, 0, slice_for(ctx->n_sm, max_rows));
}

} // extern "C"
