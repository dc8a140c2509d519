// bf_api.cu -- brute-force entry points of the C ABI (include/slamgpu.h).
// Host-buffer forms stage through the context's pinned buffer, copy H2D, run the kernels of
// bf_kernels.cu on the context's stream, copy D2H and return when the caller's buffers are filled.
#include <algorithm>
#include <chrono>
#include <functional>
#include <cmath>
#include <cstring>

#include "slamgpu_internal.cuh"

using namespace slamgpu;

namespace {

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Measured on the B200 (tools/tc_sizes.py, profiles/r01_tc_sizes.txt): the tensor-core path has a ~22 us floor
// (expansion + scan + finish launches) and overtakes the integer scan near 2500 x 2500 for one scan and near
// 3500 x 3500 for the two scans of a cross-check.
constexpr int64_t kTcMinComparisons = 6000000;
constexpr int64_t kTcMinComparisonsCross = 12000000;

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

// ---- engine choice ---------------------------------------------------------------------------
// The tensor-core scan (bf_tc_kernels.cu) pays an expansion pass and 128 x 256 tile granularity;
// the integer scan (bf_kernels.cu) wins on small sets.  SLAMGPU_BF_ENGINE=int|tc forces one.
bool use_tc(const slamgpu_ctx* ctx, int64_t nq, int64_t nt, int64_t n_pairs)
{
    if (nq <= 0 || nt <= 0 || n_pairs <= 0) return false;
    if (ctx->bf_engine == 1) return false;
    if (ctx->bf_engine == 2) return true;
    return nq * nt * n_pairs >= kTcMinComparisons && nt >= 512;
}

// Distance cuts of the tensor-core scan's filtered mode (bf_tc_kernels.cu): train rows at a distance
// >= cut cannot change the caller's result.
//   match_descriptors (src/MapMatcher.cpp:151-161): a row survives only with d0 <= max_dist, and then
//   d0 > ratio * d1 can only hold for d1 < max_dist / ratio.  Margin of 3 for the float compares.
//   match_features (src/features/FeatureExtractor.cpp:25): only mutual nearest pairs with d < max_dist.
constexpr int kNoCut = 257;
int cut_for_filter(float max_dist, float ratio)
{
    if (!(max_dist >= 0.0f) || !(ratio > 0.0f) || !std::isfinite(max_dist) || !std::isfinite(ratio)) return kNoCut;
    const double t = std::max(static_cast<double>(max_dist), static_cast<double>(max_dist) / ratio);
    return t >= 254.0 ? kNoCut : static_cast<int>(std::floor(t)) + 3;
}
int cut_for_crosscheck(float strict_max_dist)
{
    if (!(strict_max_dist >= 0.0f) || !std::isfinite(strict_max_dist)) return kNoCut;   // < 0: raw cross-check
    return strict_max_dist >= 254.0f ? kNoCut : static_cast<int>(std::ceil(strict_max_dist)) + 2;
}

// One (query, train) pair on the tensor cores: expand both sets, scan, leave 2 * splits partials.
struct TcPair {
    BfPlan plan;      // of the scan
    BfPlan fin;       // what the finish kernels see (bf_tc_partials() partials per split)
    uint2* keys = nullptr;
};
size_t tc_pair_keys(const slamgpu_ctx* ctx, const BfPlan& plan, int nq)
{
    return static_cast<size_t>(plan.splits) * bf_tc_partials(ctx) * nq;
}

// Filtered calls (a distance cut exists) run on the block-scaled FP4 kernel (bf_tc4_kernels.cu), exact
// kNN-2 on the int8 kernel (bf_tc_kernels.cu).  In -DSLAMGPU_DEBUG_VARIANTS builds SLAMGPU_TC_VARIANT bit 6 keeps filtered calls on int8 (A/B).
bool use_fp4(const slamgpu_ctx* ctx, int d_cut) { return d_cut <= 250 && !SG_TC_VARIANT(ctx->tc_variant, 64); }

BfPlan tc_plan(const slamgpu_ctx* ctx, int nq, int nt, int64_t n_pairs, bool fp4)
{
    return fp4 ? bf_tc4_make_plan(ctx, nq, nt, n_pairs) : bf_tc_make_plan(ctx, nq, nt, n_pairs);
}

int tc_expand_pair(slamgpu_ctx* ctx, const uint8_t* dq, int nq, const uint8_t* dt, int nt, bool fp4, int* pitch_rows)
{
    const int m = std::max(nq, nt);
    const int pitch = fp4 ? bf_tc4_pitch_rows(m) : bf_tc_pitch_rows(m);
    const size_t set_bytes = fp4 ? bf_tc4_expanded_bytes(1, m) : bf_tc_expanded_bytes(1, m);
    SG_TRY(ensure_device_exp(ctx, 2 * set_bytes));
    uint8_t* e = static_cast<uint8_t*>(ctx->d_exp.ptr);
    if (fp4) SG_TRY(bf_tc4_launch_expand_pair(ctx, dq, nq, dt, nt, pitch, e));     // both sets, one launch
    else SG_TRY(bf_tc_launch_expand_pair(ctx, dq, nq, dt, nt, pitch, e));
    *pitch_rows = pitch;
    return SLAMGPU_OK;
}

// scan query-set `qi` against train-set `ti` (0 or 1: the two expanded sets of tc_expand_pair)
int tc_scan_pair(slamgpu_ctx* ctx, int qi, int nq, int ti, int nt, int pitch_rows, int d_cut, bool fp4, uint2* keys,
                 TcPair& out, uint32_t* col_keys = nullptr)
{
    out.plan = tc_plan(ctx, nq, nt, 1, fp4);
    out.fin = out.plan;
    out.fin.splits = out.plan.splits * bf_tc_partials(ctx);
    out.keys = keys;
    const uint8_t* e = static_cast<const uint8_t*>(ctx->d_exp.ptr);
    const size_t set_bytes = static_cast<size_t>(pitch_rows) * (fp4 ? 128 : 256);
    if (fp4) {
        return bf_tc4_launch_scan(ctx, out.plan, e + qi * set_bytes, nq, e + ti * set_bytes, nt, pitch_rows, nullptr,
                                  nullptr, 1, nq, d_cut, keys, nullptr, col_keys);
    }
    SG_CHECK(col_keys == nullptr, "column minima come from the fp4 scan only");
    return bf_tc_launch_scan(ctx, out.plan, e + qi * set_bytes, nq, e + ti * set_bytes, nt, pitch_rows, nullptr, nullptr,
                             1, nq, d_cut, keys);
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
    if (use_tc(ctx, nq, nt, 1)) {
        int pitch = 0;
        SG_TRY(tc_expand_pair(ctx, dq, nq, dt, nt, false, &pitch));
        const BfPlan tp = tc_plan(ctx, nq, nt, 1, false);
        SG_TRY(ensure_device_ws(ctx, Carver::padded(tc_pair_keys(ctx, tp, nq) * sizeof(uint2))));
        TcPair r;
        SG_TRY(tc_scan_pair(ctx, 0, nq, 1, nt, pitch, kNoCut, false, static_cast<uint2*>(ctx->d_ws.ptr), r));
        return bf_launch_finish_raw(ctx, r.fin, r.keys, nq, nt, nq, d_idx, d_dist);
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
    if (nq >= 512 && use_tc(ctx, nq, nt, 1) &&
        (ctx->bf_engine == 2 || static_cast<int64_t>(nq) * nt >= kTcMinComparisonsCross)) {
        // tensor cores: two scans over the same expanded sets, rows (q -> t) and columns (t -> q)
        const int cut = cut_for_crosscheck(strict_max_dist);
        const bool fp4 = use_fp4(ctx, cut);
        int pitch = 0;
        SG_TRY(tc_expand_pair(ctx, dq, nq, dt, nt, fp4, &pitch));
        if (fp4 && ctx->tc4_cross_one_scan) {
            // ONE scan: only rows within the cut can be mutual nearest neighbours, and the FP4 scan decodes exactly
            // those (its hits), so every hit also takes part in its train row's column minimum (ColOut, atomicMin);
            // half the tensor-core work of the row scan + column scan below
            const BfPlan rp = tc_plan(ctx, nq, nt, 1, true);
            SG_TRY(ensure_device_ws(ctx, Carver::padded(tc_pair_keys(ctx, rp, nq) * sizeof(uint2)) + Carver::padded(sizeof(uint32_t) * nt)));
            Carver ws(ctx->d_ws.ptr);
            uint2* row_keys = ws.take<uint2>(tc_pair_keys(ctx, rp, nq));
            uint32_t* col_keys = ws.take<uint32_t>(nt);
            SG_CUDA(cudaMemsetAsync(col_keys, 0xFF, sizeof(uint32_t) * nt, ctx->stream));
            TcPair r;
            SG_TRY(tc_scan_pair(ctx, 0, nq, 1, nt, pitch, cut, true, row_keys, r, col_keys));
            return bf_launch_mutual_direct(ctx, r.fin, r.keys, nq, col_keys, nq, strict_max_dist, d_idx, d_dist);
        }
        const BfPlan rp = tc_plan(ctx, nq, nt, 1, fp4), cp = tc_plan(ctx, nt, nq, 1, fp4);
        SG_TRY(ensure_device_ws(ctx, Carver::padded(tc_pair_keys(ctx, rp, nq) * sizeof(uint2)) +
                                         Carver::padded(tc_pair_keys(ctx, cp, nt) * sizeof(uint2))));
        Carver ws(ctx->d_ws.ptr);
        uint2* row_keys = ws.take<uint2>(tc_pair_keys(ctx, rp, nq));
        uint2* col_keys = ws.take<uint2>(tc_pair_keys(ctx, cp, nt));
        TcPair r, c;
        SG_TRY(tc_scan_pair(ctx, 0, nq, 1, nt, pitch, cut, fp4, row_keys, r));
        SG_TRY(tc_scan_pair(ctx, 1, nt, 0, nq, pitch, cut, fp4, col_keys, c));
        return bf_launch_mutual(ctx, r.fin, r.keys, nq, c.fin, c.keys, nt, nq, nt, strict_max_dist, d_idx, d_dist);
    }
    // one scan: per-row top-1 partials and, in the same pass, per-column minima (atomicMin)
    const BfPlan row = bf_make_plan(ctx->n_sm, nq, nt, 1);
    const size_t n_row = bf_partial_elems(row, 1, nq);
    SG_TRY(ensure_device_ws(ctx, Carver::padded(n_row * sizeof(uint2)) + Carver::padded(sizeof(uint32_t) * nt)));
    Carver ws(ctx->d_ws.ptr);
    uint2* row_keys = ws.take<uint2>(n_row);
    uint32_t* col_keys = ws.take<uint32_t>(nt);
    SG_CUDA(cudaMemsetAsync(col_keys, 0xFF, sizeof(uint32_t) * nt, ctx->stream));
    SG_TRY(bf_launch_scan(ctx, row, dq, nq, dt, nt, nullptr, nullptr, 1, nq, false, row_keys, col_keys));
    return bf_launch_mutual_direct(ctx, row, row_keys, nq, col_keys, nq, strict_max_dist, d_idx, d_dist);
}

int match_descriptors_core(slamgpu_ctx* ctx, const uint8_t* dq, int nq, const uint8_t* dt, int nt, float max_dist,
                           float ratio, int32_t* d_out_train, int32_t* d_out_count)
{
    if (nq == 0 || nt == 0) {
        if (d_out_count) SG_CUDA(cudaMemsetAsync(d_out_count, 0, sizeof(int32_t), ctx->stream));
    }
    if (nq == 0) return SLAMGPU_OK;
    if (nt == 0) {   // src/MapMatcher.cpp:139-141: no matched key-frame keypoints -> no matches
        SG_CUDA(cudaMemsetAsync(d_out_train, 0xFF, sizeof(int32_t) * nq, ctx->stream));
        return SLAMGPU_OK;
    }
    if (use_tc(ctx, nq, nt, 1)) {
        const int cut = cut_for_filter(max_dist, ratio);
        const bool fp4 = use_fp4(ctx, cut);
        int pitch = 0;
        SG_TRY(tc_expand_pair(ctx, dq, nq, dt, nt, fp4, &pitch));
        const BfPlan tp = tc_plan(ctx, nq, nt, 1, fp4);
        SG_TRY(ensure_device_ws(ctx, Carver::padded(tc_pair_keys(ctx, tp, nq) * sizeof(uint2))));
        TcPair r;
        SG_TRY(tc_scan_pair(ctx, 0, nq, 1, nt, pitch, cut, fp4, static_cast<uint2*>(ctx->d_ws.ptr), r));
        return bf_launch_finish_filter(ctx, r.fin, r.keys, nq, nt, nullptr, nullptr, 1, nq, max_dist, ratio, d_out_train,
                                       d_out_count);
    }
    const BfPlan plan = bf_make_plan(ctx->n_sm, nq, nt, 1);
    SG_TRY(ensure_device_ws(ctx, Carver::padded(bf_partial_elems(plan, 1, nq) * sizeof(uint2))));
    Carver ws(ctx->d_ws.ptr);
    uint2* keys = ws.take<uint2>(bf_partial_elems(plan, 1, nq));
    SG_TRY(bf_launch_scan(ctx, plan, dq, nq, dt, nt, nullptr, nullptr, 1, nq, false, keys));
    return bf_launch_finish_filter(ctx, plan, keys, nq, nt, nullptr, nullptr, 1, nq, max_dist, ratio, d_out_train,
                                   d_out_count);
}

// fn(first frame, frame count) over all frames, or over the runs of consecutive frames of a sorted list
template <class F> int for_frame_runs(int n_frames, const std::vector<int>* used, F fn)
{
    if (used == nullptr) return fn(0, n_frames);
    for (size_t i = 0; i < used->size();) {
        size_t j = i + 1;
        while (j < used->size() && (*used)[j] == (*used)[j - 1] + 1) ++j;
        SG_TRY(fn((*used)[i], static_cast<int>(j - i)));
        i = j;
    }
    return SLAMGPU_OK;
}

// PUSH mode of the host-buffer batch call (pinned result lists): a slice's compaction and its way home run on a
// second stream, driven by the device alone -- finish + offsets (main stream) -> event -> compaction of the slice's
// rows into the device lists -> SM-store copy of that piece into the caller's pinned lists, while the main stream
// already scans the next slice.  The host queues everything up front and waits once at the end; the offsets reach it
// through a mapped mirror.  (Before, the host read each slice's offsets, queued two copies and waited for them: with
// eight ranks on one host its reaction time was on every rank's critical path, and the compaction -- 3.6 % of the GPU
// time -- sat between two scans.)  The rows scratch is double-buffered: finish of slice s + 2 waits for the compaction
// of slice s.  Measured on one box: resident-database batches 9,558 -> 9,692 G/s.  Moving the finish kernel to the
// second stream as well (scans back to back) was SLOWER, 9,483: its 131k CTAs get in the way of the persistent scan.
constexpr int kPushTaperFloor = 2048;
struct BatchPush {
    cudaStream_t stream;          // the second stream
    int32_t* rows2;               // second rows scratch [slice_pairs x max_rows]
    int32_t* h_query;             // the caller's pinned lists (device-addressable)
    int32_t* h_train;
    int32_t* h_offset_mirror;     // [n_pairs + 1] pinned
    cudaEvent_t* events;          // 2 per slice: finish done (main stream), compaction done (second stream)
};

// Pairs are processed in slices so that the partial-key scratch stays around L2 size.
constexpr size_t kBatchScratchBytes = size_t(256) << 20;

// Pairs of the next slice.  The host-buffer forms bring a slice's match lists back while the later slices are
// scanned, so only the LAST slice's copy is exposed: slices shrink geometrically toward the end of the call
// (each takes half of what is left, down to one wave of CTA pairs), which leaves a tail copy of a few hundred
// pairs instead of a full slice -- with eight ranks finishing a step at the same moment the full-size tail
// copies collided on the host.
int next_slice(const slamgpu_ctx* ctx, int remaining, int slice_pairs, bool taper, int floor_pairs_min = 0)
{
    if (!taper || remaining > 2 * slice_pairs) return std::min(slice_pairs, remaining);
    const int floor_pairs = std::max(std::max(ctx->n_sm * 2, 64), floor_pairs_min);
    if (remaining <= floor_pairs) return std::min(remaining, slice_pairs);
    return std::min(slice_pairs, std::max((remaining + 1) / 2, floor_pairs));
}

int batch_core(slamgpu_ctx* ctx, const uint8_t* d_desc, const int32_t* d_frame_offset, int n_frames,
               const int32_t* d_pairs, int n_pairs, int max_rows, float max_dist, float ratio, int32_t* d_out_train, int32_t* d_out_count,
               // optional ordered compaction (host-buffer form): per-slice out_train is then a scratch
               // of slice_pairs x max_rows and the lists are filled slice by slice
               int32_t* d_offset, int32_t* d_query_list, int32_t* d_train_list, int64_t capacity, int slice_pairs,
               // called after the work of slice [p0, p0 + np) has been queued on the stream
               const std::function<int(int, int)>* after_slice = nullptr,
               // sorted frames the pairs name, when they are few (host-buffer forms know them): only these get their
               // expanded copy built (a loop-closure query against a handful of candidates of a large resident database)
               const std::vector<int>* frames_used = nullptr,
               // a device-addressable host copy of the offsets (short calls: the kernels write results to host memory)
               int32_t* h_offset_mirror = nullptr,
               const BatchPush* push = nullptr,
               // the frames' FP4-expanded copy, already built (resident database): no expansion launches
               const uint8_t* exp4_ready = nullptr, int exp4_pitch = 0)
{
    if (n_pairs == 0) return SLAMGPU_OK;
    if (max_rows == 0) {
        SG_CUDA(cudaMemsetAsync(d_out_count, 0, sizeof(int32_t) * n_pairs, ctx->stream));
        if (d_offset) SG_CUDA(cudaMemsetAsync(d_offset, 0, sizeof(int32_t) * (n_pairs + 1), ctx->stream));
        return SLAMGPU_OK;
    }
    const bool tc = use_tc(ctx, max_rows, max_rows, n_pairs);
    const int cut = cut_for_filter(max_dist, ratio);
    const bool fp4 = tc && use_fp4(ctx, cut);
    BfPlan plan = tc ? tc_plan(ctx, max_rows, max_rows, std::min(n_pairs, slice_pairs), fp4)
                     : bf_make_plan(ctx->n_sm, max_rows, max_rows, std::min(n_pairs, slice_pairs));
    BfPlan fin = plan;
    int pitch = 0;
    // FP4 batches with one train split per pair: the scan filters and writes the rows itself (bf_tc4_kernels.cu: fused
    // finish); the short-call path keeps its own finish + compaction kernel
    const bool fused = fp4 && plan.splits == 1 && ctx->tc4_fuse_finish && !ctx->tc4_teams && h_offset_mirror == nullptr;
    const uint8_t* exp_base = nullptr;          // the expanded copy the scans read
    if (tc) {   // every frame expanded once per call; the scans then read only the expanded copy
        fin.splits = plan.splits * bf_tc_partials(ctx);
        uint8_t* e = nullptr;
        if (fp4 && exp4_ready != nullptr) {
            pitch = exp4_pitch;
            exp_base = exp4_ready;
        } else if (fp4) {
            pitch = bf_tc4_pitch_rows(max_rows);
            SG_TRY(ensure_device_exp(ctx, bf_tc4_expanded_bytes(n_frames, max_rows)));
            e = static_cast<uint8_t*>(ctx->d_exp.ptr);
            const size_t stride = bf_tc4_expanded_bytes(1, max_rows);
            if (frames_used && n_pairs <= 32767) {      // few frames: one launch over the pair list itself
                SG_TRY(bf_tc4_launch_expand_list(ctx, d_desc, d_frame_offset, d_pairs, 2 * n_pairs, pitch, e));
            } else {
                SG_TRY(for_frame_runs(n_frames, frames_used, [&](int f0, int nf) {
                    return bf_tc4_launch_expand(ctx, d_desc, d_frame_offset + f0, nf, 0, pitch, e + f0 * stride);
                }));
            }
        } else {
            pitch = bf_tc_pitch_rows(max_rows);
            SG_TRY(ensure_device_exp(ctx, bf_tc_expanded_bytes(n_frames, max_rows)));
            e = static_cast<uint8_t*>(ctx->d_exp.ptr);
            const size_t stride = bf_tc_expanded_bytes(1, max_rows);
            if (frames_used && n_pairs <= 32767) {
                SG_TRY(bf_tc_launch_expand_list(ctx, d_desc, d_frame_offset, d_pairs, 2 * n_pairs, max_rows, pitch, e));
            } else {
                SG_TRY(for_frame_runs(n_frames, frames_used, [&](int f0, int nf) {
                    return bf_tc_launch_expand(ctx, d_desc, d_frame_offset + f0, nf, 0, max_rows, pitch, e + f0 * stride);
                }));
            }
        }
    }
    const size_t per_pair = bf_partial_elems(fin, 1, max_rows);
    SG_TRY(ensure_device_ws(ctx, Carver::padded(per_pair * sizeof(uint2) * std::min(n_pairs, slice_pairs))));
    Carver ws(ctx->d_ws.ptr);
    uint2* keys = ws.take<uint2>(per_pair * std::min(n_pairs, slice_pairs));
    if (d_offset) SG_CUDA(cudaMemsetAsync(d_offset, 0, sizeof(int32_t), ctx->stream));
    int slice_no = 0;
    for (int p0 = 0, np = 0; p0 < n_pairs; p0 += np, ++slice_no) {
        // (push mode: no host work per slice; the taper stops at 2048 pairs -- the tail is that slice's compaction and copy)
        np = next_slice(ctx, n_pairs - p0, slice_pairs, d_offset != nullptr, push ? kPushTaperFloor : 0);
        const int32_t* pairs = d_pairs + 2 * static_cast<size_t>(p0);
        int32_t* out_train = d_offset ? d_out_train : d_out_train + static_cast<size_t>(p0) * max_rows;
        if (push && (slice_no & 1)) out_train = push->rows2;
        if (push && fused && slice_no >= 2)      // (the fused scan writes the rows itself: wait before it, not before the finish)
            SG_CUDA(cudaStreamWaitEvent(ctx->stream, push->events[2 * (slice_no - 2) + 1], 0));
        if (tc) {
            const uint8_t* e = exp_base ? exp_base : static_cast<const uint8_t*>(ctx->d_exp.ptr);
            if (fp4) {
                const Tc4Fuse fz{max_dist, ratio, out_train, d_out_count + p0};
                SG_TRY(bf_tc4_launch_scan(ctx, plan, e, 0, e, 0, pitch, d_frame_offset, pairs, np, max_rows, cut, keys,
                                          fused ? &fz : nullptr));
            } else {
                SG_TRY(bf_tc_launch_scan(ctx, plan, e, 0, e, 0, pitch, d_frame_offset, pairs, np, max_rows, cut, keys));
            }
        } else {
            SG_TRY(bf_launch_scan(ctx, plan, d_desc, 0, d_desc, 0, d_frame_offset, pairs, np, max_rows, false, keys));
        }
        if (h_offset_mirror) {      // short call, host-visible lists: out_train is a 2 x np x max_rows scratch here
            SG_TRY(bf_launch_finish_compact(ctx, fin, keys, d_frame_offset, pairs, np, max_rows, max_dist, ratio, out_train,
                                            d_out_count + p0, d_offset + p0, h_offset_mirror + p0, d_query_list,
                                            d_train_list, capacity));
            if (after_slice) SG_TRY((*after_slice)(p0, np));
            continue;
        }
        if (push) {
            cudaStream_t side = push->stream;
            if (!fused) {
                if (slice_no >= 2)      // the rows scratch of slice s - 2 must have been compacted
                    SG_CUDA(cudaStreamWaitEvent(ctx->stream, push->events[2 * (slice_no - 2) + 1], 0));
                SG_TRY(bf_launch_finish_filter(ctx, fin, keys, 0, 0, d_frame_offset, pairs, np, max_rows, max_dist, ratio,
                                               out_train, d_out_count + p0));
            }
            SG_TRY(bf_launch_scan_counts(ctx, d_out_count + p0, np, d_offset + p0, push->h_offset_mirror + p0));
            SG_CUDA(cudaEventRecord(push->events[2 * slice_no], ctx->stream));
            SG_CUDA(cudaStreamWaitEvent(side, push->events[2 * slice_no], 0));
            SG_TRY(bf_launch_compact(ctx, out_train, d_frame_offset, pairs, 0, np, max_rows, d_offset + p0, d_query_list,
                                     d_train_list, capacity, side));
            SG_TRY(bf_launch_push_lists(ctx, d_offset + p0, np, d_query_list, d_train_list, push->h_query, push->h_train,
                                        capacity, side));
            SG_CUDA(cudaEventRecord(push->events[2 * slice_no + 1], side));
            continue;
        }
        if (!fused)
            SG_TRY(bf_launch_finish_filter(ctx, fin, keys, 0, 0, d_frame_offset, pairs, np, max_rows, max_dist, ratio,
                                           out_train, d_out_count + p0));
        if (d_offset) {
            SG_TRY(bf_launch_scan_counts(ctx, d_out_count + p0, np, d_offset + p0, h_offset_mirror ? h_offset_mirror + p0 : nullptr));
            SG_TRY(bf_launch_compact(ctx, out_train, d_frame_offset, pairs, 0, np, max_rows, d_offset + p0,
                                     d_query_list, d_train_list, capacity));
        }
        if (after_slice) SG_TRY((*after_slice)(p0, np));
    }
    return SLAMGPU_OK;
}

int slice_for(const slamgpu_ctx* ctx, int max_rows)
{
    if (ctx->batch_slice > 0) return ctx->batch_slice;
    const int n_sm = ctx->n_sm;
    // one slice = enough pairs to fill the GPU a few times over, bounded by the scratch budget
    const size_t per_pair = static_cast<size_t>(max_rows) * sizeof(uint2);
    size_t s = per_pair ? kBatchScratchBytes / per_pair : 1;
    s = std::max<size_t>(s, static_cast<size_t>(n_sm) * 2);
    return static_cast<int>(std::min<size_t>(s, 1 << 20));
}


// ---- host-buffer plumbing -------------------------------------------------------------------
// Stage a host array into the pinned buffer at `h_off` and queue its H2D copy to `d_dst`.
// Pinned caller memory (slamgpu_host_alloc) skips the staging copy.
bool is_pinned(const void* p)
{
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return attr.type == cudaMemoryTypeHost;
}

int upload(slamgpu_ctx* ctx, void* d_dst, const void* h_src, size_t bytes, size_t h_off)
{
    if (bytes == 0) return SLAMGPU_OK;
    const void* src = h_src;
    if (!is_pinned(h_src)) {
        void* stage = static_cast<uint8_t*>(ctx->h_stage.ptr) + h_off;
        std::memcpy(stage, h_src, bytes);
        src = stage;
    }
    SG_CUDA(cudaMemcpyAsync(d_dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return SLAMGPU_OK;
}

struct PairIo {
    uint8_t* d_q = nullptr;
    uint8_t* d_t = nullptr;
    int32_t* d_a = nullptr;   // first output array
    int32_t* d_b = nullptr;   // second output array
    size_t out_a_bytes = 0, out_b_bytes = 0;
    bool zero_copy = false;   // d_a / d_b ARE the pinned host image: the kernels write the results straight into it
    int32_t* d_b_dev = nullptr;   // the device slot of the second output whatever the mode (atomic targets stay on the device)
};

// Outputs up to this size are written by the kernels directly into the pinned staging buffer (page-locked host
// memory is device-addressable): a single-pair call is then one upload, the kernels and one synchronisation --
// no download to queue.  The reference makes exactly such calls, one per loop candidate (src/LoopDetector.cpp:181).
constexpr size_t kZeroCopyOutBytes = size_t(256) << 10;

// device + pinned buffers for one (query, train) call with two int32 outputs, inputs uploaded
int stage_pair(slamgpu_ctx* ctx, const uint8_t* q, int nq, const uint8_t* t, int nt, size_t out_a_elems,
               size_t out_b_elems, PairIo& io)
{
    const size_t qb = Carver::padded(static_cast<size_t>(nq) * 32), tb = Carver::padded(static_cast<size_t>(nt) * 32);
    io.out_a_bytes = out_a_elems * sizeof(int32_t);
    io.out_b_bytes = out_b_elems * sizeof(int32_t);
    const size_t ob = Carver::padded(io.out_a_bytes) + Carver::padded(io.out_b_bytes);
    SG_TRY(ensure_device_io(ctx, qb + tb + ob + 256));
    SG_TRY(ensure_host_stage(ctx, qb + tb + ob + 256));
    Carver dio(ctx->d_io.ptr);
    io.d_q = dio.take<uint8_t>(static_cast<size_t>(nq) * 32);
    io.d_t = dio.take<uint8_t>(static_cast<size_t>(nt) * 32);
    io.d_a = dio.take<int32_t>(out_a_elems);
    io.d_b = dio.take<int32_t>(out_b_elems);
    io.d_b_dev = io.d_b;
    io.zero_copy = ctx->bf_zero_copy && ob <= kZeroCopyOutBytes && nq > 0 && nt > 0;    // (empty sets are answered by memsets)
    if (io.zero_copy) {       // the results' host image sits behind the inputs in the staging buffer
        uint8_t* h_out = static_cast<uint8_t*>(ctx->h_stage.ptr) + qb + tb;
        io.d_a = reinterpret_cast<int32_t*>(h_out);
        io.d_b = reinterpret_cast<int32_t*>(h_out + Carver::padded(io.out_a_bytes));
    }
    if (!is_pinned(q) && !is_pinned(t)) {     // both sets through the staging buffer: one copy for the two
        uint8_t* hs = static_cast<uint8_t*>(ctx->h_stage.ptr);
        std::memcpy(hs, q, static_cast<size_t>(nq) * 32);
        std::memcpy(hs + qb, t, static_cast<size_t>(nt) * 32);
        if (qb + tb) SG_CUDA(cudaMemcpyAsync(io.d_q, hs, qb + static_cast<size_t>(nt) * 32, cudaMemcpyHostToDevice, ctx->stream));
        return SLAMGPU_OK;
    }
    SG_TRY(upload(ctx, io.d_q, q, static_cast<size_t>(nq) * 32, 0));
    SG_TRY(upload(ctx, io.d_t, t, static_cast<size_t>(nt) * 32, qb));
    return SLAMGPU_OK;
}

// the outputs in pinned memory (one D2H copy of both unless the kernels wrote them there), then wait
int fetch_pair(slamgpu_ctx* ctx, const PairIo& io, const int32_t** h_a, const int32_t** h_b)
{
    if (io.zero_copy) {
        SG_TRY(wait_stream(ctx));
        *h_a = io.d_a;
        *h_b = io.d_b;
        return SLAMGPU_OK;
    }
    const size_t ob = Carver::padded(io.out_a_bytes) + io.out_b_bytes;
    if (ob) SG_CUDA(cudaMemcpyAsync(ctx->h_stage.ptr, io.d_a, ob, cudaMemcpyDeviceToHost, ctx->stream));
    SG_TRY(wait_stream(ctx));
    *h_a = static_cast<const int32_t*>(ctx->h_stage.ptr);
    *h_b = reinterpret_cast<const int32_t*>(static_cast<const uint8_t*>(ctx->h_stage.ptr) + Carver::padded(io.out_a_bytes));
    return SLAMGPU_OK;
}

// The resident database's FP4-expanded copy, complete for frames [0, n_frames) (see KfDbDev).
int kfdb_expanded(slamgpu_ctx* ctx, KfDbDev& d, int n_frames, int max_rows)
{
    const int pitch = bf_tc4_pitch_rows(max_rows);
    const size_t stride = bf_tc4_expanded_bytes(1, max_rows), need = stride * static_cast<size_t>(n_frames);
    if (d.exp4 == nullptr || d.exp4_pitch != pitch || need > d.exp4_bytes) {
        const bool keep = d.exp4 != nullptr && d.exp4_pitch == pitch;
        const size_t bytes = stride * static_cast<size_t>(std::max(2 * n_frames, 64));
        uint8_t* p = nullptr;
        const cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(SLAMGPU_ERR_NOMEM, std::string("expanded key-frame store allocation failed: ") + cudaGetErrorString(e));
        }
        if (keep && d.exp4_frames > 0)
            SG_CUDA(cudaMemcpyAsync(p, d.exp4, stride * static_cast<size_t>(d.exp4_frames), cudaMemcpyDeviceToDevice, ctx->stream));
        if (d.exp4) {
            SG_CUDA(cudaStreamSynchronize(ctx->stream));
            cudaFree(d.exp4);
        }
        d.exp4 = p;
        d.exp4_bytes = bytes;
        d.exp4_pitch = pitch;
        if (!keep) d.exp4_frames = 0;
    }
    if (d.exp4_frames < n_frames) {
        SG_TRY(bf_tc4_launch_expand(ctx, d.desc, d.offset + d.exp4_frames, n_frames - d.exp4_frames, 0, pitch,
                                    d.exp4 + stride * static_cast<size_t>(d.exp4_frames)));
        d.exp4_frames = n_frames;
    }
    return SLAMGPU_OK;
}

// The host-buffer batch call behind slamgpu_match_descriptors_batch (descriptors uploaded per call) and
// slamgpu_kfdb_match_pairs (d_desc_res / d_off_res: the resident database; nothing but the pair list goes up).
int batch_host(slamgpu_ctx* ctx, const uint8_t* desc, const uint8_t* d_desc_res, const int32_t* d_off_res,
               const int32_t* frame_offset, int n_frames, int max_rows, const int32_t* pairs, int n_pairs, float max_dist,
               float ratio, int32_t* out_offset, int32_t* out_query, int32_t* out_train, int64_t capacity,
               int64_t* out_total, KfDbDev* kf = nullptr)
{
    const size_t n_rows = static_cast<size_t>(frame_offset[n_frames]);
    if (max_rows == 0) {   // every frame is empty: no matches (src/MapMatcher.cpp:139-141)
        std::memset(out_offset, 0, sizeof(int32_t) * (static_cast<size_t>(n_pairs) + 1));
        return SLAMGPU_OK;
    }
    const int slice = std::min(n_pairs, slice_for(ctx, max_rows));
    const bool resident = d_desc_res != nullptr;
    const bool direct = capacity > 0 && is_pinned(out_query) && is_pinned(out_train);
    // a SHORT call into pinned result arrays (a loop-closure query: the newest key frame against a handful of candidates)
    const bool tiny = ctx->bf_zero_copy && direct && n_pairs <= 256 && n_pairs <= slice;
    const size_t b_desc = resident ? 0 : Carver::padded(n_rows * 32),
                 b_off = resident ? 0 : Carver::padded(sizeof(int32_t) * (n_frames + 1)), b_pairs = Carver::padded(sizeof(int32_t) * 2 * n_pairs);
    // batches into pinned result lists: compaction and the lists' way home on a second stream (BatchPush)
    const bool push_mode = ctx->bf_push && direct && !tiny;
    const size_t b_rows = Carver::padded(sizeof(int32_t) * static_cast<size_t>(slice) * std::max(max_rows, 1) * ((tiny || push_mode) ? 2 : 1)),
                 b_count = Carver::padded(sizeof(int32_t) * n_pairs),
                 b_offset = Carver::padded(sizeof(int32_t) * (static_cast<size_t>(n_pairs) + 1)),
                 b_list = Carver::padded(sizeof(int32_t) * static_cast<size_t>(capacity));
    SG_TRY(ensure_device_io(ctx, b_desc + b_off + b_pairs + b_rows + b_count + b_offset + 2 * b_list + 256));
    SG_TRY(ensure_host_stage(ctx, std::max(b_desc + b_off + b_pairs, b_offset + 2 * b_list) + 256));
    Carver dio(ctx->d_io.ptr);
    const uint8_t* d_desc = d_desc_res;
    const int32_t* d_off = d_off_res;
    if (!resident) {
        uint8_t* dd = dio.take<uint8_t>(n_rows * 32);
        int32_t* doff = dio.take<int32_t>(n_frames + 1);
        SG_TRY(upload(ctx, dd, desc, n_rows * 32, 0));
        SG_TRY(upload(ctx, doff, frame_offset, sizeof(int32_t) * (n_frames + 1), b_desc));
        d_desc = dd;
        d_off = doff;
    }
    int32_t* d_pairs = dio.take<int32_t>(2 * static_cast<size_t>(n_pairs));
    int32_t* d_rows = dio.take<int32_t>(static_cast<size_t>(slice) * std::max(max_rows, 1) * ((tiny || push_mode) ? 2 : 1));
    int32_t* d_count = dio.take<int32_t>(n_pairs);
    int32_t* d_offset = dio.take<int32_t>(static_cast<size_t>(n_pairs) + 1);
    int32_t* d_ql = dio.take<int32_t>(static_cast<size_t>(capacity));
    int32_t* d_tl = dio.take<int32_t>(static_cast<size_t>(capacity));
    SG_TRY(upload(ctx, d_pairs, pairs, sizeof(int32_t) * 2 * n_pairs, b_desc + b_off));
    // The lists come back slice by slice: once a slice's kernels are done its offsets are copied out on a second
    // stream, the host reads them, queues the copy of that slice's piece of the two lists and moves the previous
    // piece into the caller's arrays -- all while the GPU scans the later slices.  Pinned caller arrays
    // (slamgpu_host_alloc) receive the copies directly.
    if (ctx->copy_stream == nullptr) SG_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    int n_slices = 0;
    for (int left = n_pairs; left > 0; ++n_slices) left -= next_slice(ctx, left, slice, true);
    while (ctx->slice_events.size() < 3 * static_cast<size_t>(n_slices)) {
        cudaEvent_t e = nullptr;
        SG_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ctx->slice_events.push_back(e);
    }
    // SLAMGPU_TRACE_BATCH=1: host timestamps of this call's phases, summed per context (slamgpu_trace_read): where a
    // step's host time goes when eight ranks share one host
    const bool trace = ctx->trace_batch;
    auto now = []() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_call = trace ? now() : 0.0;
    uint8_t* hs = static_cast<uint8_t*>(ctx->h_stage.ptr);
    int32_t* h_offset = reinterpret_cast<int32_t*>(hs);
    int32_t* h_ql = direct ? out_query : reinterpret_cast<int32_t*>(hs + b_offset);
    int32_t* h_tl = direct ? out_train : reinterpret_cast<int32_t*>(hs + b_offset + b_list);
    // the uploads above may have been staged through the same pinned block: the downloads must not start
    // before they have been consumed (stream order covers it: slice 0 runs after the uploads)
    int slice_no = 0;
    std::vector<std::pair<int, int>> slices;      // (first pair, pair count) of every slice queued
    const std::function<int(int, int)> after_slice = [&](int p0, int np) -> int {
        cudaEvent_t done = ctx->slice_events[3 * slice_no], off_ready = ctx->slice_events[3 * slice_no + 1];
        ++slice_no;
        slices.emplace_back(p0, np);
        SG_CUDA(cudaEventRecord(done, ctx->stream));
        SG_CUDA(cudaStreamWaitEvent(ctx->copy_stream, done, 0));
        SG_CUDA(cudaMemcpyAsync(h_offset + p0 + 1, d_offset + p0 + 1, sizeof(int32_t) * np, cudaMemcpyDeviceToHost,
                                ctx->copy_stream));
        SG_CUDA(cudaEventRecord(off_ready, ctx->copy_stream));
        return SLAMGPU_OK;
    };
    // resident database on the FP4 engine: its expanded copy is kept with it (only new key frames are expanded)
    const uint8_t* exp4_ready = nullptr;
    int exp4_pitch = 0;
    if (kf != nullptr && ctx->kfdb_cache_exp && use_tc(ctx, max_rows, max_rows, n_pairs) &&
        use_fp4(ctx, cut_for_filter(max_dist, ratio))) {
        SG_TRY(kfdb_expanded(ctx, *kf, n_frames, max_rows));
        exp4_ready = kf->exp4;
        exp4_pitch = kf->exp4_pitch;
    }
    std::vector<int> used;
    if (n_pairs <= 256) {                               // few pairs: name the frames they touch
        used.assign(pairs, pairs + 2 * static_cast<size_t>(n_pairs));
        std::sort(used.begin(), used.end());
        used.erase(std::unique(used.begin(), used.end()), used.end());
    }
    const bool few = !used.empty() && used.size() * 2 < static_cast<size_t>(n_frames);
    if (tiny) {
        // one slice, the offsets and the two lists written by the kernels straight into host memory -- no copy
        // stream, no events, one synchronisation.
        // (the offsets' host image shares the staging block with the pair list's upload: both are stream-ordered)
        const int rc1 = batch_core(ctx, d_desc, d_off, n_frames, d_pairs, n_pairs, max_rows, max_dist, ratio, d_rows, d_count,
                                   d_offset, out_query, out_train, capacity, slice, nullptr, few ? &used : nullptr, h_offset,
                                   nullptr, exp4_ready, exp4_pitch);
        if (rc1 != SLAMGPU_OK) return rc1;
        SG_TRY(wait_stream(ctx));
        out_offset[0] = 0;
        std::memcpy(out_offset + 1, h_offset + 1, sizeof(int32_t) * static_cast<size_t>(n_pairs));
        *out_total = out_offset[n_pairs];
        if (*out_total > capacity) return fail(SLAMGPU_ERR_NOMEM, "match list capacity too small; *out_total holds the size needed");
        return SLAMGPU_OK;
    }
    if (push_mode) {
        BatchPush pm;
        pm.stream = ctx->copy_stream;
        pm.rows2 = d_rows + static_cast<size_t>(slice) * std::max(max_rows, 1);
        pm.h_query = out_query;
        pm.h_train = out_train;
        pm.h_offset_mirror = h_offset;           // (shares the staging block with the uploads: stream-ordered behind them)
        pm.events = ctx->slice_events.data();    // 3 per slice exist, 2 per slice are used
        const int rc1 = batch_core(ctx, d_desc, d_off, n_frames, d_pairs, n_pairs, max_rows, max_dist, ratio, d_rows, d_count,
                                   d_offset, d_ql, d_tl, capacity, slice, nullptr, few ? &used : nullptr, nullptr, &pm,
                                   exp4_ready, exp4_pitch);
        const double t_queued = trace ? now() : 0.0;
        const cudaError_t e1 = cudaStreamSynchronize(ctx->copy_stream), e2 = cudaStreamSynchronize(ctx->stream);
        if (rc1 != SLAMGPU_OK) return rc1;
        SG_CUDA(e1);
        SG_CUDA(e2);
        if (trace) {
            const double t_end = now();
            ctx->trace_us[0] += t_queued - t_call;
            ctx->trace_us[1] += t_end - t_queued;
            ctx->trace_us[3] += t_end - t_call;
            ctx->trace_calls++;
        }
        out_offset[0] = 0;
        std::memcpy(out_offset + 1, h_offset + 1, sizeof(int32_t) * static_cast<size_t>(n_pairs));
        *out_total = out_offset[n_pairs];
        bool wrapped = false;                    // one slice adds < 2^31 entries: a wrapped total shows as an end below its begin
        for (int p0 = 0, left = n_pairs; left > 0;) {
            const int np = next_slice(ctx, left, slice, true, kPushTaperFloor);
            wrapped = wrapped || out_offset[p0 + np] < out_offset[p0];
            p0 += np;
            left -= np;
        }
        if (wrapped) return fail(SLAMGPU_ERR_INVALID, "more than 2^31 - 1 matches in one batch call: split the pair list");
        if (*out_total > capacity) return fail(SLAMGPU_ERR_NOMEM, "match list capacity too small; *out_total holds the size needed");
        return SLAMGPU_OK;
    }
    const int rc = batch_core(ctx, d_desc, d_off, n_frames, d_pairs, n_pairs, max_rows, max_dist, ratio, d_rows, d_count,
                              d_offset, d_ql, d_tl, capacity, slice, &after_slice, few ? &used : nullptr, nullptr, nullptr,
                              exp4_ready, exp4_pitch);
    if (rc != SLAMGPU_OK) {
        cudaStreamSynchronize(ctx->copy_stream);
        return rc;
    }
    const double t_queued = trace ? now() : 0.0;       // every kernel of the call is queued; the GPU is working
    // h_offset[0] is never written or read on the host: the pinned block's first bytes may still be the source
    // of an upload queued behind an unsynchronised *_dev call (the descriptors, or the pair list of the resident
    // form), so nothing on the host may touch it before the stream has drained.  Offset 0 is the constant 0.
    bool overflow = false, wrapped = false;
    int64_t begin = 0, prev_begin = 0, prev_end = 0;
    int prev = -1;
    auto unstage = [&](int s, int64_t b, int64_t e) -> int {      // piece [b, e) of slice s into the caller's arrays
        SG_CUDA(cudaEventSynchronize(ctx->slice_events[3 * s + 2]));
        if (!direct && e > b) {
            std::memcpy(out_query + b, h_ql + b, sizeof(int32_t) * (e - b));
            std::memcpy(out_train + b, h_tl + b, sizeof(int32_t) * (e - b));
        }
        return SLAMGPU_OK;
    };
    double t_last_offsets = 0.0;
    for (int s = 0; s < slice_no; ++s) {
        const int p0 = slices[s].first, np = slices[s].second;
        SG_CUDA(cudaEventSynchronize(ctx->slice_events[3 * s + 1]));
        if (trace) t_last_offsets = now();
        const int64_t end = h_offset[p0 + np];
        // offsets are int32 on the device; one slice adds < 2^31 entries (slice_for), so a wrapped total shows up
        // as an end below its begin
        wrapped = wrapped || end < begin;
        overflow = overflow || wrapped || end > capacity;
        if (!overflow) {
            if (end > begin) {
                SG_CUDA(cudaMemcpyAsync(h_ql + begin, d_ql + begin, sizeof(int32_t) * (end - begin), cudaMemcpyDeviceToHost,
                                        ctx->copy_stream));
                SG_CUDA(cudaMemcpyAsync(h_tl + begin, d_tl + begin, sizeof(int32_t) * (end - begin), cudaMemcpyDeviceToHost,
                                        ctx->copy_stream));
            }
            SG_CUDA(cudaEventRecord(ctx->slice_events[3 * s + 2], ctx->copy_stream));
            if (prev >= 0) SG_TRY(unstage(prev, prev_begin, prev_end));
            prev = s;
            prev_begin = begin;
            prev_end = end;
        }
        begin = end;
    }
    if (prev >= 0) SG_TRY(unstage(prev, prev_begin, prev_end));
    SG_CUDA(cudaStreamSynchronize(ctx->copy_stream));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    if (trace) {
        const double t_end = now();
        ctx->trace_us[0] += t_queued - t_call;            // staging, uploads queued, every slice's kernels queued
        ctx->trace_us[1] += t_last_offsets - t_queued;    // waiting for the GPU (slices' offsets arriving; list copies queued meanwhile)
        ctx->trace_us[2] += t_end - t_last_offsets;       // the tail: the last slice's lists come down, final synchronisation
        ctx->trace_us[3] += t_end - t_call;
        ctx->trace_calls++;
    }
    out_offset[0] = 0;
    std::memcpy(out_offset + 1, h_offset + 1, sizeof(int32_t) * static_cast<size_t>(n_pairs));
    *out_total = out_offset[n_pairs];
    if (wrapped) return fail(SLAMGPU_ERR_INVALID, "more than 2^31 - 1 matches in one batch call: split the pair list");
    if (overflow) return fail(SLAMGPU_ERR_NOMEM, "match list capacity too small; *out_total holds the size needed");
    return SLAMGPU_OK;
}

void free_kfdb(KfDbDev& d)
{
    if (d.desc) cudaFree(d.desc);
    if (d.offset) cudaFree(d.offset);
    if (d.exp4) cudaFree(d.exp4);
    d = KfDbDev();
}

} // namespace

extern "C" {

int slamgpu_knn2(slamgpu_ctx* ctx, const uint8_t* query, int nq, const uint8_t* train, int nt, int32_t* out_idx,
                 int32_t* out_dist)
{
    SG_CTX(ctx);
    SG_TRY(check_sets(query, nq, train, nt));
    if (nq == 0) return SLAMGPU_OK;
    SG_CHECK(out_idx && out_dist, "null output");
    PairIo io;
    SG_TRY(stage_pair(ctx, query, nq, train, nt, 2 * static_cast<size_t>(nq), 2 * static_cast<size_t>(nq), io));
    SG_TRY(knn2_core(ctx, io.d_q, nq, io.d_t, nt, io.d_a, io.d_b));
    const int32_t *ha, *hb;
    SG_TRY(fetch_pair(ctx, io, &ha, &hb));
    std::memcpy(out_idx, ha, io.out_a_bytes);
    std::memcpy(out_dist, hb, io.out_b_bytes);
    return SLAMGPU_OK;
}

int slamgpu_crosscheck(slamgpu_ctx* ctx, const uint8_t* query, int nq, const uint8_t* train, int nt,
                       int32_t* out_idx, int32_t* out_dist)
{
    SG_CTX(ctx);
    SG_TRY(check_sets(query, nq, train, nt));
    if (nq == 0) return SLAMGPU_OK;
    SG_CHECK(out_idx && out_dist, "null output");
    PairIo io;
    SG_TRY(stage_pair(ctx, query, nq, train, nt, nq, nq, io));
    SG_TRY(crosscheck_core(ctx, io.d_q, nq, io.d_t, nt, -1.0f, io.d_a, io.d_b));
    const int32_t *ha, *hb;
    SG_TRY(fetch_pair(ctx, io, &ha, &hb));
    std::memcpy(out_idx, ha, io.out_a_bytes);
    std::memcpy(out_dist, hb, io.out_b_bytes);
    return SLAMGPU_OK;
}

int slamgpu_match_descriptors(slamgpu_ctx* ctx, const uint8_t* query, int nq, const uint8_t* train, int nt,
                              float max_dist, float ratio, int32_t* out_query, int32_t* out_train, int* out_count)
{
    SG_CTX(ctx);
    SG_TRY(check_sets(query, nq, train, nt));
    SG_CHECK(out_count != nullptr, "null out_count");
    *out_count = 0;
    if (nq == 0 || nt == 0) return SLAMGPU_OK;   // src/MapMatcher.cpp:139-141
    SG_CHECK(out_query && out_train, "null output");
    PairIo io;
    SG_TRY(stage_pair(ctx, query, nq, train, nt, nq, 1, io));
    // zero-copy: the host counts the surviving rows itself, so the device count (a memset and atomics) is left out
    SG_TRY(match_descriptors_core(ctx, io.d_q, nq, io.d_t, nt, max_dist, ratio, io.d_a, io.zero_copy ? nullptr : io.d_b_dev));
    const int32_t *h_train, *h_count;
    SG_TRY(fetch_pair(ctx, io, &h_train, &h_count));
    // marshal the per-row result into the reference's list form (ascending queryIdx, :151-161)
    int n = 0;
    for (int q = 0; q < nq; ++q) {
        if (h_train[q] >= 0) {
            out_query[n] = q;
            out_train[n] = h_train[q];
            ++n;
        }
    }
    if (!io.zero_copy && n != h_count[0]) return fail(SLAMGPU_ERR_CUDA, "match_descriptors: device count and row flags disagree");
    *out_count = n;
    return SLAMGPU_OK;
}

int slamgpu_match_features(slamgpu_ctx* ctx, const uint8_t* prev, int nprev, const uint8_t* cur, int ncur,
                           float max_dist, int32_t* out_train, int32_t* out_query, int* out_count)
{
    SG_CTX(ctx);
    SG_TRY(check_sets(cur, ncur, prev, nprev));
    SG_CHECK(out_count != nullptr, "null out_count");
    SG_CHECK(!(max_dist < 0.0f), "negative max_dist");
    *out_count = 0;
    if (ncur == 0 || nprev == 0) return SLAMGPU_OK;
    SG_CHECK(out_query && out_train, "null output");
    PairIo io;
    SG_TRY(stage_pair(ctx, cur, ncur, prev, nprev, ncur, ncur, io));
    SG_TRY(crosscheck_core(ctx, io.d_q, ncur, io.d_t, nprev, max_dist, io.d_a, io.d_b));
    const int32_t *h_idx, *h_dist;
    SG_TRY(fetch_pair(ctx, io, &h_idx, &h_dist));
    int n = 0;
    for (int q = 0; q < ncur; ++q) {   // FeatureMatch(trainIdx, queryIdx), src/features/FeatureExtractor.cpp:26
        if (h_idx[q] >= 0) {
            out_train[n] = h_idx[q];
            out_query[n] = q;
            ++n;
        }
    }
    *out_count = n;
    return SLAMGPU_OK;
}

int slamgpu_match_descriptors_batch(slamgpu_ctx* ctx, const uint8_t* desc, const int32_t* frame_offset, int n_frames,
                                    const int32_t* pairs, int n_pairs, float max_dist, float ratio,
                                    int32_t* out_offset, int32_t* out_query, int32_t* out_train, int64_t capacity,
                                    int64_t* out_total)
{
    SG_CTX(ctx);
    SG_CHECK(n_frames >= 0 && n_pairs >= 0 && capacity >= 0, "negative size");
    SG_CHECK(out_total != nullptr, "null out_total");
    *out_total = 0;
    if (n_pairs == 0) return SLAMGPU_OK;
    SG_CHECK(frame_offset && pairs && out_offset, "null pointer");
    SG_CHECK(capacity == 0 || (out_query && out_train), "null output lists");
    SG_CHECK(frame_offset[0] == 0, "frame_offset[0] must be 0");
    int max_rows = 0;
    for (int f = 0; f < n_frames; ++f) {
        SG_CHECK(frame_offset[f + 1] >= frame_offset[f], "frame_offset must be non-decreasing");
        max_rows = std::max(max_rows, frame_offset[f + 1] - frame_offset[f]);
    }
    SG_CHECK(max_rows <= SLAMGPU_MAX_TRAIN, "frame larger than SLAMGPU_MAX_TRAIN");
    for (int p = 0; p < 2 * n_pairs; ++p) SG_CHECK(pairs[p] >= 0 && pairs[p] < n_frames, "pair names an unknown frame");
    SG_CHECK(frame_offset[n_frames] == 0 || desc != nullptr, "null descriptors");
    return batch_host(ctx, desc, nullptr, nullptr, frame_offset, n_frames, max_rows, pairs, n_pairs, max_dist, ratio,
                      out_offset, out_query, out_train, capacity, out_total);
}

// ---- resident key-frame database ---------------------------------------------------------------

int slamgpu_kfdb_append(slamgpu_ctx* ctx, int db_id, const uint8_t* desc, int n_rows, int* out_frame)
{
    SG_CTX(ctx);
    SG_CHECK(n_rows >= 0, "negative row count");
    SG_CHECK(n_rows == 0 || desc != nullptr, "null descriptors");
    SG_CHECK(n_rows <= SLAMGPU_MAX_TRAIN, "frame larger than SLAMGPU_MAX_TRAIN");
    KfDbDev& d = ctx->kf_dbs[db_id];
    if (d.h_offset.empty()) d.h_offset.push_back(0);
    const int n_frames = static_cast<int>(d.h_offset.size()) - 1;
    const int64_t rows = d.h_offset.back();
    SG_CHECK(rows + n_rows <= 0x7FFFFFFF, "key-frame database too large");
    if (n_frames + 1 > d.cap_frames) {      // grow geometrically; the live part moves device-to-device
        const int cap = std::max(1024, 2 * d.cap_frames);
        int32_t* p = nullptr;
        SG_CUDA(cudaMalloc(&p, (static_cast<size_t>(cap) + 1) * sizeof(int32_t)));
        if (d.offset) {
            SG_CUDA(cudaMemcpyAsync(p, d.offset, (static_cast<size_t>(n_frames) + 1) * 4, cudaMemcpyDeviceToDevice, ctx->stream));
            SG_CUDA(cudaStreamSynchronize(ctx->stream));
            cudaFree(d.offset);
        } else {
            SG_CUDA(cudaMemsetAsync(p, 0, 4, ctx->stream));
        }
        d.offset = p;
        d.cap_frames = cap;
    }
    if (rows + n_rows > d.cap_rows) {
        const int64_t cap = std::max<int64_t>(std::max<int64_t>(1 << 20, rows + n_rows), 2 * d.cap_rows);
        uint8_t* p = nullptr;
        const cudaError_t e = cudaMalloc(&p, static_cast<size_t>(cap) * 32);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(SLAMGPU_ERR_NOMEM, std::string("key-frame database growth failed: ") + cudaGetErrorString(e));
        }
        if (rows) {
            SG_CUDA(cudaMemcpyAsync(p, d.desc, static_cast<size_t>(rows) * 32, cudaMemcpyDeviceToDevice, ctx->stream));
            SG_CUDA(cudaStreamSynchronize(ctx->stream));
        }
        if (d.desc) cudaFree(d.desc);
        d.desc = p;
        d.cap_rows = cap;
    }
    const int32_t end = static_cast<int32_t>(rows + n_rows);
    if (n_rows) {
        SG_TRY(ensure_host_stage(ctx, static_cast<size_t>(n_rows) * 32 + 256));
        SG_TRY(upload(ctx, d.desc + static_cast<size_t>(rows) * 32, desc, static_cast<size_t>(n_rows) * 32, 0));
    }
    SG_CUDA(cudaMemcpyAsync(d.offset + n_frames + 1, &end, sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));       // the caller's rows and `end` have been consumed
    d.h_offset.push_back(end);
    d.max_rows = std::max(d.max_rows, n_rows);
    if (out_frame) *out_frame = n_frames;
    return SLAMGPU_OK;
}

int slamgpu_kfdb_count(slamgpu_ctx* ctx, int db_id, int* n_frames, int64_t* n_rows)
{
    SG_CHECK(ctx != nullptr, "null context");
    auto it = ctx->kf_dbs.find(db_id);
    const bool have = it != ctx->kf_dbs.end() && !it->second.h_offset.empty();
    if (n_frames) *n_frames = have ? static_cast<int>(it->second.h_offset.size()) - 1 : 0;
    if (n_rows) *n_rows = have ? it->second.h_offset.back() : 0;
    return SLAMGPU_OK;
}

int slamgpu_kfdb_release(slamgpu_ctx* ctx, int db_id)
{
    SG_CTX(ctx);
    auto it = ctx->kf_dbs.find(db_id);
    if (it == ctx->kf_dbs.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown key-frame database id");
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    free_kfdb(it->second);
    ctx->kf_dbs.erase(it);
    return SLAMGPU_OK;
}

int slamgpu_kfdb_match_pairs(slamgpu_ctx* ctx, int db_id, const int32_t* pairs, int n_pairs, float max_dist, float ratio,
                             int32_t* out_offset, int32_t* out_query, int32_t* out_train, int64_t capacity,
                             int64_t* out_total)
{
    SG_CTX(ctx);
    SG_CHECK(n_pairs >= 0 && capacity >= 0, "negative size");
    SG_CHECK(out_total != nullptr, "null out_total");
    *out_total = 0;
    if (n_pairs == 0) return SLAMGPU_OK;
    SG_CHECK(pairs && out_offset, "null pointer");
    SG_CHECK(capacity == 0 || (out_query && out_train), "null output lists");
    auto it = ctx->kf_dbs.find(db_id);
    if (it == ctx->kf_dbs.end() || it->second.h_offset.size() < 2)
        return fail(SLAMGPU_ERR_NOTFOUND, "unknown or empty key-frame database id");
    KfDbDev& d = it->second;
    const int n_frames = static_cast<int>(d.h_offset.size()) - 1;
    for (int p = 0; p < 2 * n_pairs; ++p) SG_CHECK(pairs[p] >= 0 && pairs[p] < n_frames, "pair names an unknown frame");
    return batch_host(ctx, nullptr, d.desc, d.offset, d.h_offset.data(), n_frames, d.max_rows, pairs, n_pairs, max_dist,
                      ratio, out_offset, out_query, out_train, capacity, out_total, &d);
}

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
    return batch_core(ctx, d_desc, d_frame_offset, n_frames, d_pairs, n_pairs, max_rows, max_dist, ratio, d_out_train,
                      d_out_count, nullptr, nullptr, nullptr, 0, slice_for(ctx, max_rows));
}

} // extern "C"

namespace slamgpu {
void free_all_kfdbs(slamgpu_ctx* ctx)
{
    for (auto& kv : ctx->kf_dbs) free_kfdb(kv.second);
    ctx->kf_dbs.clear();
}
} // namespace slamgpu
