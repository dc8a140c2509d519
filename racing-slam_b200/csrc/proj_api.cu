// proj_api.cu -- projection-guided entry points of the C ABI (include/slamgpu.h):
// resident frame / map-point state and the stream-ordered match_projection calls.  Kernels are
// in proj_kernels.cu.  Uploads go through the context's pinned staging buffer in one H2D copy
// per call; results come back in one D2H copy.
#include <algorithm>
#include <chrono>
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

void free_frame(FrameDev& f)
{
    if (f.pending && f.pending->worker.joinable()) f.pending->worker.join();   // a KD build still running
    if (f.blob) cudaFree(f.blob);
    f = FrameDev();
}
int finalize_frame(slamgpu_ctx* ctx, FrameDev& f);
int settle_frame(slamgpu_ctx* ctx, FrameDev& f, bool* speculative);
void join_frame(FrameDev& f);
size_t kd_image_bytes(size_t n, size_t n_cells);
void build_kd_image(FrameKdPending& k);

// One allocation carved for cap_n points / cap_obs observations; the first five arrays are in the
// order and padding of the staging image so that an exact-size upload is one copy.
size_t points_image_bytes(size_t N, size_t O)
{
    return Carver::padded(N * 12) + Carver::padded(N * 12) + Carver::padded(N * 8) + Carver::padded((N + 1) * 4) +
           Carver::padded(O * 32);
}
int alloc_points(PointSetDev& p, int cap_n, int64_t cap_obs)
{
    const size_t N = static_cast<size_t>(cap_n), O = static_cast<size_t>(cap_obs);
    const cudaError_t e = cudaMalloc(&p.blob, points_image_bytes(N, O) + Carver::padded(N) + 256);
    if (e != cudaSuccess) {
        cudaGetLastError();
        p.blob = nullptr;
        return fail(SLAMGPU_ERR_NOMEM, std::string("point-set allocation failed: ") + cudaGetErrorString(e));
    }
    Carver dv(p.blob);
    p.pos = dv.take<float>(N * 3);
    p.normal = dv.take<float>(N * 3);
    p.range = dv.take<float>(N * 2);
    p.obs_offset = dv.take<int32_t>(N + 1);
    p.obs_desc = dv.take<uint4>(O * 2);
    p.pt_matched = dv.take<uint8_t>(N);
    p.obs_end = p.obs_offset + 1;      // CSR
    p.cap_n = cap_n;
    p.cap_obs = cap_obs;
    return SLAMGPU_OK;
}

int check_params(const slamgpu_proj_params* p)
{
    SG_CHECK(p != nullptr, "null projection params");
    SG_CHECK(p->width >= 0 && p->height >= 0, "negative image size");
    SG_CHECK(p->radius >= 0.0f, "negative search radius");
    SG_CHECK(p->max_dist == p->max_dist, "max_dist is NaN");
    return SLAMGPU_OK;
}

// accept Hamming d iff (double)d < max_dist  (src/MapMatcher.cpp:78,88)  <=>  d < ceil(max_dist)
int dist_threshold(float max_dist)
{
    const double c = std::ceil(static_cast<double>(max_dist));
    if (c <= 0.0) return 0;
    if (c > 257.0) return 257;
    return static_cast<int>(c);
}

void fill_common(ProjCommonArgs& c, const PointSetDev& ps, const slamgpu_proj_params& p, const int32_t* d_order,
                 int n_visit)
{
    c.n_points = ps.n;
    c.pos = ps.pos;
    c.normal = ps.normal;
    c.range = ps.range;
    c.obs_offset = ps.obs_offset;
    c.obs_end = ps.obs_end;
    c.obs_desc = ps.obs_desc;
    c.order = d_order;
    c.n_visit = n_visit;
    c.width_f = static_cast<float>(p.width);      // int -> float promotion of src/Camera.cpp:36
    c.height_f = static_cast<float>(p.height);
    c.radius = p.radius;
    volatile float r2 = p.radius * p.radius;      // src/KDTree.cpp:48
    c.r2 = r2;
    c.min_cos = p.min_cos;
    c.nearer = p.nearer;
    c.further = p.further;
    c.thr = dist_threshold(p.max_dist);
    c.replace = p.replace ? 1 : 0;
    c.dbg_uv = nullptr;
    c.dbg_kp = nullptr;
    c.dbg_d = nullptr;
    c.stats = nullptr;
}

void fill_frame(ProjFrameArgs& f, const FrameDev& fd, const float* pose, const slamgpu_proj_params& p)
{
    f.nkp = fd.nkp;
    f.kd_xy = fd.kd_xy;
    f.kd_idx = fd.kd_idx;
    f.desc = fd.desc;
    f.matched = fd.matched;
    f.xy = fd.xy;
    f.proposals = fd.proposals;
    f.cell_items = fd.cell_items;
    f.kd_ready = fd.pending ? 0 : 1;                    // the KD image is resident once finalize_frame has run
    f.cell_start = f.kd_ready ? fd.cell_start : fd.cell_start_spec;
    f.cell_rec = f.kd_ready ? fd.cell_rec : fd.cell_rec_spec;
    f.ties = nullptr;
    f.ties_out = nullptr;
    f.gx0 = fd.gx0;
    f.gy0 = fd.gy0;
    f.inv_cell = fd.inv_cell;
    f.gw = fd.gw;
    f.gh = fd.gh;
    f.skip = nullptr;
    f.out_point = nullptr;
    f.out_dist = nullptr;
    host_projection_matrix(pose, p.fx, p.fy, p.cx, p.cy, f.M);
    host_camera_center(pose, f.C);
}

// One call = any number of resident frames against one point set.
int match_projection_impl(slamgpu_ctx* ctx, const int* frame_ids, int n_frames, int set_id, const float* poses,
                          const slamgpu_proj_params* params, const uint8_t* skip, const int32_t* order, int n_order,
                          int32_t* out_point, int32_t* out_dist, float* dbg_uv, int32_t* dbg_kp, int32_t* dbg_d,
                          int64_t* stats, const int32_t* matched_points = nullptr, int n_matched = 0)
{
    SG_CTX(ctx);
    SG_TRY(check_params(params));
    SG_CHECK(n_frames >= 0, "negative frame count");
    if (n_frames == 0) return SLAMGPU_OK;
    SG_CHECK(frame_ids != nullptr && poses != nullptr, "null frame list / poses");
    auto ps_it = ctx->point_sets.find(set_id);
    if (ps_it == ctx->point_sets.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown point set id");
    const PointSetDev& ps = ps_it->second;
    SG_CHECK(order != nullptr || n_order == 0 || n_order == ps.n, "n_order without an order array");
    // the sparse form of `skip` (one frame only): the points with frame.is_matched(point), as an index list
    SG_CHECK(n_matched >= 0 && (n_matched == 0 || matched_points != nullptr), "bad matched-point list");
    SG_CHECK(n_matched == 0 || (skip == nullptr && n_frames == 1), "matched_points excludes skip and batches");
    for (int i = 0; i < n_matched; ++i)
        SG_CHECK(matched_points[i] >= 0 && matched_points[i] < ps.n, "matched_points entry outside the point set");
    const int n_visit = order ? n_order : ps.n;
    SG_CHECK(n_visit >= 0, "negative visit count");
    if (order) {
        for (int i = 0; i < n_order; ++i)
            SG_CHECK(order[i] >= -1 && order[i] < ps.n, "order entry outside the point set (-1 = nullptr entry)");
    }

    std::vector<const FrameDev*> fds(n_frames);
    bool speculative = false;
    size_t total_kp = 0;
    int max_nkp = 0;
    for (int i = 0; i < n_frames; ++i) {
        auto it = ctx->frames.find(frame_ids[i]);
        if (it == ctx->frames.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown frame id");
        for (int j = 0; j < i; ++j)
            SG_CHECK(frame_ids[j] != frame_ids[i], "a frame may appear only once per projection batch");
        SG_TRY(settle_frame(ctx, it->second, &speculative));   // the KD-tree may still be building on its host thread
        fds[i] = &it->second;
        total_kp += static_cast<size_t>(it->second.nkp);
        max_nkp = std::max(max_nkp, it->second.nkp);
    }
    SG_CHECK(total_kp == 0 || (out_point != nullptr && out_dist != nullptr), "null output");
    if (total_kp == 0 && !(dbg_uv || dbg_kp || dbg_d || stats)) return SLAMGPU_OK;      // no keypoint can be matched

    const bool want_dbg = dbg_uv || dbg_kp || dbg_d || stats;
    const size_t dbg_rows = static_cast<size_t>(n_frames) * n_visit;

    // ---- device io layout: [skip][order][frame args][out_point][out_dist][dbg...] ----
    const size_t skip_bytes = skip ? Carver::padded(static_cast<size_t>(n_frames) * ps.n) : 0;
    const size_t order_bytes = order ? Carver::padded(sizeof(int32_t) * n_order) : 0;
    const size_t fargs_bytes = n_frames > 1 ? Carver::padded(sizeof(ProjFrameArgs) * n_frames) : 0;
    const size_t mlist_bytes = n_matched ? Carver::padded(sizeof(int32_t) * n_matched) : 0;
    const size_t up_bytes = skip_bytes + order_bytes + fargs_bytes + mlist_bytes;
    const size_t out_bytes = Carver::padded(sizeof(int32_t) * total_kp) * 2 + 256;    // + the tie counter
    size_t dbg_bytes = 0;
    if (want_dbg)
        dbg_bytes = Carver::padded(sizeof(float) * 2 * dbg_rows) + 2 * Carver::padded(sizeof(int32_t) * dbg_rows) +
                    Carver::padded(2 * sizeof(unsigned long long));
    SG_TRY(ensure_device_io(ctx, up_bytes + out_bytes + dbg_bytes + 256));
    SG_TRY(ensure_host_stage(ctx, up_bytes + out_bytes + dbg_bytes + 256));
    // ZERO COPY for the small per-frame calls (one frame, no diagnostics, under 1 MB of lists and results): the
    // kernels read the skip / order / matched lists from, and write the results to, the pinned staging buffer
    // itself (page-locked host memory is device-addressable).  The call is then two kernel launches and one
    // synchronisation instead of copy + launches + copy + synchronisation: about 20 us less on a 47 us call.
    const bool zero_copy = ctx->proj_zero_copy && n_frames == 1 && !want_dbg && up_bytes + out_bytes <= (size_t(1) << 20);

    Carver dio(ctx->d_io.ptr), hio(ctx->h_stage.ptr);
    uint8_t* d_skip = skip ? dio.take<uint8_t>(static_cast<size_t>(n_frames) * ps.n) : nullptr;
    uint8_t* h_skip = skip ? hio.take<uint8_t>(static_cast<size_t>(n_frames) * ps.n) : nullptr;
    int32_t* d_order = order ? dio.take<int32_t>(n_order) : nullptr;
    int32_t* h_order = order ? hio.take<int32_t>(n_order) : nullptr;
    ProjFrameArgs* d_fargs = n_frames > 1 ? dio.take<ProjFrameArgs>(n_frames) : nullptr;
    ProjFrameArgs* h_fargs_stage = n_frames > 1 ? hio.take<ProjFrameArgs>(n_frames) : nullptr;
    int32_t* d_mlist = n_matched ? dio.take<int32_t>(n_matched) : nullptr;
    int32_t* h_mlist = n_matched ? hio.take<int32_t>(n_matched) : nullptr;
    int32_t* d_out_point = dio.take<int32_t>(total_kp);
    int32_t* d_out_dist = dio.take<int32_t>(total_kp);
    unsigned int* d_ties = dio.take<unsigned int>(1);
    // the host image of the results sits behind the inputs in the staging buffer (zero copy: the kernels' target)
    uint8_t* h_out = static_cast<uint8_t*>(ctx->h_stage.ptr) + up_bytes;
    if (zero_copy) {
        d_skip = h_skip;
        d_order = h_order;
        d_mlist = h_mlist;
        d_out_point = reinterpret_cast<int32_t*>(h_out);
        d_out_dist = reinterpret_cast<int32_t*>(h_out + Carver::padded(sizeof(int32_t) * total_kp));
    }
    float* d_dbg_uv = nullptr;
    int32_t *d_dbg_kp = nullptr, *d_dbg_d = nullptr;
    unsigned long long* d_stats = nullptr;
    if (want_dbg) {
        d_dbg_uv = dio.take<float>(2 * dbg_rows);
        d_dbg_kp = dio.take<int32_t>(dbg_rows);
        d_dbg_d = dio.take<int32_t>(dbg_rows);
        d_stats = dio.take<unsigned long long>(2);
    }

    ProjCommonArgs common;
    fill_common(common, ps, *params, d_order, n_visit);
    if (want_dbg) {
        common.dbg_uv = d_dbg_uv;
        common.dbg_kp = d_dbg_kp;
        common.dbg_d = d_dbg_d;
        common.stats = stats ? d_stats : nullptr;
    }

    std::vector<ProjFrameArgs> fargs(n_frames);
    size_t kp_off = 0;
    for (int i = 0; i < n_frames; ++i) {
        fill_frame(fargs[i], *fds[i], poses + 16 * static_cast<size_t>(i), *params);
        fargs[i].skip = skip ? d_skip + static_cast<size_t>(i) * ps.n : (matched_points ? ps.pt_matched : nullptr);
        fargs[i].out_point = d_out_point + kp_off;
        fargs[i].out_dist = d_out_dist + kp_off;
        fargs[i].ties = d_ties;
        if (zero_copy && speculative)
            fargs[i].ties_out = reinterpret_cast<unsigned int*>(h_out + 2 * Carver::padded(sizeof(int32_t) * total_kp));
        kp_off += static_cast<size_t>(fds[i]->nkp);
    }
    if (speculative) SG_CUDA(cudaMemsetAsync(d_ties, 0, sizeof(unsigned int), ctx->stream));

    // ---- upload (one copy: the three pieces are contiguous in both arenas) ----
    if (up_bytes) {
        if (skip) std::memcpy(h_skip, skip, static_cast<size_t>(n_frames) * ps.n);
        if (order) std::memcpy(h_order, order, sizeof(int32_t) * n_order);
        if (n_frames > 1) std::memcpy(h_fargs_stage, fargs.data(), sizeof(ProjFrameArgs) * n_frames);
        if (n_matched) std::memcpy(h_mlist, matched_points, sizeof(int32_t) * n_matched);
        if (!zero_copy) SG_CUDA(cudaMemcpyAsync(ctx->d_io.ptr, ctx->h_stage.ptr, up_bytes, cudaMemcpyHostToDevice, ctx->stream));
    }
    if (matched_points) SG_TRY(proj_launch_flags(ctx, d_mlist, n_matched, ps.pt_matched, ps.n));
    if (want_dbg) {   // rows the kernel never reaches read as NaN / -1
        SG_CUDA(cudaMemsetAsync(d_dbg_uv, 0xFF, dbg_bytes - Carver::padded(2 * sizeof(unsigned long long)), ctx->stream));
        SG_CUDA(cudaMemsetAsync(d_stats, 0, 2 * sizeof(unsigned long long), ctx->stream));
    }

    SG_TRY(proj_launch(ctx, common, fargs.data(), d_fargs, n_frames, max_nkp));

    // ---- download: results (and diagnostics) are contiguous on the device ----
    const size_t down_bytes = out_bytes + dbg_bytes;
    if (down_bytes && !zero_copy)
        SG_CUDA(cudaMemcpyAsync(h_out, d_out_point, down_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    SG_TRY(wait_stream(ctx));
    const uint8_t* hb = h_out;
    const size_t o_pad = Carver::padded(sizeof(int32_t) * total_kp);
    if (speculative) {
        unsigned int ties = 0;
        std::memcpy(&ties, hb + 2 * o_pad, sizeof(ties));
        ctx->spec_calls++;
        if (ties == 0 && ctx->spec_tie_streak > 0) ctx->spec_tie_streak--;
        if (ties != 0) {
            ctx->spec_tie_streak = 8;
            // some point's best distance was shared by two keypoints: only the KD visit order can say which one the
            // reference keeps (src/KDTree.cpp:52-82, src/MapMatcher.cpp:88).  Wait for the tree and repeat the call.
            ctx->spec_repeats++;
            for (int i = 0; i < n_frames; ++i) SG_TRY(finalize_frame(ctx, ctx->frames[frame_ids[i]]));
            return match_projection_impl(ctx, frame_ids, n_frames, set_id, poses, params, skip, order, n_order, out_point,
                                         out_dist, dbg_uv, dbg_kp, dbg_d, stats, matched_points, n_matched);
        }
    }
    if (total_kp) {
        std::memcpy(out_point, hb, sizeof(int32_t) * total_kp);
        std::memcpy(out_dist, hb + o_pad, sizeof(int32_t) * total_kp);
    }
    if (want_dbg) {
        const uint8_t* p = hb + 2 * o_pad + 256;
        if (dbg_uv) std::memcpy(dbg_uv, p, sizeof(float) * 2 * dbg_rows);
        p += Carver::padded(sizeof(float) * 2 * dbg_rows);
        if (dbg_kp) std::memcpy(dbg_kp, p, sizeof(int32_t) * dbg_rows);
        p += Carver::padded(sizeof(int32_t) * dbg_rows);
        if (dbg_d) std::memcpy(dbg_d, p, sizeof(int32_t) * dbg_rows);
        p += Carver::padded(sizeof(int32_t) * dbg_rows);
        if (stats) {
            unsigned long long s[2];
            std::memcpy(s, p, sizeof(s));
            stats[0] = static_cast<int64_t>(s[0]);
            stats[1] = static_cast<int64_t>(s[1]);
        }
    }
    return SLAMGPU_OK;
}

// Bytes of the KD image: [kd_xy][kd_idx][cell_start][cell_items][cell_rec], each padded like the device blob.
size_t kd_image_bytes(size_t n, size_t n_cells)
{
    return Carver::padded(n * sizeof(float2)) + Carver::padded(n * sizeof(int32_t)) +
           Carver::padded((n_cells + 1) * sizeof(int32_t)) + Carver::padded(n * sizeof(int32_t)) +
           Carver::padded(n * sizeof(int4));
}

// The part of the candidate grid that does not need the tree: [cell_start][cell_rec_spec], records in keypoint
// order within a cell, (x bits, y bits, keypoint index, keypoint index).  Built inside slamgpu_frame_upload (a
// counting sort, tens of microseconds) so that a match call made before the KD-tree is ready can run at once.
size_t grid_image_bytes(size_t n, size_t n_cells)
{
    return Carver::padded((n_cells + 1) * sizeof(int32_t)) + Carver::padded(n * sizeof(int4));
}
// floor() for the cell index: |v| is far below 2^31 here (the grid has at most 2^20 cells).  std::floor without
// -msse4.1 is a libm call, and two of them per keypoint made this the largest item of slamgpu_frame_upload (85 us
// for 8000 keypoints).
static inline int floor_to_int(float v)
{
    const int i = static_cast<int>(v);
    return v < static_cast<float>(i) ? i - 1 : i;
}

void build_grid_image(const float* xy, size_t n, float gx0, float gy0, float inv_cell, int gw, int gh, uint8_t* image,
                      std::vector<int32_t>& scratch)
{
    const size_t n_cells = static_cast<size_t>(gw) * gh;
    Carver hv(image);
    int32_t* h_cstart = hv.take<int32_t>(n_cells + 1);
    int4* h_crec = hv.take<int4>(n);
    std::memset(h_cstart, 0, (n_cells + 1) * sizeof(int32_t));
    scratch.resize(n + n_cells);                        // [cell of keypoint k][write cursor of cell c]
    int32_t* cell = scratch.data();
    int32_t* cursor = scratch.data() + n;
    for (size_t k = 0; k < n; ++k) {                    // the same expression as build_kd_image / proj_kernels.cu: grid_cell
        int cxi = floor_to_int((xy[2 * k] - gx0) * inv_cell);
        int cyi = floor_to_int((xy[2 * k + 1] - gy0) * inv_cell);
        cxi = std::min(std::max(cxi, 0), gw - 1);
        cyi = std::min(std::max(cyi, 0), gh - 1);
        cell[k] = cyi * gw + cxi;
        h_cstart[cell[k] + 1]++;
    }
    for (size_t c = 0; c < n_cells; ++c) {
        cursor[c] = h_cstart[c];
        h_cstart[c + 1] += h_cstart[c];
    }
    for (size_t k = 0; k < n; ++k) {
        int32_t xb, yb;
        std::memcpy(&xb, &xy[2 * k], 4);
        std::memcpy(&yb, &xy[2 * k + 1], 4);
        h_crec[cursor[cell[k]]++] = make_int4(xb, yb, static_cast<int32_t>(k), static_cast<int32_t>(k));
    }
}

// Everything the device keeps in KD order, from the permutation (host side; may run on the frame's worker thread).
void build_kd_image(FrameKdPending& k)
{
    const size_t n = k.xy.size() / 2;
    const size_t n_cells = static_cast<size_t>(k.gw) * k.gh;
    Carver hv(k.image.data());
    float* h_kdxy = reinterpret_cast<float*>(hv.take<float2>(n));
    int32_t* h_kdidx = hv.take<int32_t>(n);
    int32_t* h_cstart = hv.take<int32_t>(n_cells + 1);
    int32_t* h_citems = hv.take<int32_t>(n);
    int4* h_crec = hv.take<int4>(n);
    const float* xy = k.xy.data();
    for (size_t m = 0; m < n; ++m) {
        const uint32_t p = k.perm[m];
        h_kdxy[2 * m] = xy[2 * p];
        h_kdxy[2 * m + 1] = xy[2 * p + 1];
        h_kdidx[m] = static_cast<int32_t>(p);
    }
    // counting sort of the KD positions by cell; the device maps coordinates to cells with the
    // same float expression (proj_kernels.cu: grid_cell), and both are monotone in the coordinate
    std::memset(h_cstart, 0, (n_cells + 1) * sizeof(int32_t));
    auto cell_of = [&](float x, float y) {
        int cxi = floor_to_int((x - k.gx0) * k.inv_cell);
        int cyi = floor_to_int((y - k.gy0) * k.inv_cell);
        cxi = std::min(std::max(cxi, 0), k.gw - 1);
        cyi = std::min(std::max(cyi, 0), k.gh - 1);
        return static_cast<size_t>(cyi) * k.gw + cxi;
    };
    std::vector<int32_t> cell(n);
    for (size_t m = 0; m < n; ++m) {
        cell[m] = static_cast<int32_t>(cell_of(h_kdxy[2 * m], h_kdxy[2 * m + 1]));
        h_cstart[cell[m] + 1]++;
    }
    for (size_t c = 0; c < n_cells; ++c) h_cstart[c + 1] += h_cstart[c];
    {
        std::vector<int32_t> cursor(h_cstart, h_cstart + n_cells);
        for (size_t m = 0; m < n; ++m) h_citems[cursor[cell[m]]++] = static_cast<int32_t>(m);
    }
    // one 16-byte record per grid slot: (x bits, y bits, keypoint index, KD position) -- everything the
    // candidate test needs in a single load
    for (size_t i = 0; i < n; ++i) {
        const int32_t m = h_citems[i];
        int32_t xb, yb;
        std::memcpy(&xb, &h_kdxy[2 * m], 4);
        std::memcpy(&yb, &h_kdxy[2 * m + 1], 4);
        h_crec[i] = make_int4(xb, yb, h_kdidx[m], m);
    }
}

void join_frame(FrameDev& f)
{
    if (f.pending && f.pending->worker.joinable()) f.pending->worker.join();
}

// Make the frame's KD image resident: wait for the builder, one H2D copy.  Every entry point that reads the
// tree calls this first; it is a no-op once done.
int finalize_frame(slamgpu_ctx* ctx, FrameDev& f)
{
    if (!f.pending) return SLAMGPU_OK;
    join_frame(f);
    if (f.pending->lazy && f.pending->done.load(std::memory_order_acquire) == 0) {
        // the build was put off until somebody needs the tree: now
        FrameKdPending& k = *f.pending;
        const int nk = static_cast<int>(k.xy.size() / 2);
        k.perm.resize(static_cast<size_t>(nk));
        slamgpu_kd_build_host(k.xy.data(), nk, k.perm.data());
        k.image.resize(f.kd_image_bytes);
        build_kd_image(k);
        k.done.store(1, std::memory_order_release);
    }
    f.perm_host = f.pending->perm;
    if (f.kd_image_bytes) {
        SG_TRY(ensure_host_stage(ctx, f.kd_image_bytes + 256));
        SG_CUDA(cudaStreamSynchronize(ctx->stream));            // staging is reused
        std::memcpy(ctx->h_stage.ptr, f.pending->image.data(), f.kd_image_bytes);
        SG_CUDA(cudaMemcpyAsync(f.kd_image, ctx->h_stage.ptr, f.kd_image_bytes, cudaMemcpyHostToDevice, ctx->stream));
        SG_TRY(wait_stream(ctx));
    }
    f.pending.reset();
    return SLAMGPU_OK;
}

// Before a match call: make the KD image resident if its builder has finished; if it is still running, leave the
// frame pending and let the call run SPECULATIVELY on the tree-free grid (ties counted, not settled) -- unless a
// kernel form that walks the tree has been asked for (SLAMGPU_PROJ_WALK / SLAMGPU_PROJ_V2) or SLAMGPU_KD_SPECULATE=0.
int settle_frame(slamgpu_ctx* ctx, FrameDev& f, bool* speculative)
{
    if (!f.pending) return SLAMGPU_OK;
    const bool done = f.pending->done.load(std::memory_order_acquire) != 0;
    if (done || !ctx->kd_speculate || ctx->proj_walk || ctx->proj_v2) return finalize_frame(ctx, f);
    *speculative = true;
    return SLAMGPU_OK;
}

} // namespace

extern "C" {

int slamgpu_frame_upload(slamgpu_ctx* ctx, int frame_id, int nkp, const float* xy, const uint8_t* desc,
                         const uint32_t* kd_perm, const uint8_t* matched)
{
    SG_CTX(ctx);
    SG_CHECK(nkp >= 0 && nkp < (1 << 26), "keypoint count out of range");
    SG_CHECK(nkp == 0 || (xy != nullptr && desc != nullptr), "null keypoints / descriptors");
    const size_t n = static_cast<size_t>(nkp);

    std::vector<uint32_t> perm;
    if (kd_perm) {
        perm.resize(n);
        std::vector<uint8_t> seen(n, 0);
        for (int i = 0; i < nkp; ++i) {
            SG_CHECK(kd_perm[i] < static_cast<uint32_t>(nkp) && !seen[kd_perm[i]], "kd_perm is not a permutation");
            seen[kd_perm[i]] = 1;
            perm[i] = kd_perm[i];
        }
    }

    FrameDev f;
    f.nkp = nkp;
    // uniform grid over the keypoints' bounding box: 16 px cells (coarser when that would exceed 1 Mi cells)
    float cell = ctx->grid_cell;
    if (nkp > 0) {
        // bounding box and finiteness in one branch-free pass (vectorises: 8000 keypoints in ~3 us instead of ~19 with
        // a test and an early return per coordinate).  x - x is 0 for a finite x and NaN for NaN / +-inf.
        float x0 = xy[0], x1 = xy[0], y0 = xy[1], y1 = xy[1], bad = 0.0f;
        for (int i = 0; i < nkp; ++i) {
            const float x = xy[2 * i], y = xy[2 * i + 1];
            x0 = x < x0 ? x : x0; x1 = x > x1 ? x : x1;
            y0 = y < y0 ? y : y0; y1 = y > y1 ? y : y1;
            bad += (x - x) + (y - y);
        }
        SG_CHECK(bad == 0.0f, "non-finite keypoint coordinate");
        f.gx0 = x0;
        f.gy0 = y0;
        for (;;) {
            const double gw = std::floor((double(x1) - x0) / cell) + 1.0, gh = std::floor((double(y1) - y0) / cell) + 1.0;
            if (gw * gh <= double(1 << 20)) {
                f.gw = static_cast<int>(gw);
                f.gh = static_cast<int>(gh);
                break;
            }
            cell *= 2.0f;
        }
        f.inv_cell = 1.0f / cell;
    }

    // a slot that is uploaded again (the tracker's current frame, every frame) keeps its allocation
    // when it is large enough: cudaFree + cudaMalloc cost more than the upload itself
    struct Held {   // freed on every early return unless the new frame adopts it
        void* p = nullptr;
        ~Held() { if (p) cudaFree(p); }
    } held;
    void*& old_blob = held.p;
    size_t old_bytes = 0;
    auto it = ctx->frames.find(frame_id);
    if (it != ctx->frames.end()) {
        SG_CUDA(cudaStreamSynchronize(ctx->stream));
        join_frame(it->second);                       // a KD build still running on the old contents
        old_blob = it->second.blob;
        old_bytes = it->second.blob_bytes;
        it->second.blob = nullptr;
        ctx->frames.erase(it);
    }

    // device blob: [desc][matched][xy][grid without the tree: cell_start', cell_rec_spec] (uploaded now)
    // [kd_xy][kd_idx][cell_start][cell_items][cell_rec] (the KD image, uploaded when the tree is built)
    // [proposals] (device only).  cell_start' == cell_start (the grid does not depend on the tree); the kernels
    // read the copy that came with the records they use.
    const size_t n_cells = static_cast<size_t>(f.gw) * f.gh;
    const size_t b_desc = Carver::padded(n * 32), b_matched = Carver::padded(n), b_xy = Carver::padded(n * sizeof(float2));
    const size_t b_grid = grid_image_bytes(n, n_cells);
    const size_t now_bytes = b_desc + b_matched + b_xy + b_grid;
    const size_t kd_bytes = kd_image_bytes(n, n_cells);
    const size_t total = now_bytes + kd_bytes + Carver::padded(n * sizeof(unsigned long long)) + 256;
    if (old_blob && old_bytes >= total) {
        f.blob = old_blob;
        f.blob_bytes = old_bytes;
        old_blob = nullptr;
    } else {
        if (old_blob) cudaFree(old_blob);
        old_blob = nullptr;
        const cudaError_t e = cudaMalloc(&f.blob, total);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(SLAMGPU_ERR_NOMEM, std::string("frame allocation failed: ") + cudaGetErrorString(e));
        }
        f.blob_bytes = total;
    }
    struct NewBlob {   // freed on every error return below; disarmed when the frame is stored (on failure the
        FrameDev* f;   // slot is left absent rather than half-built)
        ~NewBlob() { if (f && f->blob) cudaFree(f->blob); }
    } new_blob{&f};
    Carver dv(f.blob);
    f.desc = dv.take<uint4>(2 * n);
    f.matched = dv.take<uint8_t>(n);
    f.xy = dv.take<float2>(n);
    f.cell_start_spec = dv.take<int32_t>(n_cells + 1);
    f.cell_rec_spec = dv.take<int4>(n);
    f.kd_image = dv.base + dv.off;
    f.kd_xy = dv.take<float2>(n);
    f.kd_idx = dv.take<int32_t>(n);
    f.cell_start = dv.take<int32_t>(n_cells + 1);
    f.cell_items = dv.take<int32_t>(n);
    f.cell_rec = dv.take<int4>(n);
    f.proposals = dv.take<unsigned long long>(n);
    f.kd_image_bytes = kd_bytes;

    if (nkp > 0) {
        SG_TRY(ensure_host_stage(ctx, std::max(now_bytes, kd_bytes) + 256));
        // staging is free to overwrite only after earlier copies out of it have completed
        SG_CUDA(cudaStreamSynchronize(ctx->stream));
        Carver hv(ctx->h_stage.ptr);
        uint8_t* h_desc = reinterpret_cast<uint8_t*>(hv.take<uint4>(2 * n));
        uint8_t* h_matched = hv.take<uint8_t>(n);
        float* h_xy = reinterpret_cast<float*>(hv.take<float2>(n));
        std::memcpy(h_desc, desc, n * 32);
        if (matched) {
            for (size_t i = 0; i < n; ++i) h_matched[i] = matched[i] ? 1 : 0;
        } else {
            std::memset(h_matched, 0, n);
        }
        std::memcpy(h_xy, xy, n * sizeof(float2));
        // the tree-free grid: one device launch behind the upload for frames of ordinary size, else a host counting sort
        const bool grid_on_device = ctx->grid_on_device && n <= (size_t(1) << 16) && n_cells * sizeof(int) <= (size_t(200) << 10);
        if (!grid_on_device)
            build_grid_image(xy, n, f.gx0, f.gy0, f.inv_cell, f.gw, f.gh, static_cast<uint8_t*>(ctx->h_stage.ptr) + hv.off, ctx->grid_scratch);
        SG_CUDA(cudaMemcpyAsync(f.blob, ctx->h_stage.ptr, grid_on_device ? now_bytes - b_grid : now_bytes, cudaMemcpyHostToDevice,
                                ctx->stream));
        SG_CUDA(cudaEventRecord(ctx->sync_event, ctx->stream));     // the copy out of the staging buffer: all this call waits for
        if (grid_on_device) {
            bool built = false;
            SG_TRY(proj_launch_grid_build(ctx, f.xy, nkp, f.gx0, f.gy0, f.inv_cell, f.gw, f.gh, f.cell_start_spec, f.cell_rec_spec,
                                          &built));
            SG_CHECK(built, "grid build refused a frame it was sized for");
        }
        SG_CUDA(cudaMemsetAsync(f.proposals, 0xFF, n * sizeof(unsigned long long), ctx->stream));
        // the staging buffer is reused by the next call: wait for the copy out of it (the grid build and the reset of
        // the proposal slots run on behind it; whatever uses them is stream-ordered)
        SG_CUDA(cudaEventSynchronize(ctx->sync_event));

        // The KD-tree and everything ordered by it.  With the caller's permutation (or SLAMGPU_KD_ASYNC=0) it is
        // built here; otherwise on a host thread that the first call needing the tree joins, so that the build
        // (the largest item of a tracked frame) overlaps whatever the caller does in between -- in the reference
        // KLT tracking and pose seeding sit between Frame construction and matching (src/Tracker.cpp:66-88).
        // LAZY: with speculation on, almost no call ever needs the tree (only a tie at some point's minimum does), so
        // it is not built at all until one does -- unless recent calls did hit ties (then the builder thread is
        // started as before, to have the tree ready when the repeat comes).
        const bool lazy = !kd_perm && ctx->kd_async && ctx->kd_speculate && ctx->kd_lazy && ctx->spec_tie_streak == 0 &&
                          !ctx->proj_walk && !ctx->proj_v2;
        auto pending = std::make_shared<FrameKdPending>();
        pending->xy.assign(xy, xy + 2 * n);
        pending->lazy = lazy;
        if (!lazy) pending->image.resize(kd_bytes);
        pending->gx0 = f.gx0; pending->gy0 = f.gy0; pending->inv_cell = f.inv_cell; pending->gw = f.gw; pending->gh = f.gh;
        if (kd_perm) pending->perm = perm;
        const int delay_ms = ctx->kd_delay_ms;
        auto work = [delay_ms](FrameKdPending* k, bool have_perm) {
            if (delay_ms > 0) std::this_thread::sleep_for(std::chrono::milliseconds(delay_ms));   // tests: force speculation
            const int nk = static_cast<int>(k->xy.size() / 2);
            if (!have_perm) {
                k->perm.resize(static_cast<size_t>(nk));
                slamgpu_kd_build_host(k->xy.data(), nk, k->perm.data());
            }
            build_kd_image(*k);
            k->done.store(1, std::memory_order_release);
        };
        f.pending = pending;
        if (kd_perm || !ctx->kd_async) {
            work(pending.get(), kd_perm != nullptr);
            SG_TRY(finalize_frame(ctx, f));
        } else if (!lazy) {
            pending->worker = std::thread(work, pending.get(), false);
        }
    }
    new_blob.f = nullptr;
    ctx->frames[frame_id] = std::move(f);
    return SLAMGPU_OK;
}

int slamgpu_frame_set_matched(slamgpu_ctx* ctx, int frame_id, const uint8_t* matched)
{
    SG_CTX(ctx);
    auto it = ctx->frames.find(frame_id);
    if (it == ctx->frames.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown frame id");
    FrameDev& f = it->second;
    if (f.nkp == 0) return SLAMGPU_OK;
    const size_t n = static_cast<size_t>(f.nkp);
    if (matched == nullptr) {
        SG_CUDA(cudaMemsetAsync(f.matched, 0, n, ctx->stream));
        return SLAMGPU_OK;
    }
    SG_TRY(ensure_host_stage(ctx, n + 256));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    uint8_t* h = static_cast<uint8_t*>(ctx->h_stage.ptr);
    for (size_t i = 0; i < n; ++i) h[i] = matched[i] ? 1 : 0;
    SG_CUDA(cudaMemcpyAsync(f.matched, h, n, cudaMemcpyHostToDevice, ctx->stream));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    return SLAMGPU_OK;
}

int slamgpu_frame_release(slamgpu_ctx* ctx, int frame_id)
{
    SG_CTX(ctx);
    auto it = ctx->frames.find(frame_id);
    if (it == ctx->frames.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown frame id");
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    free_frame(it->second);
    ctx->frames.erase(it);
    return SLAMGPU_OK;
}

int slamgpu_frame_get_kd_perm(slamgpu_ctx* ctx, int frame_id, uint32_t* out_perm)
{
    SG_CHECK(ctx != nullptr, "null context");
    auto it = ctx->frames.find(frame_id);
    if (it == ctx->frames.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown frame id");
    SG_CHECK(it->second.nkp == 0 || out_perm != nullptr, "null output");
    SG_CUDA(cudaSetDevice(ctx->device));
    SG_TRY(finalize_frame(ctx, it->second));
    std::copy(it->second.perm_host.begin(), it->second.perm_host.end(), out_perm);
    return SLAMGPU_OK;
}

int slamgpu_points_upload(slamgpu_ctx* ctx, int set_id, int n, const float* pos, const float* normal,
                          const float* range, const int32_t* obs_offset, const uint8_t* obs_desc,
                          const float* obs_center)
{
    SG_CTX(ctx);
    SG_CHECK(n >= 0, "negative point count");
    SG_CHECK(n == 0 || (pos != nullptr && obs_offset != nullptr), "null positions / observation offsets");
    const bool derive = (normal == nullptr || range == nullptr);
    SG_CHECK(n == 0 || !derive || obs_center != nullptr, "normal/range missing and no obs_center to derive them");
    int64_t n_obs = 0;
    if (n > 0) {
        SG_CHECK(obs_offset[0] == 0, "obs_offset[0] must be 0");
        for (int i = 0; i < n; ++i) SG_CHECK(obs_offset[i + 1] >= obs_offset[i], "obs_offset must be non-decreasing");
        n_obs = obs_offset[n];
        SG_CHECK(n_obs == 0 || obs_desc != nullptr, "null observation descriptors");
    }

    auto it = ctx->point_sets.find(set_id);
    if (it != ctx->point_sets.end()) {
        SG_CUDA(cudaStreamSynchronize(ctx->stream));
        free_points(it->second);
        ctx->point_sets.erase(it);
    }
    PointSetDev p;
    p.n = n;
    p.n_obs = n_obs;
    const size_t N = static_cast<size_t>(n), O = static_cast<size_t>(n_obs);
    const size_t up = points_image_bytes(N, O);
    const size_t b_center = derive ? Carver::padded(O * 12) : 0;
    SG_TRY(alloc_points(p, n, n_obs));
    if (n == 0) cudaMemsetAsync(p.obs_offset, 0, sizeof(int32_t), ctx->stream);   // slamgpu_points_append starts here

    if (n > 0) {
        int rc = ensure_host_stage(ctx, up + b_center + 256);
        if (rc == SLAMGPU_OK && derive) rc = ensure_device_io(ctx, b_center + 256);
        if (rc != SLAMGPU_OK) {
            cudaFree(p.blob);
            return rc;
        }
        SG_CUDA(cudaStreamSynchronize(ctx->stream));
        Carver hv(ctx->h_stage.ptr);
        float* h_pos = hv.take<float>(N * 3);
        float* h_nrm = hv.take<float>(N * 3);
        float* h_rng = hv.take<float>(N * 2);
        int32_t* h_off = hv.take<int32_t>(N + 1);
        uint8_t* h_desc = reinterpret_cast<uint8_t*>(hv.take<uint4>(O * 2));
        std::memcpy(h_pos, pos, N * 12);
        if (!derive) {
            std::memcpy(h_nrm, normal, N * 12);
            std::memcpy(h_rng, range, N * 8);
        }
        std::memcpy(h_off, obs_offset, (N + 1) * 4);
        if (O) std::memcpy(h_desc, obs_desc, O * 32);
        SG_CUDA(cudaMemcpyAsync(p.blob, ctx->h_stage.ptr, up, cudaMemcpyHostToDevice, ctx->stream));
        if (derive) {
            float* h_center = hv.take<float>(O * 3);
            if (O) std::memcpy(h_center, obs_center, O * 12);
            float* d_center = static_cast<float*>(ctx->d_io.ptr);
            if (O) SG_CUDA(cudaMemcpyAsync(d_center, h_center, O * 12, cudaMemcpyHostToDevice, ctx->stream));
            rc = proj_launch_point_stats(ctx, p.pos, n, p.obs_offset, d_center, p.normal, p.range);
            if (rc != SLAMGPU_OK) {
                cudaFree(p.blob);
                return rc;
            }
        }
        SG_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    ctx->point_sets[set_id] = p;
    return SLAMGPU_OK;
}

int slamgpu_points_append(slamgpu_ctx* ctx, int set_id, int n_new, const float* pos, const float* normal,
                          const float* range, const int32_t* obs_offset, const uint8_t* obs_desc,
                          const float* obs_center)
{
    SG_CTX(ctx);
    SG_CHECK(n_new >= 0, "negative point count");
    auto it = ctx->point_sets.find(set_id);
    if (it == ctx->point_sets.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown point set id");
    SG_CHECK(!it->second.is_mutable, "a mutable point set grows through slamgpu_points_append_observed");
    if (n_new == 0) return SLAMGPU_OK;
    SG_CHECK(pos != nullptr && obs_offset != nullptr, "null positions / observation offsets");
    const bool derive = (normal == nullptr || range == nullptr);
    SG_CHECK(!derive || obs_center != nullptr, "normal/range missing and no obs_center to derive them");
    SG_CHECK(obs_offset[0] == 0, "obs_offset[0] must be 0 (offsets are relative to the appended block)");
    for (int i = 0; i < n_new; ++i) SG_CHECK(obs_offset[i + 1] >= obs_offset[i], "obs_offset must be non-decreasing");
    const int64_t o_new = obs_offset[n_new];
    SG_CHECK(o_new == 0 || obs_desc != nullptr, "null observation descriptors");
    PointSetDev& p = it->second;
    SG_CHECK(static_cast<int64_t>(p.n) + n_new <= (1 << 30) && p.n_obs + o_new <= 0x7FFFFFFF, "point set too large");

    // grow geometrically: new allocation, device-to-device copies of the live part, old one freed.  A set that
    // is appended to is a map: its first growth reserves room for 256k points / 1M observations (45 MB of the
    // 180 GB; SLAMGPU_MAP_RESERVE = points), so a tracking run of BASELINE C3's size never reallocates (each cudaMalloc + cudaFree pair
    // stalls the frame for milliseconds).
    if (p.n + n_new > p.cap_n || p.n_obs + o_new > p.cap_obs) {
        PointSetDev g = p;
        const int cap_n = std::max({p.n + n_new, 2 * p.cap_n, ctx->map_reserve});
        const int64_t cap_obs = std::max<int64_t>({p.n_obs + o_new, 2 * p.cap_obs, int64_t(4) * ctx->map_reserve});
        SG_TRY(alloc_points(g, cap_n, cap_obs));
        const size_t N = static_cast<size_t>(p.n), O = static_cast<size_t>(p.n_obs);
        cudaError_t e = cudaSuccess;
        if (p.blob) {
            if (N) e = cudaMemcpyAsync(g.pos, p.pos, N * 12, cudaMemcpyDeviceToDevice, ctx->stream);
            if (N && e == cudaSuccess) e = cudaMemcpyAsync(g.normal, p.normal, N * 12, cudaMemcpyDeviceToDevice, ctx->stream);
            if (N && e == cudaSuccess) e = cudaMemcpyAsync(g.range, p.range, N * 8, cudaMemcpyDeviceToDevice, ctx->stream);
            if (e == cudaSuccess) e = cudaMemcpyAsync(g.obs_offset, p.obs_offset, (N + 1) * 4, cudaMemcpyDeviceToDevice, ctx->stream);
            if (O && e == cudaSuccess) e = cudaMemcpyAsync(g.obs_desc, p.obs_desc, O * 32, cudaMemcpyDeviceToDevice, ctx->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        } else {
            e = cudaMemsetAsync(g.obs_offset, 0, 4, ctx->stream);
        }
        if (e != cudaSuccess) {
            cudaFree(g.blob);
            return fail(SLAMGPU_ERR_CUDA, std::string("point-set growth: ") + cudaGetErrorString(e));
        }
        if (p.blob) cudaFree(p.blob);
        p = g;
    }

    const size_t M = static_cast<size_t>(n_new), Q = static_cast<size_t>(o_new);
    const size_t image = points_image_bytes(M, Q);
    const size_t b_center = derive ? Carver::padded(Q * 12) : 0;
    SG_TRY(ensure_host_stage(ctx, image + b_center + 256));
    if (derive) SG_TRY(ensure_device_io(ctx, b_center + 256));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));   // staging is reused
    Carver hv(ctx->h_stage.ptr);
    float* h_pos = hv.take<float>(M * 3);
    float* h_nrm = hv.take<float>(M * 3);
    float* h_rng = hv.take<float>(M * 2);
    int32_t* h_off = hv.take<int32_t>(M + 1);
    uint8_t* h_desc = reinterpret_cast<uint8_t*>(hv.take<uint4>(Q * 2));
    std::memcpy(h_pos, pos, M * 12);
    const int32_t base = static_cast<int32_t>(p.n_obs);
    for (size_t i = 0; i < M; ++i) h_off[i] = base + obs_offset[i + 1];   // entries n+1 .. n+n_new
    if (Q) std::memcpy(h_desc, obs_desc, Q * 32);
    const size_t n0 = static_cast<size_t>(p.n), o0 = static_cast<size_t>(p.n_obs);
    SG_CUDA(cudaMemcpyAsync(p.pos + 3 * n0, h_pos, M * 12, cudaMemcpyHostToDevice, ctx->stream));
    SG_CUDA(cudaMemcpyAsync(p.obs_offset + n0 + 1, h_off, M * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (Q) SG_CUDA(cudaMemcpyAsync(p.obs_desc + 2 * o0, h_desc, Q * 32, cudaMemcpyHostToDevice, ctx->stream));
    if (!derive) {
        std::memcpy(h_nrm, normal, M * 12);
        std::memcpy(h_rng, range, M * 8);
        SG_CUDA(cudaMemcpyAsync(p.normal + 3 * n0, h_nrm, M * 12, cudaMemcpyHostToDevice, ctx->stream));
        SG_CUDA(cudaMemcpyAsync(p.range + 2 * n0, h_rng, M * 8, cudaMemcpyHostToDevice, ctx->stream));
    } else {
        float* h_center = hv.take<float>(Q * 3);
        if (Q) std::memcpy(h_center, obs_center, Q * 12);
        float* d_center = static_cast<float*>(ctx->d_io.ptr);
        if (Q) SG_CUDA(cudaMemcpyAsync(d_center, h_center, Q * 12, cudaMemcpyHostToDevice, ctx->stream));
        // the kernel indexes observations by their absolute offset: bias the centre pointer by the old total
        SG_TRY(proj_launch_point_stats(ctx, p.pos + 3 * n0, n_new, p.obs_offset + n0, d_center - 3 * o0,
                                       p.normal + 3 * n0, p.range + 2 * n0));
    }
    SG_TRY(wait_stream(ctx));
    p.n += n_new;
    p.n_obs += o_new;
    return SLAMGPU_OK;
}

int slamgpu_speculation_count(slamgpu_ctx* ctx, int64_t* calls, int64_t* repeats)
{
    SG_CHECK(ctx != nullptr, "null context");
    if (calls) *calls = ctx->spec_calls;
    if (repeats) *repeats = ctx->spec_repeats;
    return SLAMGPU_OK;
}

int slamgpu_points_count(slamgpu_ctx* ctx, int set_id, int* n_points, int64_t* n_observations)
{
    SG_CHECK(ctx != nullptr, "null context");
    auto it = ctx->point_sets.find(set_id);
    if (it == ctx->point_sets.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown point set id");
    if (n_points) *n_points = it->second.n;
    if (n_observations) *n_observations = it->second.n_obs;
    return SLAMGPU_OK;
}

int slamgpu_points_release(slamgpu_ctx* ctx, int set_id)
{
    SG_CTX(ctx);
    auto it = ctx->point_sets.find(set_id);
    if (it == ctx->point_sets.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown point set id");
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    free_points(it->second);
    ctx->point_sets.erase(it);
    return SLAMGPU_OK;
}

int slamgpu_points_get_stats(slamgpu_ctx* ctx, int set_id, float* out_normal, float* out_range)
{
    SG_CTX(ctx);
    auto it = ctx->point_sets.find(set_id);
    if (it == ctx->point_sets.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown point set id");
    const PointSetDev& p = it->second;
    if (p.n == 0) return SLAMGPU_OK;
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    if (out_normal) SG_CUDA(cudaMemcpy(out_normal, p.normal, sizeof(float) * 3 * p.n, cudaMemcpyDeviceToHost));
    if (out_range) SG_CUDA(cudaMemcpy(out_range, p.range, sizeof(float) * 2 * p.n, cudaMemcpyDeviceToHost));
    return SLAMGPU_OK;
}

int slamgpu_match_projection(slamgpu_ctx* ctx, int frame_id, int set_id, const float* pose_colmajor16,
                             const slamgpu_proj_params* params, const uint8_t* skip, const int32_t* order,
                             int n_order, int32_t* out_point, int32_t* out_dist)
{
    return match_projection_impl(ctx, &frame_id, 1, set_id, pose_colmajor16, params, skip, order, n_order, out_point,
                                 out_dist, nullptr, nullptr, nullptr, nullptr);
}

int slamgpu_match_projection_ex(slamgpu_ctx* ctx, int frame_id, int set_id, const float* pose_colmajor16,
                                const slamgpu_proj_params* params, const uint8_t* skip, const int32_t* order,
                                int n_order, int32_t* out_point, int32_t* out_dist, float* dbg_uv, int32_t* dbg_kp,
                                int32_t* dbg_d, int64_t* stats)
{
    return match_projection_impl(ctx, &frame_id, 1, set_id, pose_colmajor16, params, skip, order, n_order, out_point,
                                 out_dist, dbg_uv, dbg_kp, dbg_d, stats);
}

int slamgpu_match_projection_lists(slamgpu_ctx* ctx, int frame_id, int set_id, const float* pose_colmajor16,
                                   const slamgpu_proj_params* params, const int32_t* matched_points, int n_matched,
                                   const int32_t* order, int n_order, int32_t* out_point, int32_t* out_dist)
{
    static const int32_t none = 0;
    return match_projection_impl(ctx, &frame_id, 1, set_id, pose_colmajor16, params, nullptr, order, n_order, out_point,
                                 out_dist, nullptr, nullptr, nullptr, nullptr, matched_points ? matched_points : &none,
                                 n_matched);
}

int slamgpu_track_match(slamgpu_ctx* ctx, int frame_id, int set_id, const float* pose_colmajor16,
                        const slamgpu_proj_params* params, const int32_t* matched_points, int n_matched,
                        const int32_t* kf_points, int n_kf, int32_t* out_kf_point, int32_t* out_kf_dist,
                        int32_t* out_map_point, int32_t* out_map_dist)
{
    SG_CTX(ctx);
    SG_TRY(check_params(params));
    SG_CHECK(pose_colmajor16 != nullptr, "null pose");
    SG_CHECK(n_matched >= 0 && n_kf >= 0, "negative list length");
    SG_CHECK((n_matched == 0 || matched_points) && (n_kf == 0 || kf_points), "null index list");
    auto f_it = ctx->frames.find(frame_id);
    if (f_it == ctx->frames.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown frame id");
    auto ps_it = ctx->point_sets.find(set_id);
    if (ps_it == ctx->point_sets.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown point set id");
    bool speculative = false;
    SG_TRY(settle_frame(ctx, f_it->second, &speculative));      // the KD-tree may still be building on its host thread
    const FrameDev& fd = f_it->second;
    const PointSetDev& ps = ps_it->second;
    for (int i = 0; i < n_matched; ++i)
        SG_CHECK(matched_points[i] >= 0 && matched_points[i] < ps.n, "matched_points entry outside the point set");
    for (int i = 0; i < n_kf; ++i) {
        SG_CHECK(kf_points[i] >= 0 && kf_points[i] < ps.n, "kf_points entry outside the point set");
        SG_CHECK(i == 0 || kf_points[i] > kf_points[i - 1], "kf_points must be strictly ascending (map order)");
    }
    const size_t nkp = static_cast<size_t>(fd.nkp);
    SG_CHECK(nkp == 0 || (out_kf_point && out_kf_dist && out_map_point && out_map_dist), "null output");

    const size_t b_m = Carver::padded(sizeof(int32_t) * n_matched), b_k = Carver::padded(sizeof(int32_t) * n_kf);
    const size_t b_out = Carver::padded(sizeof(int32_t) * (4 * nkp + 1));     // + the tie counter of a speculative call
    SG_TRY(ensure_device_io(ctx, b_m + b_k + b_out + Carver::padded(nkp) + 256));
    SG_TRY(ensure_host_stage(ctx, b_m + b_k + b_out + 256));
    Carver dio(ctx->d_io.ptr), hio(ctx->h_stage.ptr);
    int32_t* d_m = dio.take<int32_t>(n_matched);
    int32_t* d_k = dio.take<int32_t>(n_kf);
    int32_t* d_out = dio.take<int32_t>(4 * nkp + 1);
    unsigned int* d_ties = reinterpret_cast<unsigned int*>(d_out + 4 * nkp);
    uint8_t* d_flags_backup = dio.take<uint8_t>(nkp);
    int32_t* h_m = hio.take<int32_t>(n_matched);
    int32_t* h_k = hio.take<int32_t>(n_kf);
    int32_t* h_out = hio.take<int32_t>(4 * nkp + 1);       // the results' host image, behind the uploaded lists
    // ZERO COPY for the results of a frame of ordinary size: the two emit kernels write the four arrays (and the
    // tie count) straight into the pinned image -- no download to queue behind the kernels
    const bool zero_copy = ctx->proj_zero_copy && nkp > 0 && b_out <= (size_t(1) << 20);
    int32_t* k_out = zero_copy ? h_out : d_out;            // where the kernels write
    if (n_matched) std::memcpy(h_m, matched_points, sizeof(int32_t) * n_matched);
    if (n_kf) std::memcpy(h_k, kf_points, sizeof(int32_t) * n_kf);
    if (b_m + b_k) SG_CUDA(cudaMemcpyAsync(ctx->d_io.ptr, ctx->h_stage.ptr, b_m + b_k, cudaMemcpyHostToDevice, ctx->stream));

    ProjCommonArgs common;
    fill_common(common, ps, *params, nullptr, ps.n);
    ProjFrameArgs fa;
    fill_frame(fa, fd, pose_colmajor16, *params);
    fa.ties = d_ties;
    if (speculative) {
        // the call rewrites the frame's keypoint flags (Frame::add_map_match on the device): keep the old ones in
        // case the speculation has to be repeated with the tree
        SG_CUDA(cudaMemsetAsync(d_ties, 0, sizeof(unsigned int), ctx->stream));
        if (nkp) SG_CUDA(cudaMemcpyAsync(d_flags_backup, fd.matched, nkp, cudaMemcpyDeviceToDevice, ctx->stream));
        if (zero_copy) fa.ties_out = reinterpret_cast<unsigned int*>(h_out + 4 * nkp);
    }
    SG_TRY(proj_launch_track(ctx, common, fa, ps.pt_matched, fd.matched, d_m, n_matched, d_k, n_kf, k_out));
    if (nkp && !zero_copy) SG_CUDA(cudaMemcpyAsync(h_out, d_out, sizeof(int32_t) * (4 * nkp + 1), cudaMemcpyDeviceToHost, ctx->stream));
    SG_TRY(wait_stream(ctx));
    if (speculative && nkp) {
        unsigned int ties = 0;
        std::memcpy(&ties, h_out + 4 * nkp, sizeof(ties));
        ctx->spec_calls++;
        if (ties == 0 && ctx->spec_tie_streak > 0) ctx->spec_tie_streak--;
        if (ties != 0) {      // a tie at some point's minimum: the KD visit order decides; wait for the tree, repeat
            ctx->spec_tie_streak = 8;
            ctx->spec_repeats++;
            SG_TRY(finalize_frame(ctx, f_it->second));
            SG_CUDA(cudaMemcpyAsync(fd.matched, d_flags_backup, nkp, cudaMemcpyDeviceToDevice, ctx->stream));
            fill_frame(fa, fd, pose_colmajor16, *params);
            SG_TRY(proj_launch_track(ctx, common, fa, ps.pt_matched, fd.matched, d_m, n_matched, d_k, n_kf, k_out));
            if (!zero_copy) SG_CUDA(cudaMemcpyAsync(h_out, d_out, sizeof(int32_t) * 4 * nkp, cudaMemcpyDeviceToHost, ctx->stream));
            SG_TRY(wait_stream(ctx));
        }
    }
    if (nkp) {
        const int32_t* h = h_out;
        std::memcpy(out_kf_point, h, sizeof(int32_t) * nkp);
        std::memcpy(out_kf_dist, h + nkp, sizeof(int32_t) * nkp);
        std::memcpy(out_map_point, h + 2 * nkp, sizeof(int32_t) * nkp);
        std::memcpy(out_map_dist, h + 3 * nkp, sizeof(int32_t) * nkp);
    }
    return SLAMGPU_OK;
}

int slamgpu_match_projection_batch(slamgpu_ctx* ctx, const int* frame_ids, int n_frames, int set_id,
                                   const float* poses_colmajor16, const slamgpu_proj_params* params,
                                   const uint8_t* skip, const int32_t* order, int n_order, int32_t* out_point,
                                   int32_t* out_dist)
{
    return match_projection_impl(ctx, frame_ids, n_frames, set_id, poses_colmajor16, params, skip, order, n_order,
                                 out_point, out_dist, nullptr, nullptr, nullptr, nullptr);
}

} // extern "C"
