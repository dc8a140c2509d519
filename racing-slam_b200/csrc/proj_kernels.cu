// proj_kernels.cu -- projection-guided map-point search for sm_100a.
//
// Replaces match_via_reproject (src/MapMatcher.cpp:45-98) and its drivers (:107-127,165-175):
//   K4  proj_match_kernel   one thread per visited map point: Camera::project (src/Camera.cpp:25-32),
//                           is_in_image (:34-37), viewing-angle and distance gates
//                           (src/MapMatcher.cpp:62-73), KDTree2D::radius_search in the reference's
//                           visit order (src/KDTree.cpp:52-82), candidates x observations Hamming
//                           (src/MapMatcher.cpp:80-93), then a 64-bit atomicMin of
//                           (distance << 32 | visit rank) on the chosen keypoint (:95-97)
//   K5  proj_emit_kernel    per keypoint: decode the winning proposal (accepted_matches, :34-43)
//   N2  point_stats_kernel  avg_viewing_normal / observed_distance_range (src/MapPoint.cpp:24-45)
//
// Float parity: every float operation is an explicit round-to-nearest intrinsic in the oracle's
// order, so nvcc can never contract a multiply-add into an FMA (the reference builds without
// FMA: CMakeLists.txt:1-7).  Roofline: HBM/L2 bytes, in practice latency.
#include "slamgpu_internal.cuh"

namespace slamgpu {

namespace {

constexpr int kProjThreads = 128;
constexpr int kStackDepth = 40;   // KD depth <= log2(nkp) + 1; far-children pending <= depth

__device__ __forceinline__ float sum3(float a0, float a1, float a2) { return __fadd_rn(a0, __fadd_rn(a1, a2)); }

// Eigen normalized(): z = squaredNorm(); z > 0 ? v / sqrt(z) : v
__device__ __forceinline__ void normalized3(const float v[3], float out[3])
{
    const float z = sum3(__fmul_rn(v[0], v[0]), __fmul_rn(v[1], v[1]), __fmul_rn(v[2], v[2]));
    if (z > 0.0f) {
        const float n = __fsqrt_rn(z);
        out[0] = __fdiv_rn(v[0], n);
        out[1] = __fdiv_rn(v[1], n);
        out[2] = __fdiv_rn(v[2], n);
    } else {
        out[0] = v[0]; out[1] = v[1]; out[2] = v[2];
    }
}
__device__ __forceinline__ float norm3(const float v[3])
{
    return __fsqrt_rn(sum3(__fmul_rn(v[0], v[0]), __fmul_rn(v[1], v[1]), __fmul_rn(v[2], v[2])));
}

__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// A point that passed P0-P4, as the gate pass of the two-pass form hands it to the search pass.
struct ProjSurvivor {
    float u, w;
    int v, o_begin;
    int o_end, cx0, cx1, cy0;
    int n_rows, pad0, pad1, pad2;
};
static_assert(sizeof(ProjSurvivor) == 48, "three 16-byte loads");

struct ProjKernelArgs {
    ProjCommonArgs c;
    ProjFrameArgs inline_frame;       // used when frames == nullptr (single-frame call: no upload)
    const ProjFrameArgs* frames;      // device array for the batched form
    ProjSurvivor* surv = nullptr;     // two-pass form: [n_frames][n_visit] survivors and their count per frame
    int* surv_count = nullptr;
};

__global__ void __launch_bounds__(kProjThreads) proj_match_kernel(const ProjKernelArgs a)
{
    const ProjFrameArgs& f = a.frames ? a.frames[blockIdx.y] : a.inline_frame;
    const ProjCommonArgs& c = a.c;
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= c.n_visit) return;
    // debug rows are per (frame, visit)
    const size_t dbg_row = static_cast<size_t>(blockIdx.y) * c.n_visit + v;
    if (c.dbg_uv) {
        c.dbg_uv[2 * dbg_row] = __int_as_float(0x7FC00000);
        c.dbg_uv[2 * dbg_row + 1] = __int_as_float(0x7FC00000);
    }
    if (c.dbg_kp) c.dbg_kp[dbg_row] = -1;
    if (c.dbg_d) c.dbg_d[dbg_row] = -1;

    const int pi = c.order ? c.order[v] : v;
    if (pi < 0) return;                               // match_for_fuse nullptr, src/MapMatcher.cpp:119-121
    if (f.skip && f.skip[pi]) return;                 // :53 is_matched(point) / :169 !is_observed_by

    const float X[3] = {c.pos[3 * pi], c.pos[3 * pi + 1], c.pos[3 * pi + 2]};

    // P1  Camera::project: M[:, :3] * X summed a0 + (a1 + a2), then + M[:, 3]; true division
    float h[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        h[i] = sum3(__fmul_rn(f.M[4 * i], X[0]), __fmul_rn(f.M[4 * i + 1], X[1]), __fmul_rn(f.M[4 * i + 2], X[2]));
        h[i] = __fadd_rn(h[i], f.M[4 * i + 3]);
    }
    float u, w;
    if (h[2] < 0.0f) {
        u = -1.0f; w = -1.0f;                          // src/Camera.cpp:28-30
    } else {
        u = __fdiv_rn(h[0], h[2]);
        w = __fdiv_rn(h[1], h[2]);
    }
    if (c.dbg_uv) {
        c.dbg_uv[2 * dbg_row] = u;
        c.dbg_uv[2 * dbg_row + 1] = w;
    }
    // P2  is_in_image (NaN fails every comparison, as on the host)
    if (!(u >= 0.0f && u < c.width_f && w >= 0.0f && w < c.height_f)) return;
    if (c.stats) {
        if (fabsf(u) < 1e-4f || fabsf(__fsub_rn(u, c.width_f)) < 1e-4f || fabsf(w) < 1e-4f ||
            fabsf(__fsub_rn(w, c.height_f)) < 1e-4f)
            atomicAdd(&c.stats[0], 1ull);
    }

    // P3  viewing-angle gate
    const float ray[3] = {__fsub_rn(X[0], f.C[0]), __fsub_rn(X[1], f.C[1]), __fsub_rn(X[2], f.C[2])};
    float rn[3];
    normalized3(ray, rn);
    const float nx = c.normal[3 * pi], ny = c.normal[3 * pi + 1], nz = c.normal[3 * pi + 2];
    if (sum3(__fmul_rn(nx, rn[0]), __fmul_rn(ny, rn[1]), __fmul_rn(nz, rn[2])) < c.min_cos) return;
    // P4  distance gate
    const float nearest = c.range[2 * pi], furthest = c.range[2 * pi + 1];
    const float distance = norm3(ray);
    if (distance < __fdiv_rn(nearest, c.nearer) || distance > __fmul_rn(furthest, c.further)) return;

    if (c.stats) {   // diagnostic only: keypoints within 1e-4 px of the search radius
        unsigned long long cnt = 0;
        for (int k = 0; k < f.nkp; ++k) {
            const float2 p = f.xy[k];
            const float dx = __fsub_rn(p.x, u), dy = __fsub_rn(p.y, w);
            const float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
            if (fabsf(__fsub_rn(d, c.radius)) < 1e-4f) cnt++;
        }
        if (cnt) atomicAdd(&c.stats[1], cnt);
    }

    // P5 + P6  KD radius search in reference visit order, Hamming against every observation
    const int o_begin = c.obs_offset[pi], o_end = c.obs_end[pi];
    int best = c.thr;          // accept d < thr  <=>  (double)d < max_descriptor_distance
    int best_kp = 0;

    int st_s[kStackDepth], st_e[kStackDepth];   // st_e packs (depth << 26 | end)
    int sp = 0;
    int s = 0, e = f.nkp, depth = 0;
    while (true) {
        while (s < e) {
            const int mid = (s + e) >> 1;                          // src/KDTree.cpp:25
            const float2 p = f.kd_xy[mid];
            const float dx = __fsub_rn(p.x, u);                    // :61
            const float dy = __fsub_rn(p.y, w);                    // :62
            const float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));   // :63
            if (d2 <= c.r2) {                                      // :65
                const int kp = f.kd_idx[mid];
                if (c.replace || !f.matched[kp]) {                 // src/MapMatcher.cpp:81
                    const uint4 a0 = __ldg(&f.desc[2 * kp]), a1 = __ldg(&f.desc[2 * kp + 1]);
                    for (int o = o_begin; o < o_end; ++o) {        // :86
                        const uint4 b0 = __ldg(&c.obs_desc[2 * static_cast<size_t>(o)]);
                        const uint4 b1 = __ldg(&c.obs_desc[2 * static_cast<size_t>(o) + 1]);
                        const int d = hamming256(a0, a1, b0, b1);
                        if (d < best) {                            // :88 strict: first in visit order wins
                            best = d;
                            best_kp = kp;
                        }
                    }
                }
            }
            const float delta = (depth & 1) ? dy : dx;             // :69-70
            const bool left_near = delta > 0.0f;                   // :73
            const int fs = left_near ? mid + 1 : s, fe = left_near ? e : mid;
            const int ns = left_near ? s : mid + 1, ne = left_near ? mid : e;
            if (__fmul_rn(delta, delta) <= c.r2 && fs < fe && sp < kStackDepth) {   // :79
                st_s[sp] = fs;
                st_e[sp] = ((depth + 1) << 26) | fe;
                sp++;
            }
            s = ns; e = ne; depth++;
        }
        if (sp == 0) break;
        sp--;
        s = st_s[sp];
        e = st_e[sp] & ((1 << 26) - 1);
        depth = st_e[sp] >> 26;
    }

    if (best < c.thr) {
        // P7  proposals[kp] = min over points of (distance, visit rank): the first point in visit
        // order attaining the minimum wins, as in the sequential loop (src/MapMatcher.cpp:95-97)
        const unsigned long long key = (static_cast<unsigned long long>(best) << 32) | static_cast<unsigned>(v);
        atomicMin(&f.proposals[best_kp], key);
        if (c.dbg_kp) c.dbg_kp[dbg_row] = best_kp;
        if (c.dbg_d) c.dbg_d[dbg_row] = best;
    }
}

// ---------------------------------------------------------------------------------------------
// K4, warp-per-point form (default).  ncu on the thread-per-point kernel above showed 5.4 of 32
// lanes active per instruction (every lane walks its own KD path) and 2 warps per SM
// (profiles/r01_ncu_proj_match_walk.txt).  Here one WARP owns one visited point:
//   * projection and gates are warp-uniform;
//   * candidates come from the frame's uniform grid (the cells the radius box touches are
//     contiguous per cell row), one candidate per lane, tested with the reference's own float
//     expression dx*dx + dy*dy <= r*r (src/KDTree.cpp:61-67) -- a KD radius search is exact, so
//     the candidate SET is the same;
//   * each lane takes the Hamming minimum of its candidate over the point's observations
//     (observation rows are warp-uniform loads);
//   * the winner is the minimum distance; ties are broken by the KD VISIT ORDER, recovered from
//     the tree itself: the path root -> candidate, one symbol per level (0 = the node itself,
//     1 = near child, 2 = far child, near/far decided by the sign of delta exactly as in
//     src/KDTree.cpp:69-76), compared lexicographically = pre-order "node, near subtree, far
//     subtree" = the order radius_search pushes results, so the strict '<' of
//     src/MapMatcher.cpp:88 keeps the same candidate.  Keys are only computed on ties.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long kd_visit_key(const float2* __restrict__ kd_xy, int n, int m, float u,
                                                           float w)
{
    unsigned long long key = 0;
    int s = 0, e = n, depth = 0;
    while (s < e && depth < 31) {
        const int mid = (s + e) >> 1;
        if (mid == m) break;                                     // symbol 0 and all-zero tail
        const float2 p = kd_xy[mid];
        const float delta = (depth & 1) ? __fsub_rn(p.y, w) : __fsub_rn(p.x, u);   // src/KDTree.cpp:61-62,69-70
        const bool left_near = delta > 0.0f;                     // :73
        const bool go_left = m < mid;
        key |= static_cast<unsigned long long>(go_left == left_near ? 1 : 2) << (62 - 2 * depth);
        if (go_left) e = mid; else s = mid + 1;
        depth++;
    }
    return key;
}

__device__ __forceinline__ int grid_cell(float x, float x0, float inv_cell, int n)
{
    const int c = static_cast<int>(floorf(__fmul_rn(__fsub_rn(x, x0), inv_cell)));
    return min(max(c, 0), n - 1);
}

constexpr int kGroupThreads = 128;

// LANES = 32: one warp per point (small calls: latency).  LANES = 8: four points per warp (large
// batches: fewer idle lanes).  All collectives carry the group's lane mask.
template <int LANES>
__global__ void __launch_bounds__(kGroupThreads) proj_match_warp_kernel(const ProjKernelArgs a)
{
    const ProjFrameArgs& f = a.frames ? a.frames[blockIdx.y] : a.inline_frame;
    const ProjCommonArgs& c = a.c;
    const int lane = threadIdx.x % LANES;
    const unsigned gmask = LANES == 32 ? 0xFFFFFFFFu : (((1u << LANES) - 1u) << ((threadIdx.x & 31) / LANES * LANES));
    const int v = blockIdx.x * (kGroupThreads / LANES) + threadIdx.x / LANES;
    if (v >= c.n_visit) return;
    const size_t dbg_row = static_cast<size_t>(blockIdx.y) * c.n_visit + v;

    const int pi = c.order ? c.order[v] : v;
    if (pi < 0) return;                               // match_for_fuse nullptr, src/MapMatcher.cpp:119-121
    if (f.skip && f.skip[pi]) return;                 // :53 is_matched(point) / :169 !is_observed_by

    const float X[3] = {c.pos[3 * pi], c.pos[3 * pi + 1], c.pos[3 * pi + 2]};
    float h[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {                     // P1, same order as the kernel above
        h[i] = sum3(__fmul_rn(f.M[4 * i], X[0]), __fmul_rn(f.M[4 * i + 1], X[1]), __fmul_rn(f.M[4 * i + 2], X[2]));
        h[i] = __fadd_rn(h[i], f.M[4 * i + 3]);
    }
    float u, w;
    if (h[2] < 0.0f) {
        u = -1.0f; w = -1.0f;
    } else {
        u = __fdiv_rn(h[0], h[2]);
        w = __fdiv_rn(h[1], h[2]);
    }
    if (c.dbg_uv && lane == 0) {
        c.dbg_uv[2 * dbg_row] = u;
        c.dbg_uv[2 * dbg_row + 1] = w;
    }
    if (!(u >= 0.0f && u < c.width_f && w >= 0.0f && w < c.height_f)) return;      // P2
    if (c.stats && lane == 0) {
        if (fabsf(u) < 1e-4f || fabsf(__fsub_rn(u, c.width_f)) < 1e-4f || fabsf(w) < 1e-4f ||
            fabsf(__fsub_rn(w, c.height_f)) < 1e-4f)
            atomicAdd(&c.stats[0], 1ull);
    }
    const float ray[3] = {__fsub_rn(X[0], f.C[0]), __fsub_rn(X[1], f.C[1]), __fsub_rn(X[2], f.C[2])};
    float rn[3];
    normalized3(ray, rn);
    const float nx = c.normal[3 * pi], ny = c.normal[3 * pi + 1], nz = c.normal[3 * pi + 2];
    if (sum3(__fmul_rn(nx, rn[0]), __fmul_rn(ny, rn[1]), __fmul_rn(nz, rn[2])) < c.min_cos) return;   // P3
    const float nearest = c.range[2 * pi], furthest = c.range[2 * pi + 1];
    const float distance = norm3(ray);
    if (distance < __fdiv_rn(nearest, c.nearer) || distance > __fmul_rn(furthest, c.further)) return;  // P4

    if (c.stats) {   // diagnostic only: keypoints within 1e-4 px of the search radius
        unsigned long long cnt = 0;
        for (int k = lane; k < f.nkp; k += LANES) {
            const float2 p = f.xy[k];
            const float dx = __fsub_rn(p.x, u), dy = __fsub_rn(p.y, w);
            const float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
            if (fabsf(__fsub_rn(d, c.radius)) < 1e-4f) cnt++;
        }
        if (cnt) atomicAdd(&c.stats[1], cnt);
    }

    // P5 + P6: candidates from the grid, one per lane
    const int o_begin = c.obs_offset[pi], o_end = c.obs_end[pi];
    const float pad = c.radius + 1.0f;                           // conservative box; the exact test is d2 <= r2
    const int cx0 = grid_cell(u - pad, f.gx0, f.inv_cell, f.gw), cx1 = grid_cell(u + pad, f.gx0, f.inv_cell, f.gw);
    const int cy0 = grid_cell(w - pad, f.gy0, f.inv_cell, f.gh), cy1 = grid_cell(w + pad, f.gy0, f.inv_cell, f.gh);
    int best = c.thr;                                            // lane-local best: accept d < thr
    int best_m = -1;
    for (int cy = cy0; cy <= cy1; ++cy) {
        const int i_begin = f.cell_start[cy * f.gw + cx0], i_end = f.cell_start[cy * f.gw + cx1 + 1];
        for (int i = i_begin + lane; i < i_end; i += LANES) {
            const int m = f.cell_items[i];
            const float2 p = f.kd_xy[m];
            const float dx = __fsub_rn(p.x, u);                  // src/KDTree.cpp:61
            const float dy = __fsub_rn(p.y, w);                  // :62
            const float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));   // :63
            if (!(d2 <= c.r2)) continue;                         // :65
            const int kp = f.kd_idx[m];
            if (!c.replace && f.matched[kp]) continue;           // src/MapMatcher.cpp:81
            const uint4 a0 = __ldg(&f.desc[2 * kp]), a1 = __ldg(&f.desc[2 * kp + 1]);
            int dmin = 1 << 20;
            for (int o = o_begin; o < o_end; ++o) {              // :86
                const uint4 b0 = __ldg(&c.obs_desc[2 * static_cast<size_t>(o)]);
                const uint4 b1 = __ldg(&c.obs_desc[2 * static_cast<size_t>(o) + 1]);
                dmin = min(dmin, hamming256(a0, a1, b0, b1));
            }
            if (dmin < best) {
                best = dmin;
                best_m = m;
            } else if (dmin == best && best_m >= 0) {            // tie inside one lane: earlier visit wins
                if (kd_visit_key(f.kd_xy, f.nkp, m, u, w) < kd_visit_key(f.kd_xy, f.nkp, best_m, u, w)) best_m = m;
            }
        }
    }
    // warp winner: minimum distance, then earliest KD visit
    const int wbest = __reduce_min_sync(gmask, best_m >= 0 ? best : (1 << 20));
    if (wbest >= c.thr) return;                                  // nothing under max_descriptor_distance
    const bool tied = best_m >= 0 && best == wbest;
    const unsigned ballot = __ballot_sync(gmask, tied);
    int win_lane = __ffs(ballot) - 1;                            // lane index within the warp
    if (__popc(ballot) > 1) {
        unsigned long long key = tied ? kd_visit_key(f.kd_xy, f.nkp, best_m, u, w) : ~0ull;
        // 64-bit warp minimum by two 32-bit reductions
        const unsigned hi = __reduce_min_sync(gmask, static_cast<unsigned>(key >> 32));
        const unsigned lo = __reduce_min_sync(gmask, (key >> 32) == hi ? static_cast<unsigned>(key) : 0xFFFFFFFFu);
        const unsigned long long wkey = (static_cast<unsigned long long>(hi) << 32) | lo;
        win_lane = __ffs(__ballot_sync(gmask, tied && key == wkey)) - 1;
    }
    if (static_cast<int>(threadIdx.x & 31) == win_lane) {
        const int best_kp = f.kd_idx[best_m];
        // P7: first point in visit order attaining the minimum wins (src/MapMatcher.cpp:95-97)
        const unsigned long long key = (static_cast<unsigned long long>(wbest) << 32) | static_cast<unsigned>(v);
        atomicMin(&f.proposals[best_kp], key);
        if (c.dbg_kp) c.dbg_kp[dbg_row] = best_kp;
        if (c.dbg_d) c.dbg_d[dbg_row] = wbest;
    }
}

// ---------------------------------------------------------------------------------------------
// K4, group-per-point form v3 (default).  ncu on the kernel above (profiles/r01_ncu_full_prof_proj_match_v2.txt):
// 0.18 instructions issued per scheduler cycle, 66 % of the stall cycles on L1TEX scoreboards -- the time is
// the DEPTH of the dependent-load chain (order -> point -> per cell row: cell_start -> cell_items -> kd_xy ->
// kd_idx -> matched -> desc -> observations), about 18 round trips through L2 for a three-row search.
// Same arithmetic, same candidates, same tie rule; the chain is cut to 5 round trips:
//   1. order[v]
//   2. everything indexed by the point at once (skip flag, position, normal, range, observation offsets)
//   3. the cell-row bounds, one row per lane, and the point's first observation descriptors (kept in
//      registers, then parked in shared memory for the group)
//   4. one 16-byte record per grid slot (x, y, keypoint index, KD position); the rows are flattened so
//      that the lanes share all slots evenly
//   5. the keypoint's matched flag and descriptor together
// ---------------------------------------------------------------------------------------------
constexpr int kObsStaged = 8;     // observation descriptors of a point parked in shared memory (the rest: global)

// The search half of K4 for one point and its group of LANES lanes (P5-P7): the point's staged observation
// descriptors go to shared memory, the cell rows [rb, re) (one row per lane) are flattened over the lanes, every
// slot is tested, the group's best candidate proposes.  Shared by the fused kernel below and by the two-pass
// form (proj_search_kernel).
template <int LANES, bool QUEUE>
__device__ __forceinline__ void group_search(const ProjFrameArgs& f, const ProjCommonArgs& c, const int lane, const int grp,
                                             const unsigned gmask, const float u, const float w, const int v,
                                             const size_t dbg_row, const int o_begin, const int o_end, const int n_staged,
                                             const int cx0, const int cx1, const int cy0, const int n_rows, int rb, int re,
                                             const uint4* stage, uint4 (*s_obs)[2 * kObsStaged], int2 (*s_queue)[2 * LANES])
{
    constexpr int STAGE_REGS = (2 * kObsStaged + LANES - 1) / LANES;
    if (c.stats) {   // diagnostic only: keypoints within 1e-4 px of the search radius
        unsigned long long cnt = 0;
        for (int k = lane; k < f.nkp; k += LANES) {
            const float2 p = f.xy[k];
            const float dx = __fsub_rn(p.x, u), dy = __fsub_rn(p.y, w);
            const float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
            if (fabsf(__fsub_rn(d, c.radius)) < 1e-4f) cnt++;
        }
        if (cnt) atomicAdd(&c.stats[1], cnt);
    }

#pragma unroll
    for (int r = 0; r < STAGE_REGS; ++r) {
        const int k = lane + r * LANES;
        if (k < 2 * kObsStaged) s_obs[grp][k] = stage[r];
    }
    __syncwarp(gmask);

    // P5 + P6: candidates from the grid, the rows flattened over the lanes
    int best = c.thr;                                            // lane-local best: accept d < thr
    int best_m = -1;
    bool lane_tie = false;                                       // speculative call only: this lane's best is shared
    // P6 for one candidate: every observation descriptor of the point against the keypoint's.  m = the candidate's
    // KD position, or -- in a speculative call made before the KD-tree has been built (f.kd_ready == 0) -- its
    // keypoint index: then ties cannot be settled, they are only COUNTED (f.ties), and the caller repeats the call
    // with the tree if the count is not zero.  Ties at the minimum are rare (two keypoints in the search radius
    // whose best Hamming distance to the point's observations is equal AND under the maximum distance).
    auto score = [&](const int kp, const int m) {
        const unsigned kp_matched = f.matched[kp];               // round trip 5: flag and descriptor together
        const uint4 a0 = __ldg(&f.desc[2 * kp]), a1 = __ldg(&f.desc[2 * kp + 1]);
        if (!c.replace && kp_matched) return;                    // src/MapMatcher.cpp:81
        int dmin = 1 << 20;
        for (int o = 0; o < n_staged; ++o)                       // :86
            dmin = min(dmin, hamming256(a0, a1, s_obs[grp][2 * o], s_obs[grp][2 * o + 1]));
        for (int o = o_begin + kObsStaged; o < o_end; ++o) {
            const uint4 b0 = __ldg(&c.obs_desc[2 * static_cast<size_t>(o)]);
            const uint4 b1 = __ldg(&c.obs_desc[2 * static_cast<size_t>(o) + 1]);
            dmin = min(dmin, hamming256(a0, a1, b0, b1));
        }
        if (dmin < best) {
            best = dmin;
            best_m = m;
            lane_tie = false;
        } else if (dmin == best && best_m >= 0) {                // tie inside one lane: earlier visit wins
            if (!f.kd_ready) lane_tie = true;                    // speculative call: the tree is not there yet
            else if (kd_visit_key(f.kd_xy, f.nkp, m, u, w) < kd_visit_key(f.kd_xy, f.nkp, best_m, u, w)) best_m = m;
        }
    };
    // QUEUE: the slots that pass the radius test (about 4 in 10 of the box) are packed into a per-group queue
    // and scored LANES at a time, so that the Hamming loop runs with full groups instead of 3 lanes in 8.
    // The order candidates are scored in does not matter: ties are settled by the KD visit key.
    const int gshift = (threadIdx.x & 31) / LANES * LANES;
    int qn = 0;                                                  // queue fill, the same in every lane of the group
    for (int row0 = 0; row0 < n_rows; row0 += LANES) {
        if (row0 > 0) {                                          // more rows than lanes (huge radius): next batch
            rb = re = 0;
            if (row0 + lane < n_rows) {
                rb = f.cell_start[(cy0 + row0 + lane) * f.gw + cx0];
                re = f.cell_start[(cy0 + row0 + lane) * f.gw + cx1 + 1];
            }
        }
        const int len = re - rb;
        int incl = len;                                          // inclusive prefix sum of the row lengths
#pragma unroll
        for (int d = 1; d < LANES; d <<= 1) {
            const int t = __shfl_up_sync(gmask, incl, d, LANES);
            if (lane >= d) incl += t;
        }
        const int total = __shfl_sync(gmask, incl, LANES - 1, LANES);
        const int excl = incl - len;                             // lanes past the last row hold `total`
        const int delta = rb - excl;
        for (int base = 0; base < total; base += LANES) {
            const int t = base + lane;
            // which row does flat index t fall in?  The last row whose exclusive prefix is <= t (an empty row
            // shares its prefix with the next one, which then wins): binary search over the lanes' prefixes
            int r = 0;
#pragma unroll
            for (int step = LANES / 2; step > 0; step >>= 1) {
                const int e = __shfl_sync(gmask, excl, r + step, LANES);
                if (t >= e) r += step;
            }
            const int slot = t + __shfl_sync(gmask, delta, r, LANES);
            bool pass = false;
            int kp = 0, m = 0;
            if (t < total) {
                // round trip 4.  (Prefetching the next batch's records here was measured slower: 80 registers
                // instead of 72 cost more occupancy than the overlap gained; profiles/r01_proj_variants_v3b.txt.)
                const int4 rec = __ldg(&f.cell_rec[slot]);
                const float dx = __fsub_rn(__int_as_float(rec.x), u);                // src/KDTree.cpp:61
                const float dy = __fsub_rn(__int_as_float(rec.y), w);                // :62
                const float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));    // :63
                pass = d2 <= c.r2;                               // :65
                kp = rec.z;
                m = f.kd_ready ? rec.w : rec.z;
            }
            if (!QUEUE) {
                if (pass) score(kp, m);
                continue;
            }
            const unsigned hits = (__ballot_sync(gmask, pass) >> gshift) & (LANES == 32 ? 0xFFFFFFFFu : ((1u << LANES) - 1u));
            if (pass) s_queue[grp][qn + __popc(hits & ((1u << lane) - 1u))] = make_int2(kp, m);
            qn += __popc(hits);
            __syncwarp(gmask);
            if (qn >= LANES) {                                   // a full group of candidates: score them
                const int2 cand = s_queue[grp][lane];
                const int2 left = s_queue[grp][LANES + lane];    // read before anyone overwrites
                __syncwarp(gmask);
                qn -= LANES;
                if (lane < qn) s_queue[grp][lane] = left;
                score(cand.x, cand.y);
                __syncwarp(gmask);
            }
        }
    }
    if (QUEUE && lane < qn) {
        const int2 cand = s_queue[grp][lane];
        score(cand.x, cand.y);
    }
    // group winner: minimum distance, then earliest KD visit
    const int wbest = __reduce_min_sync(gmask, best_m >= 0 ? best : (1 << 20));
    if (wbest >= c.thr) return;                                  // nothing under max_descriptor_distance
    const bool tied = best_m >= 0 && best == wbest;
    const unsigned ballot = __ballot_sync(gmask, tied);
    int win_lane = __ffs(ballot) - 1;                            // lane index within the warp
    if (!f.kd_ready) {                                           // speculative: count what the tree would have settled
        const unsigned lt = __ballot_sync(gmask, tied && lane_tie);
        if (static_cast<int>(threadIdx.x & 31) == win_lane && (__popc(ballot) > 1 || lt != 0u)) atomicAdd(f.ties, 1u);
    } else if (__popc(ballot) > 1) {
        unsigned long long key = tied ? kd_visit_key(f.kd_xy, f.nkp, best_m, u, w) : ~0ull;
        const unsigned hi = __reduce_min_sync(gmask, static_cast<unsigned>(key >> 32));
        const unsigned lo = __reduce_min_sync(gmask, (key >> 32) == hi ? static_cast<unsigned>(key) : 0xFFFFFFFFu);
        const unsigned long long wkey = (static_cast<unsigned long long>(hi) << 32) | lo;
        win_lane = __ffs(__ballot_sync(gmask, tied && key == wkey)) - 1;
    }
    if (static_cast<int>(threadIdx.x & 31) == win_lane) {
        const int best_kp = f.kd_ready ? f.kd_idx[best_m] : best_m;
        // P7: first point in visit order attaining the minimum wins (src/MapMatcher.cpp:95-97)
        const unsigned long long key = (static_cast<unsigned long long>(wbest) << 32) | static_cast<unsigned>(v);
        atomicMin(&f.proposals[best_kp], key);
        if (c.dbg_kp) c.dbg_kp[dbg_row] = best_kp;
        if (c.dbg_d) c.dbg_d[dbg_row] = wbest;
    }
}

template <int LANES, bool QUEUE>
__device__ __forceinline__ void proj_match_group_body(const ProjKernelArgs& a, const ProjFrameArgs& f,
                                                      uint4 (*s_obs)[2 * kObsStaged], int2 (*s_queue)[2 * LANES])
{
    constexpr int GROUPS = kGroupThreads / LANES;
    constexpr int STAGE_REGS = (2 * kObsStaged + LANES - 1) / LANES;
    const ProjCommonArgs& c = a.c;
    const int lane = threadIdx.x % LANES;
    const int grp = threadIdx.x / LANES;
    const unsigned gmask = LANES == 32 ? 0xFFFFFFFFu : (((1u << LANES) - 1u) << ((threadIdx.x & 31) / LANES * LANES));
    const int v = blockIdx.x * GROUPS + grp;
    if (v >= c.n_visit) return;
    const size_t dbg_row = static_cast<size_t>(blockIdx.y) * c.n_visit + v;

    const int pi = c.order ? c.order[v] : v;
    if (pi < 0) return;                               // match_for_fuse nullptr, src/MapMatcher.cpp:119-121
    // round trip 2: every load indexed by the point is issued before the first test
    const unsigned skipped = f.skip ? f.skip[pi] : 0u;
    const float X[3] = {c.pos[3 * pi], c.pos[3 * pi + 1], c.pos[3 * pi + 2]};
    const float nx = c.normal[3 * pi], ny = c.normal[3 * pi + 1], nz = c.normal[3 * pi + 2];
    const float nearest = c.range[2 * pi], furthest = c.range[2 * pi + 1];
    const int o_begin = c.obs_offset[pi], o_end = c.obs_end[pi];
    if (skipped) return;                              // :53 is_matched(point) / :169 !is_observed_by

    float h[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {                     // P1, same order as the kernels above
        h[i] = sum3(__fmul_rn(f.M[4 * i], X[0]), __fmul_rn(f.M[4 * i + 1], X[1]), __fmul_rn(f.M[4 * i + 2], X[2]));
        h[i] = __fadd_rn(h[i], f.M[4 * i + 3]);
    }
    float u, w;
    if (h[2] < 0.0f) {
        u = -1.0f; w = -1.0f;
    } else {
        u = __fdiv_rn(h[0], h[2]);
        w = __fdiv_rn(h[1], h[2]);
    }
    if (c.dbg_uv && lane == 0) {
        c.dbg_uv[2 * dbg_row] = u;
        c.dbg_uv[2 * dbg_row + 1] = w;
    }
    if (!(u >= 0.0f && u < c.width_f && w >= 0.0f && w < c.height_f)) return;      // P2

    // round trip 3a: the first observation descriptors, one or two 16-byte halves per lane
    const int n_obs = o_end - o_begin;
    const int n_staged = min(n_obs, kObsStaged);
    uint4 stage[STAGE_REGS];
#pragma unroll
    for (int r = 0; r < STAGE_REGS; ++r) {
        const int k = lane + r * LANES;
        stage[r] = k < 2 * n_staged ? __ldg(&c.obs_desc[2 * static_cast<size_t>(o_begin) + k]) : make_uint4(0, 0, 0, 0);
    }
    // round trip 3b: bounds of the cell rows the search box touches, one row per lane
    const float pad = c.radius + 1.0f;                           // conservative box; the exact test is d2 <= r2
    const int cx0 = grid_cell(u - pad, f.gx0, f.inv_cell, f.gw), cx1 = grid_cell(u + pad, f.gx0, f.inv_cell, f.gw);
    const int cy0 = grid_cell(w - pad, f.gy0, f.inv_cell, f.gh), cy1 = grid_cell(w + pad, f.gy0, f.inv_cell, f.gh);
    const int n_rows = cy1 - cy0 + 1;
    int rb = 0, re = 0;
    if (lane < n_rows) {
        rb = f.cell_start[(cy0 + lane) * f.gw + cx0];
        re = f.cell_start[(cy0 + lane) * f.gw + cx1 + 1];
    }

    if (c.stats && lane == 0) {
        if (fabsf(u) < 1e-4f || fabsf(__fsub_rn(u, c.width_f)) < 1e-4f || fabsf(w) < 1e-4f ||
            fabsf(__fsub_rn(w, c.height_f)) < 1e-4f)
            atomicAdd(&c.stats[0], 1ull);
    }
    const float ray[3] = {__fsub_rn(X[0], f.C[0]), __fsub_rn(X[1], f.C[1]), __fsub_rn(X[2], f.C[2])};
    float rn[3];
    normalized3(ray, rn);
    if (sum3(__fmul_rn(nx, rn[0]), __fmul_rn(ny, rn[1]), __fmul_rn(nz, rn[2])) < c.min_cos) return;   // P3
    const float distance = norm3(ray);
    if (distance < __fdiv_rn(nearest, c.nearer) || distance > __fmul_rn(furthest, c.further)) return;  // P4

    group_search<LANES, QUEUE>(f, c, lane, grp, gmask, u, w, v, dbg_row, o_begin, o_end, n_staged, cx0, cx1, cy0, n_rows,
                               rb, re, stage, s_obs, s_queue);
}

template <int LANES, bool QUEUE>
__global__ void __launch_bounds__(kGroupThreads) proj_match_group_kernel(const ProjKernelArgs a)
{
    constexpr int GROUPS = kGroupThreads / LANES;
    __shared__ uint4 s_obs[GROUPS][2 * kObsStaged];
    __shared__ int2 s_queue[QUEUE ? GROUPS : 1][2 * LANES];
    const ProjFrameArgs& f = a.frames ? a.frames[blockIdx.y] : a.inline_frame;
    proj_match_group_body<LANES, QUEUE>(a, f, s_obs, s_queue);
}
// (Letting the last CTA of this kernel run the emit step -- one launch instead of two for a per-frame call -- was
// tried and measured 2.5 us SLOWER on the C2 call than the separate 8-CTA emit launch: the tail is a chain of two
// round trips on one SM behind a fence and a counter in every CTA, and a dependent launch costs only 2.3 us.)

// ---------------------------------------------------------------------------------------------
// K4, two-pass form (large visit lists).  In the fused kernel every lane of a group repeats the point's
// projection and gates (IEEE divisions and square roots are long instruction sequences): one warp
// instruction serves 4 points, and on a 200k-point map the kernel issues on 56 % of the cycles
// (profiles/r01_ncu_full_proj_match_group_200k.txt).  Here the gate pass runs one THREAD per point (one warp
// instruction serves 32 points, the point arrays are read coalesced) and compacts the survivors; the search
// pass gives every survivor a group of lanes and runs the same group_search.  Same arithmetic, same
// candidates, same tie rule; the visit rank travels with the survivor, so the order of the compaction cannot
// change a result.  Small visit lists stay on the fused kernel (one launch less, shorter dependent chain).
// ---------------------------------------------------------------------------------------------
constexpr int kGateThreads = 256;

__global__ void __launch_bounds__(kGateThreads) proj_gate_kernel(const ProjKernelArgs a)
{
    const ProjFrameArgs& f = a.frames ? a.frames[blockIdx.y] : a.inline_frame;
    const ProjCommonArgs& c = a.c;
    const int v = blockIdx.x * kGateThreads + threadIdx.x;
    bool alive = v < c.n_visit;
    int pi = -1;
    if (alive) {
        pi = c.order ? c.order[v] : v;
        alive = pi >= 0;                                         // match_for_fuse nullptr, src/MapMatcher.cpp:119-121
    }
    ProjSurvivor s;
    s.u = s.w = 0.0f;
    s.v = v;
    s.o_begin = s.o_end = s.cx0 = s.cx1 = s.cy0 = s.n_rows = s.pad0 = s.pad1 = s.pad2 = 0;
    if (alive) {
        const unsigned skipped = f.skip ? f.skip[pi] : 0u;
        const float X[3] = {c.pos[3 * pi], c.pos[3 * pi + 1], c.pos[3 * pi + 2]};
        const float nx = c.normal[3 * pi], ny = c.normal[3 * pi + 1], nz = c.normal[3 * pi + 2];
        const float nearest = c.range[2 * pi], furthest = c.range[2 * pi + 1];
        s.o_begin = c.obs_offset[pi];
        s.o_end = c.obs_end[pi];
        alive = !skipped;                                        // :53 is_matched(point) / :169 !is_observed_by
        float h[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {                            // P1, same order as the kernels above
            h[i] = sum3(__fmul_rn(f.M[4 * i], X[0]), __fmul_rn(f.M[4 * i + 1], X[1]), __fmul_rn(f.M[4 * i + 2], X[2]));
            h[i] = __fadd_rn(h[i], f.M[4 * i + 3]);
        }
        float u, w;
        if (h[2] < 0.0f) {
            u = -1.0f; w = -1.0f;
        } else {
            u = __fdiv_rn(h[0], h[2]);
            w = __fdiv_rn(h[1], h[2]);
        }
        s.u = u;
        s.w = w;
        if (alive && c.dbg_uv) {
            const size_t dbg_row = static_cast<size_t>(blockIdx.y) * c.n_visit + v;
            c.dbg_uv[2 * dbg_row] = u;
            c.dbg_uv[2 * dbg_row + 1] = w;
        }
        alive = alive && (u >= 0.0f && u < c.width_f && w >= 0.0f && w < c.height_f);      // P2
        if (alive) {
            if (c.stats) {
                if (fabsf(u) < 1e-4f || fabsf(__fsub_rn(u, c.width_f)) < 1e-4f || fabsf(w) < 1e-4f ||
                    fabsf(__fsub_rn(w, c.height_f)) < 1e-4f)
                    atomicAdd(&c.stats[0], 1ull);
            }
            const float ray[3] = {__fsub_rn(X[0], f.C[0]), __fsub_rn(X[1], f.C[1]), __fsub_rn(X[2], f.C[2])};
            float rn[3];
            normalized3(ray, rn);
            if (sum3(__fmul_rn(nx, rn[0]), __fmul_rn(ny, rn[1]), __fmul_rn(nz, rn[2])) < c.min_cos) alive = false;   // P3
            const float distance = norm3(ray);
            if (distance < __fdiv_rn(nearest, c.nearer) || distance > __fmul_rn(furthest, c.further)) alive = false;  // P4
            const float pad = c.radius + 1.0f;                   // conservative box; the exact test is d2 <= r2
            s.cx0 = grid_cell(u - pad, f.gx0, f.inv_cell, f.gw);
            s.cx1 = grid_cell(u + pad, f.gx0, f.inv_cell, f.gw);
            s.cy0 = grid_cell(w - pad, f.gy0, f.inv_cell, f.gh);
            s.n_rows = grid_cell(w + pad, f.gy0, f.inv_cell, f.gh) - s.cy0 + 1;
        }
    }
    // warp-aggregated append to this frame's survivor list
    const unsigned live = __ballot_sync(0xFFFFFFFFu, alive);
    if (live == 0u) return;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(live) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(&a.surv_count[blockIdx.y], __popc(live));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    if (alive) {
        ProjSurvivor* dst = a.surv + static_cast<size_t>(blockIdx.y) * c.n_visit + base + __popc(live & ((1u << lane) - 1u));
        uint4* d4 = reinterpret_cast<uint4*>(dst);
        const uint4* s4 = reinterpret_cast<const uint4*>(&s);
        d4[0] = s4[0];
        d4[1] = s4[1];
        d4[2] = s4[2];
    }
}

// 8 CTAs per SM (64 registers, 20 bytes spilled) instead of the 7 that 72 registers allow: the pass is
// latency-bound (issue slots 48 % busy at 37 % occupancy), C4 274 -> 255 us, 200k map 69 -> 65.5 us.
// Measured and dropped: prefetch.global.L1 of a candidate's descriptor and flag when it is queued
// (C4 260 -> 353 us).
template <int LANES, bool QUEUE>
__global__ void __launch_bounds__(kGroupThreads, 8) proj_search_kernel(const ProjKernelArgs a)
{
    constexpr int GROUPS = kGroupThreads / LANES;
    constexpr int STAGE_REGS = (2 * kObsStaged + LANES - 1) / LANES;
    __shared__ uint4 s_obs[GROUPS][2 * kObsStaged];
    __shared__ int2 s_queue[QUEUE ? GROUPS : 1][2 * LANES];
    const ProjFrameArgs& f = a.frames ? a.frames[blockIdx.y] : a.inline_frame;
    const ProjCommonArgs& c = a.c;
    const int lane = threadIdx.x % LANES;
    const int grp = threadIdx.x / LANES;
    const unsigned gmask = LANES == 32 ? 0xFFFFFFFFu : (((1u << LANES) - 1u) << ((threadIdx.x & 31) / LANES * LANES));
    const int idx = blockIdx.x * GROUPS + grp;
    if (idx >= a.surv_count[blockIdx.y]) return;
    const uint4* s4 = reinterpret_cast<const uint4*>(a.surv + static_cast<size_t>(blockIdx.y) * c.n_visit + idx);
    const uint4 r0 = __ldcg(s4), r1 = __ldcg(s4 + 1), r2 = __ldcg(s4 + 2);
    const float u = __uint_as_float(r0.x), w = __uint_as_float(r0.y);
    const int v = static_cast<int>(r0.z), o_begin = static_cast<int>(r0.w), o_end = static_cast<int>(r1.x);
    const int cx0 = static_cast<int>(r1.y), cx1 = static_cast<int>(r1.z), cy0 = static_cast<int>(r1.w);
    const int n_rows = static_cast<int>(r2.x);
    const size_t dbg_row = static_cast<size_t>(blockIdx.y) * c.n_visit + v;
    const int n_staged = min(o_end - o_begin, kObsStaged);
    uint4 stage[STAGE_REGS];
#pragma unroll
    for (int r = 0; r < STAGE_REGS; ++r) {
        const int k = lane + r * LANES;
        stage[r] = k < 2 * n_staged ? __ldg(&c.obs_desc[2 * static_cast<size_t>(o_begin) + k]) : make_uint4(0, 0, 0, 0);
    }
    int rb = 0, re = 0;
    if (lane < n_rows) {
        rb = f.cell_start[(cy0 + lane) * f.gw + cx0];
        re = f.cell_start[(cy0 + lane) * f.gw + cx1 + 1];
    }
    group_search<LANES, QUEUE>(f, c, lane, grp, gmask, u, w, v, dbg_row, o_begin, o_end, n_staged, cx0, cx1, cy0, n_rows,
                               rb, re, stage, s_obs, s_queue);
}

// Measured and dropped (profiles/r01_proj_variants_v4.txt): a cooperative form in which every lane first
// projects and gates its own point and the group then walks its survivors one by one.  It removes the 8x
// redundant projection arithmetic but was no faster on 200k points (103 vs 87 us) and 2.5x slower on 10k:
// the kernel is bound by the latency of its dependent loads at the occupancy its registers allow, not by
// instruction issue.

__device__ __forceinline__ void emit_keypoint(const ProjKernelArgs& a, const ProjFrameArgs& f, int kp)
{
    const unsigned long long key = __ldcg(&f.proposals[kp]);          // written by atomics of other CTAs: read at L2
    int point = -1, dist = -1;
    if (key != ~0ull) {
        const int v = static_cast<int>(key & 0xFFFFFFFFull);
        point = a.c.order ? a.c.order[v] : v;
        dist = static_cast<int>(key >> 32);
        f.proposals[kp] = ~0ull;          // the slots are all-ones at rest (set at upload): no reset pass before a call
    }
    f.out_point[kp] = point;
    f.out_dist[kp] = dist;
    if (kp == 0 && f.ties_out != nullptr) *f.ties_out = __ldcg(f.ties);      // zero-copy call: the tie count to host memory
}

__global__ void proj_emit_kernel(const ProjKernelArgs a)
{
    const ProjFrameArgs& f = a.frames ? a.frames[blockIdx.y] : a.inline_frame;
    const int kp = blockIdx.x * blockDim.x + threadIdx.x;
    if (kp >= f.nkp) return;
    emit_keypoint(a, f, kp);
}

// Emit + Frame::add_map_match (src/Frame.cpp:80-102) on the device.  A proposal comes from a point that
// is not yet matched in the frame and lands on a keypoint that is not yet matched (replace = 0), and a
// point proposes to one keypoint only, so `previous == nullptr` and the sweep of :90-98 finds nothing:
// the update is two flags.  The proposal slot is reset for the next stage.
__global__ void proj_emit_apply_kernel(const ProjKernelArgs a, uint8_t* pt_matched, uint8_t* kp_matched)
{
    const ProjFrameArgs& f = a.inline_frame;
    const int kp = blockIdx.x * blockDim.x + threadIdx.x;
    if (kp >= f.nkp) return;
    const unsigned long long key = f.proposals[kp];
    int point = -1, dist = -1;
    if (key != ~0ull) {
        const int v = static_cast<int>(key & 0xFFFFFFFFull);
        point = a.c.order ? a.c.order[v] : v;
        dist = static_cast<int>(key >> 32);
        kp_matched[kp] = 1;
        pt_matched[point] = 1;
        f.proposals[kp] = ~0ull;
    }
    f.out_point[kp] = point;
    f.out_dist[kp] = dist;
    if (kp == 0 && f.ties_out != nullptr) *f.ties_out = *f.ties;      // zero-copy call: the tie count so far to host memory
}

__global__ void proj_scatter_flags_kernel(const int32_t* idx, int n, uint8_t* flags)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[idx[i]] = 1;
}

// src/MapPoint.cpp:24-45, one thread per point, observations in the host's iteration order.
__global__ void point_stats_kernel(const float* pos, int n, const int32_t* obs_offset, const float* obs_center,
                                   float* normal, float* range)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float X[3] = {pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]};
    float acc[3] = {0.0f, 0.0f, 0.0f};
    float nearest = 3.402823466e+38f, furthest = 0.0f;
    for (int o = obs_offset[i]; o < obs_offset[i + 1]; ++o) {
        const float* C = obs_center + 3 * static_cast<size_t>(o);
        const float d[3] = {__fsub_rn(X[0], C[0]), __fsub_rn(X[1], C[1]), __fsub_rn(X[2], C[2])};
        float v[3], vv[3];
        normalized3(d, v);
        normalized3(v, vv);
        acc[0] = __fadd_rn(acc[0], vv[0]);
        acc[1] = __fadd_rn(acc[1], vv[1]);
        acc[2] = __fadd_rn(acc[2], vv[2]);
        const float dist = norm3(d);
        nearest = dist < nearest ? dist : nearest;     // std::min(nearest, dist)
        furthest = furthest < dist ? dist : furthest;  // std::max(furthest, dist)
    }
    float nrm[3];
    normalized3(acc, nrm);
    normal[3 * i] = nrm[0]; normal[3 * i + 1] = nrm[1]; normal[3 * i + 2] = nrm[2];
    range[2 * i] = nearest;
    range[2 * i + 1] = furthest;
}

__global__ void kd_gather_kernel(const float2* xy, const int32_t* perm, int n, float2* kd_xy)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) kd_xy[i] = xy[perm[i]];
}

} // namespace

// K4 launch: kernel variant by context switches (see the comments on the kernels).
static int launch_match(slamgpu_ctx* ctx, const ProjKernelArgs& a, int n_visit, int n_frames)
{
    ProfScope prof(ctx, SLAMGPU_KERNEL_PROJ_MATCH);
    if (ctx->proj_walk) {   // thread-per-point KD walk, kept for A/B checks (SLAMGPU_PROJ_WALK=1)
        const dim3 gp((n_visit + kProjThreads - 1) / kProjThreads, n_frames);
        proj_match_kernel<<<gp, kProjThreads, 0, ctx->stream>>>(a);
        ctx->launches++;
        return SLAMGPU_OK;
    }
    // 8 lanes per point: as fast as a full warp on C2 (16 us) and 1.8x faster on C4-sized batches; 4 lanes
    // are no faster on large maps and slower on small ones (tools/time_proj.py,
    // profiles/r01_proj_variants_v3*.txt); SLAMGPU_PROJ_LANES=32 overrides
    const int lanes = ctx->proj_lanes ? ctx->proj_lanes : 8;
    const int per_cta = kGroupThreads / lanes;
    const dim3 gp((n_visit + per_cta - 1) / per_cta, n_frames);
    const long long visits = static_cast<long long>(n_visit) * n_frames;
    const bool two_pass = !ctx->proj_v2 && (ctx->proj_two_pass == 1 || (ctx->proj_two_pass == 0 && visits >= 32768));
    if (two_pass) {
        const size_t b_surv = Carver::padded(static_cast<size_t>(visits) * sizeof(ProjSurvivor));
        SG_TRY(ensure_device_ws(ctx, b_surv + Carver::padded(sizeof(int) * n_frames) + 256));
        ProjKernelArgs b = a;
        Carver ws(ctx->d_ws.ptr);
        b.surv = ws.take<ProjSurvivor>(static_cast<size_t>(visits));
        b.surv_count = ws.take<int>(n_frames);
        SG_CUDA(cudaMemsetAsync(b.surv_count, 0, sizeof(int) * n_frames, ctx->stream));
        const dim3 gg((n_visit + kGateThreads - 1) / kGateThreads, n_frames);
        proj_gate_kernel<<<gg, kGateThreads, 0, ctx->stream>>>(b);
        if (lanes == 8 && ctx->proj_queue) proj_search_kernel<8, true><<<gp, kGroupThreads, 0, ctx->stream>>>(b);
        else if (lanes == 8) proj_search_kernel<8, false><<<gp, kGroupThreads, 0, ctx->stream>>>(b);
        else if (ctx->proj_queue) proj_search_kernel<32, true><<<gp, kGroupThreads, 0, ctx->stream>>>(b);
        else proj_search_kernel<32, false><<<gp, kGroupThreads, 0, ctx->stream>>>(b);
        ctx->launches += 2;
        return SLAMGPU_OK;
    }
    if (ctx->proj_v2) {             // previous form, kept for A/B checks (SLAMGPU_PROJ_V2=1)
        if (lanes == 8) proj_match_warp_kernel<8><<<gp, kGroupThreads, 0, ctx->stream>>>(a);
        else proj_match_warp_kernel<32><<<gp, kGroupThreads, 0, ctx->stream>>>(a);
    } else if (lanes == 8) {
        if (ctx->proj_queue) proj_match_group_kernel<8, true><<<gp, kGroupThreads, 0, ctx->stream>>>(a);
        else proj_match_group_kernel<8, false><<<gp, kGroupThreads, 0, ctx->stream>>>(a);
    } else {
        if (ctx->proj_queue) proj_match_group_kernel<32, true><<<gp, kGroupThreads, 0, ctx->stream>>>(a);
        else proj_match_group_kernel<32, false><<<gp, kGroupThreads, 0, ctx->stream>>>(a);
    }
    ctx->launches++;
    return SLAMGPU_OK;
}

int proj_launch(slamgpu_ctx* ctx, const ProjCommonArgs& common, const ProjFrameArgs* h_frames,
                const ProjFrameArgs* d_frames, int n_frames, int max_nkp)
{
    if (n_frames <= 0) return SLAMGPU_OK;
    ProjKernelArgs a;
    a.c = common;
    a.inline_frame = h_frames[0];
    a.frames = n_frames == 1 ? nullptr : d_frames;
    SG_CHECK(n_frames <= 65535, "too many frames in one projection batch");
    if (max_nkp > 0) {
        const dim3 gk((max_nkp + 255) / 256, n_frames);
        if (common.n_visit > 0) SG_TRY(launch_match(ctx, a, common.n_visit, n_frames));
        proj_emit_kernel<<<gk, 256, 0, ctx->stream>>>(a);
        ctx->launches++;
    }
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int proj_launch_track(slamgpu_ctx* ctx, const ProjCommonArgs& common, const ProjFrameArgs& frame, uint8_t* d_pt_matched,
                      uint8_t* d_kp_matched, const int32_t* d_matched_points, int n_matched, const int32_t* d_kf_order,
                      int n_kf, int32_t* d_out)
{
    const int nkp = frame.nkp, n = common.n_points;
    if (n > 0) SG_CUDA(cudaMemsetAsync(d_pt_matched, 0, static_cast<size_t>(n), ctx->stream));
    if (n_matched > 0) {
        proj_scatter_flags_kernel<<<(n_matched + 255) / 256, 256, 0, ctx->stream>>>(d_matched_points, n_matched,
                                                                                    d_pt_matched);
        ctx->launches++;
    }
    if (nkp <= 0) return SLAMGPU_OK;
    ProjKernelArgs a;
    a.c = common;
    a.c.replace = 0;                     // match(): is_matched(keypoint) candidates are skipped, src/MapMatcher.cpp:173
    a.inline_frame = frame;
    a.inline_frame.skip = d_pt_matched;  // frame.is_matched(point), :53
    a.frames = nullptr;
    const dim3 gk((nkp + 255) / 256, 1);
    // stage 1: match_key_frame -- the points the last key frame observes, in map order (:169)
    a.c.order = d_kf_order;
    a.c.n_visit = n_kf;
    a.inline_frame.out_point = d_out;
    a.inline_frame.out_dist = d_out + nkp;
    if (n_kf > 0) SG_TRY(launch_match(ctx, a, n_kf, 1));
    proj_emit_apply_kernel<<<gk, 256, 0, ctx->stream>>>(a, d_pt_matched, d_kp_matched);
    ctx->launches++;
    // stage 2: match_map -- every point (:107-110); stage 1's matches are already in the masks
    a.c.order = nullptr;
    a.c.n_visit = n;
    a.inline_frame.out_point = d_out + 2 * nkp;
    a.inline_frame.out_dist = d_out + 3 * nkp;
    if (n > 0) SG_TRY(launch_match(ctx, a, n, 1));
    proj_emit_apply_kernel<<<gk, 256, 0, ctx->stream>>>(a, d_pt_matched, d_kp_matched);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

// flags[0..n_flags) = 0, then flags[idx[i]] = 1: frame.is_matched(point) from a sparse list (src/MapMatcher.cpp:53)
int proj_launch_flags(slamgpu_ctx* ctx, const int32_t* d_idx, int n, uint8_t* d_flags, int n_flags)
{
    if (n_flags > 0) SG_CUDA(cudaMemsetAsync(d_flags, 0, static_cast<size_t>(n_flags), ctx->stream));
    if (n > 0) {
        proj_scatter_flags_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(d_idx, n, d_flags);
        ctx->launches++;
        SG_CUDA(cudaGetLastError());
    }
    return SLAMGPU_OK;
}

int proj_launch_point_stats(slamgpu_ctx* ctx, const float* d_pos, int n, const int32_t* d_obs_offset,
                            const float* d_obs_center, float* d_normal, float* d_range)
{
    if (n <= 0) return SLAMGPU_OK;
    point_stats_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(d_pos, n, d_obs_offset, d_obs_center, d_normal,
                                                                 d_range);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

// ---------------------------------------------------------------------------------------------
// The candidate grid of a freshly uploaded frame, built on the device (one CTA: a frame has a few thousand
// keypoints and a few thousand cells).  cell_start[c] .. cell_start[c + 1] are the slots of cell c, a slot holds
// (x bits, y bits, keypoint index, keypoint index).  The slots of a cell are in NO particular order: the grid
// without the tree serves only speculative calls, whose result does not depend on the order candidates are
// scored in (a tie at a point's minimum is counted, whichever candidate came first, and the call is repeated with
// the tree: group_search).  On the host the same counting sort cost 85 us per 8000-keypoint frame, the largest
// item of slamgpu_frame_upload; here the host only queues one launch behind the upload.
// Shared memory: one counter per cell (count, then exclusive prefix, then write cursor).
// ---------------------------------------------------------------------------------------------
constexpr int kGridBuildThreads = 1024;

__global__ void __launch_bounds__(kGridBuildThreads) grid_build_kernel(const float2* __restrict__ xy, int n, float gx0,
                                                                       float gy0, float inv_cell, int gw, int gh,
                                                                       int32_t* cell_start, int4* cell_rec)
{
    extern __shared__ int s_cell[];                    // [n_cells]
    __shared__ int s_warp[32];
    const int n_cells = gw * gh;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int c = tid; c < n_cells; c += kGridBuildThreads) s_cell[c] = 0;
    __syncthreads();
    for (int k = tid; k < n; k += kGridBuildThreads) {
        const float2 p = xy[k];
        atomicAdd(&s_cell[grid_cell(p.y, gy0, inv_cell, gh) * gw + grid_cell(p.x, gx0, inv_cell, gw)], 1);
    }
    __syncthreads();
    // exclusive scan: every thread owns a run of consecutive cells
    const int per = (n_cells + kGridBuildThreads - 1) / kGridBuildThreads;
    const int c0 = min(tid * per, n_cells), c1 = min(c0 + per, n_cells);
    int sum = 0;
    for (int c = c0; c < c1; ++c) sum += s_cell[c];
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xFFFFFFFFu, w, o);
            if (lane >= o) w += y;
        }
        s_warp[lane] = w;
    }
    __syncthreads();
    int run = (warp ? s_warp[warp - 1] : 0) + incl - sum;
    for (int c = c0; c < c1; ++c) {
        const int cnt = s_cell[c];
        s_cell[c] = run;                               // from here on: the cell's write cursor
        cell_start[c] = run;
        run += cnt;
    }
    if (tid == 0) cell_start[n_cells] = n;
    __syncthreads();
    for (int k = tid; k < n; k += kGridBuildThreads) {
        const float2 p = xy[k];
        const int slot = atomicAdd(&s_cell[grid_cell(p.y, gy0, inv_cell, gh) * gw + grid_cell(p.x, gx0, inv_cell, gw)], 1);
        cell_rec[slot] = make_int4(__float_as_int(p.x), __float_as_int(p.y), k, k);
    }
}

// true in *done when the frame is of a size the single-CTA build takes (else the caller builds the grid on the host)
int proj_launch_grid_build(slamgpu_ctx* ctx, const float2* d_xy, int n, float gx0, float gy0, float inv_cell, int gw, int gh,
                           int32_t* d_cell_start, int4* d_cell_rec, bool* done)
{
    constexpr size_t kMaxSmem = size_t(200) << 10;
    const size_t smem = static_cast<size_t>(gw) * gh * sizeof(int);
    *done = false;
    if (n <= 0 || n > (1 << 16) || smem > kMaxSmem) return SLAMGPU_OK;
    if (smem > (size_t(48) << 10) && !ctx->grid_attr_set) {
        SG_CUDA(cudaFuncSetAttribute(grid_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kMaxSmem)));
        ctx->grid_attr_set = true;
    }
    grid_build_kernel<<<1, kGridBuildThreads, smem, ctx->stream>>>(d_xy, n, gx0, gy0, inv_cell, gw, gh, d_cell_start, d_cell_rec);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    *done = true;
    return SLAMGPU_OK;
}

int proj_launch_kd_gather(slamgpu_ctx* ctx, const float2* d_xy, const int32_t* d_perm, int n, float2* d_kd_xy)
{
    if (n <= 0) return SLAMGPU_OK;
    kd_gather_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(d_xy, d_perm, n, d_kd_xy);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

} // namespace slamgpu
