// map_api.cu -- the MUTABLE resident map (include/slamgpu.h "Mutable map"; SURVEY 8f N1 / N2 "with incremental
// updates").  The reference changes existing map points all the time, between the calls of the matcher:
//   Map::associate / disassociate (src/Map.cpp:97-124)            a point gains / loses an observation
//   Map::remove_point (src/Map.cpp:63-76), Mapper::cull_points    a point is erased from the MIDDLE of Map's vector
//     (src/Mapper.cpp:396-430)                                    (later points shift down: map order = visit rank)
//   Map::fuse (src/Map.cpp:78-95), Mapper::fuse_match (:176-196)  observations move onto the kept point, the other dies
//   MapPoint::set_position (src/MapPoint.cpp:19-22)               bundle adjustment (src/Optimization.cpp:371,534),
//                                                                 re-anchoring (src/Mapper.cpp:391)
//   Frame::set_pose of an observer                                every dependent avg_viewing_normal /
//                                                                 observed_distance_range changes (src/MapPoint.cpp:24-45)
// A mutable set keeps, beside what match_via_reproject reads, the OBSERVER of every observation (a small integer
// the host assigns to each key frame) and a table of observer camera centres, so that normals and ranges are
// re-derived on the device for exactly the points an update touches -- in the host's observation order, which
// the caller restates whenever a point's observation set changes (unordered_map iteration order, src/MapPoint.h:38).
//
// Slots keep Map's vector order: new points are appended at the end (Map::create_point pushes back,
// src/Map.cpp:20-61), a removed point becomes a tombstone (no observations, NaN position: it fails is_in_image and
// proposes nothing) and slamgpu_points_compact squeezes the tombstones out without reordering, so "lowest slot"
// stays "first in map order", the proposal tie-break of src/MapMatcher.cpp:95-97.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>

#include "slamgpu_internal.cuh"

using namespace slamgpu;

namespace {

#define SG_CTX(ctx)                                 \
    do {                                            \
        SG_CHECK((ctx) != nullptr, "null context"); \
        SG_CUDA(cudaSetDevice((ctx)->device));      \
    } while (0)

// ---- kernels ----------------------------------------------------------------------------------

__device__ __forceinline__ float sum3(float a0, float a1, float a2) { return __fadd_rn(a0, __fadd_rn(a1, a2)); }
// Eigen normalized(): z = squaredNorm(); z > 0 ? v / sqrt(z) : v   (same as proj_kernels.cu)
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

struct MapDev {
    float* pos;
    float* normal;
    float* range;
    int32_t* obs_begin;
    int32_t* obs_end;
    uint4* obs_desc;
    int32_t* obs_kf;
    float* kf_center;
    uint8_t* kf_dirty;
};

// MapPoint::avg_viewing_normal / observed_distance_range (src/MapPoint.cpp:24-45) for the slots idx[0..n) (idx null:
// slots 0..n), observers' centres from the table.  only_dirty: skip a point none of whose observers has moved.
__global__ void map_stats_kernel(const MapDev m, const int32_t* idx, int n, int only_dirty)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int i = idx ? idx[t] : t;
    const int o0 = m.obs_begin[i], o1 = m.obs_end[i];
    if (only_dirty) {
        bool dirty = false;
        for (int o = o0; o < o1; ++o) dirty = dirty || m.kf_dirty[m.obs_kf[o]] != 0;
        if (!dirty) return;
    }
    const float X[3] = {m.pos[3 * i], m.pos[3 * i + 1], m.pos[3 * i + 2]};
    float acc[3] = {0.0f, 0.0f, 0.0f};
    float nearest = 3.402823466e+38f, furthest = 0.0f;
    for (int o = o0; o < o1; ++o) {
        const float* C = m.kf_center + 3 * static_cast<size_t>(m.obs_kf[o]);
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
    m.normal[3 * i] = nrm[0]; m.normal[3 * i + 1] = nrm[1]; m.normal[3 * i + 2] = nrm[2];
    m.range[2 * i] = nearest;
    m.range[2 * i + 1] = furthest;
}

// per listed slot: new position and / or new observation run
__global__ void map_scatter_points_kernel(const MapDev m, const int32_t* idx, int n, const float* pos,
                                          const int32_t* begin, const int32_t* end)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int i = idx[t];
    if (pos) {
        m.pos[3 * i] = pos[3 * t];
        m.pos[3 * i + 1] = pos[3 * t + 1];
        m.pos[3 * i + 2] = pos[3 * t + 2];
    }
    if (begin) {
        m.obs_begin[i] = begin[t];
        m.obs_end[i] = end[t];
    }
}

// staged observations -> their arena slots (two threads per observation: one 16-byte half each)
__global__ void map_scatter_obs_kernel(const MapDev m, const uint4* src_desc, const int32_t* src_kf,
                                       const int32_t* dst, int n_obs)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2 * n_obs) return;
    const int o = t >> 1, h = t & 1;
    const size_t d = static_cast<size_t>(dst[o]);
    m.obs_desc[2 * d + h] = src_desc[2 * static_cast<size_t>(o) + h];
    if (h == 0) m.obs_kf[d] = src_kf[o];
}

__global__ void map_set_centers_kernel(const MapDev m, const int32_t* ids, int n, const float* centers)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int k = ids[t];
    m.kf_center[3 * k] = centers[3 * t];
    m.kf_center[3 * k + 1] = centers[3 * t + 1];
    m.kf_center[3 * k + 2] = centers[3 * t + 2];
    m.kf_dirty[k] = 1;
}

// repack: new slot j takes old slot src_slot[j]; its run moves to new_begin[j]
__global__ void map_gather_points_kernel(const MapDev m, const int32_t* src_slot, const int32_t* new_begin, int n_new,
                                         float* t_pos, float* t_normal, float* t_range, int32_t* t_begin, int32_t* t_end)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_new) return;
    const int s = src_slot[j];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        t_pos[3 * j + k] = m.pos[3 * s + k];
        t_normal[3 * j + k] = m.normal[3 * s + k];
    }
    t_range[2 * j] = m.range[2 * s];
    t_range[2 * j + 1] = m.range[2 * s + 1];
    t_begin[j] = new_begin[j];
    t_end[j] = new_begin[j] + (m.obs_end[s] - m.obs_begin[s]);
}
// one group of 8 threads per new slot copies its run (runs are a handful of observations)
__global__ void map_gather_obs_kernel(const MapDev m, const int32_t* src_slot, const int32_t* new_begin, int n_new,
                                      uint4* t_desc, int32_t* t_kf)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = t >> 3, lane = t & 7;
    if (j >= n_new) return;
    const int s = src_slot[j];
    const int ob = m.obs_begin[s], cnt = m.obs_end[s] - ob, nb = new_begin[j];
    for (int k = lane; k < 2 * cnt; k += 8) t_desc[2 * static_cast<size_t>(nb) + k] = m.obs_desc[2 * static_cast<size_t>(ob) + k];
    for (int k = lane; k < cnt; k += 8) t_kf[nb + k] = m.obs_kf[ob + k];
}

// ---- host side ----------------------------------------------------------------------------------

MapDev dev_of(const PointSetDev& p)
{
    MapDev m;
    m.pos = p.pos;
    m.normal = p.normal;
    m.range = p.range;
    m.obs_begin = p.obs_offset;
    m.obs_end = p.obs_end_w;
    m.obs_desc = p.obs_desc;
    m.obs_kf = p.obs_kf;
    m.kf_center = p.kf_center;
    m.kf_dirty = p.kf_dirty;
    return m;
}

size_t pts_bytes(size_t N)
{
    return 2 * Carver::padded(N * 12) + Carver::padded(N * 8) + 2 * Carver::padded(N * 4) + Carver::padded(N) + 256;
}
void carve_pts(PointSetDev& p, void* blob, size_t N)
{
    Carver dv(blob);
    p.pos = dv.take<float>(N * 3);
    p.normal = dv.take<float>(N * 3);
    p.range = dv.take<float>(N * 2);
    p.obs_offset = dv.take<int32_t>(N);
    p.obs_end_w = dv.take<int32_t>(N);
    p.obs_end = p.obs_end_w;
    p.pt_matched = dv.take<uint8_t>(N);
}
size_t obs_bytes(size_t O) { return Carver::padded(O * 32) + Carver::padded(O * 4) + 256; }
void carve_obs(PointSetDev& p, void* blob, size_t O)
{
    Carver dv(blob);
    p.obs_desc = dv.take<uint4>(O * 2);
    p.obs_kf = dv.take<int32_t>(O);
}
size_t kf_bytes(size_t K) { return Carver::padded(K * 12) + Carver::padded(K) + 256; }
void carve_kf(PointSetDev& p, void* blob, size_t K)
{
    Carver dv(blob);
    p.kf_center = dv.take<float>(K * 3);
    p.kf_dirty = dv.take<uint8_t>(K);
}

int dev_alloc(void** out, size_t bytes, const char* what)
{
    const cudaError_t e = cudaMalloc(out, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        *out = nullptr;
        return fail(SLAMGPU_ERR_NOMEM, std::string(what) + ": " + cudaGetErrorString(e));
    }
    return SLAMGPU_OK;
}

// Grow the slot arrays / the observation arena / the observer table geometrically (device-to-device move of the
// live prefix).  A map is created with a reserve, so a tracking run does not come here.
int grow_pts(slamgpu_ctx* ctx, PointSetDev& p, int need)
{
    if (need <= p.cap_n) return SLAMGPU_OK;
    const int cap = static_cast<int>(std::min<int64_t>(std::max<int64_t>(need, 2 * static_cast<int64_t>(p.cap_n)), 1 << 30));
    void* blob = nullptr;
    SG_TRY(dev_alloc(&blob, pts_bytes(cap), "map slot growth failed"));
    PointSetDev old = p;
    carve_pts(p, blob, cap);
    const size_t n = static_cast<size_t>(old.n);
    cudaError_t e = cudaSuccess;
    if (n) {
        e = cudaMemcpyAsync(p.pos, old.pos, n * 12, cudaMemcpyDeviceToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(p.normal, old.normal, n * 12, cudaMemcpyDeviceToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(p.range, old.range, n * 8, cudaMemcpyDeviceToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(p.obs_offset, old.obs_offset, n * 4, cudaMemcpyDeviceToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(p.obs_end_w, old.obs_end_w, n * 4, cudaMemcpyDeviceToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    }
    if (e != cudaSuccess) {
        cudaFree(blob);
        const bool m = p.is_mutable;
        p = old;
        p.is_mutable = m;
        return fail(SLAMGPU_ERR_CUDA, std::string("map slot growth: ") + cudaGetErrorString(e));
    }
    if (old.pts_blob) cudaFree(old.pts_blob);
    p.pts_blob = blob;
    p.cap_n = cap;
    return SLAMGPU_OK;
}
int grow_obs(slamgpu_ctx* ctx, PointSetDev& p, int64_t need)
{
    if (need <= p.cap_obs) return SLAMGPU_OK;
    SG_CHECK(need <= 0x7FFFFFFF, "observation arena too large");
    const int64_t cap = std::min<int64_t>(std::max<int64_t>(need, 2 * p.cap_obs), 0x7FFFFFFF);
    void* blob = nullptr;
    SG_TRY(dev_alloc(&blob, obs_bytes(static_cast<size_t>(cap)), "map observation arena growth failed"));
    uint4* od = p.obs_desc;
    int32_t* ok = p.obs_kf;
    carve_obs(p, blob, static_cast<size_t>(cap));
    const size_t top = static_cast<size_t>(p.arena_top);
    cudaError_t e = cudaSuccess;
    if (top) {
        e = cudaMemcpyAsync(p.obs_desc, od, top * 32, cudaMemcpyDeviceToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(p.obs_kf, ok, top * 4, cudaMemcpyDeviceToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    }
    if (e != cudaSuccess) {
        cudaFree(blob);
        p.obs_desc = od;
        p.obs_kf = ok;
        return fail(SLAMGPU_ERR_CUDA, std::string("map observation arena growth: ") + cudaGetErrorString(e));
    }
    if (p.obs_blob) cudaFree(p.obs_blob);
    p.obs_blob = blob;
    p.cap_obs = cap;
    return SLAMGPU_OK;
}
int grow_kf(slamgpu_ctx* ctx, PointSetDev& p, int need)
{
    if (need <= p.cap_kf) return SLAMGPU_OK;
    const int cap = std::max(need, 2 * p.cap_kf);
    void* blob = nullptr;
    SG_TRY(dev_alloc(&blob, kf_bytes(cap), "map observer table growth failed"));
    float* oc = p.kf_center;
    carve_kf(p, blob, cap);
    cudaError_t e = cudaMemsetAsync(p.kf_dirty, 0, cap, ctx->stream);
    if (e == cudaSuccess && p.n_kf) e = cudaMemcpyAsync(p.kf_center, oc, static_cast<size_t>(p.n_kf) * 12, cudaMemcpyDeviceToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) {
        cudaFree(blob);
        return fail(SLAMGPU_ERR_CUDA, std::string("map observer table growth: ") + cudaGetErrorString(e));
    }
    if (p.kf_blob) cudaFree(p.kf_blob);
    p.kf_blob = blob;
    p.cap_kf = cap;
    return SLAMGPU_OK;
}

int find_map(slamgpu_ctx* ctx, int set_id, PointSetDev** out)
{
    auto it = ctx->point_sets.find(set_id);
    if (it == ctx->point_sets.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown point set id");
    SG_CHECK(it->second.is_mutable, "not a mutable point set (make it with slamgpu_points_create)");
    *out = &it->second;
    return SLAMGPU_OK;
}

// obs_offset [n+1] relative to the block, observers known to the table
int check_runs(const PointSetDev& p, int n, const int32_t* obs_offset, const uint8_t* obs_desc, const int32_t* obs_observer)
{
    SG_CHECK(obs_offset != nullptr, "null observation offsets");
    SG_CHECK(obs_offset[0] == 0, "obs_offset[0] must be 0 (offsets are relative to the block)");
    for (int i = 0; i < n; ++i) SG_CHECK(obs_offset[i + 1] >= obs_offset[i], "obs_offset must be non-decreasing");
    const int q = obs_offset[n];
    SG_CHECK(q == 0 || (obs_desc != nullptr && obs_observer != nullptr), "null observation descriptors / observers");
    for (int o = 0; o < q; ++o)
        SG_CHECK(obs_observer[o] >= 0 && obs_observer[o] < p.n_kf,
                 "observation names an observer without a centre (call slamgpu_observer_centers_update first)");
    return SLAMGPU_OK;
}

// distinct live slots
int check_slots(PointSetDev& p, const int32_t* idx, int n, std::vector<uint8_t>& mark)
{
    SG_CHECK(idx != nullptr, "null slot list");
    mark.assign(static_cast<size_t>(p.n), 0);
    for (int i = 0; i < n; ++i) {
        SG_CHECK(idx[i] >= 0 && idx[i] < p.n, "slot outside the point set");
        SG_CHECK(p.h_alive[idx[i]], "slot names a removed point");
        SG_CHECK(!mark[idx[i]], "a slot may appear only once per call");
        mark[idx[i]] = 1;
    }
    return SLAMGPU_OK;
}

void launch_stats(slamgpu_ctx* ctx, const PointSetDev& p, const int32_t* d_idx, int n, bool only_dirty)
{
    if (n <= 0) return;
    map_stats_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(dev_of(p), d_idx, n, only_dirty ? 1 : 0);
    ctx->launches++;
}

// Squeeze the garbage out of the observation arena (and, drop_dead, the tombstones out of the slots), keeping
// the order of both.  remap (may be null): [slots before] -> new slot or -1.
int repack(slamgpu_ctx* ctx, PointSetDev& p, bool drop_dead, int32_t* remap)
{
    const int n_old = p.n;
    std::vector<int32_t> src_slot, new_begin;
    src_slot.reserve(static_cast<size_t>(n_old));
    new_begin.reserve(static_cast<size_t>(n_old));
    int64_t top = 0;
    for (int s = 0; s < n_old; ++s) {
        const bool keep = p.h_alive[s] || !drop_dead;
        if (remap) remap[s] = keep ? static_cast<int32_t>(src_slot.size()) : -1;
        if (!keep) continue;
        src_slot.push_back(s);
        new_begin.push_back(static_cast<int32_t>(top));
        top += p.h_count[s];
    }
    const int n_new = static_cast<int>(src_slot.size());
    const size_t N = static_cast<size_t>(n_new), O = static_cast<size_t>(top);
    if (n_new > 0) {
        const size_t b_list = 2 * Carver::padded(N * 4);
        const size_t b_tmp = 2 * Carver::padded(N * 12) + Carver::padded(N * 8) + 2 * Carver::padded(N * 4) +
                             Carver::padded(O * 32) + Carver::padded(O * 4);
        SG_TRY(ensure_host_stage(ctx, b_list + 256));
        SG_TRY(ensure_device_ws(ctx, b_list + b_tmp + 256));
        SG_CUDA(cudaStreamSynchronize(ctx->stream));   // staging is reused
        Carver hv(ctx->h_stage.ptr), dv(ctx->d_ws.ptr);
        int32_t* h_src = hv.take<int32_t>(N);
        int32_t* h_nb = hv.take<int32_t>(N);
        std::memcpy(h_src, src_slot.data(), N * 4);
        std::memcpy(h_nb, new_begin.data(), N * 4);
        int32_t* d_src = dv.take<int32_t>(N);
        int32_t* d_nb = dv.take<int32_t>(N);
        float* t_pos = dv.take<float>(N * 3);
        float* t_normal = dv.take<float>(N * 3);
        float* t_range = dv.take<float>(N * 2);
        int32_t* t_begin = dv.take<int32_t>(N);
        int32_t* t_end = dv.take<int32_t>(N);
        uint4* t_desc = dv.take<uint4>(O * 2);
        int32_t* t_kf = dv.take<int32_t>(O);
        SG_CUDA(cudaMemcpyAsync(d_src, h_src, b_list, cudaMemcpyHostToDevice, ctx->stream));
        const MapDev m = dev_of(p);
        map_gather_points_kernel<<<(n_new + 127) / 128, 128, 0, ctx->stream>>>(m, d_src, d_nb, n_new, t_pos, t_normal,
                                                                             t_range, t_begin, t_end);
        map_gather_obs_kernel<<<(n_new * 8 + 127) / 128, 128, 0, ctx->stream>>>(m, d_src, d_nb, n_new, t_desc, t_kf);
        ctx->launches += 2;
        SG_CUDA(cudaGetLastError());
        SG_CUDA(cudaMemcpyAsync(p.pos, t_pos, N * 12, cudaMemcpyDeviceToDevice, ctx->stream));
        SG_CUDA(cudaMemcpyAsync(p.normal, t_normal, N * 12, cudaMemcpyDeviceToDevice, ctx->stream));
        SG_CUDA(cudaMemcpyAsync(p.range, t_range, N * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        SG_CUDA(cudaMemcpyAsync(p.obs_offset, t_begin, N * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        SG_CUDA(cudaMemcpyAsync(p.obs_end_w, t_end, N * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        if (O) {
            SG_CUDA(cudaMemcpyAsync(p.obs_desc, t_desc, O * 32, cudaMemcpyDeviceToDevice, ctx->stream));
            SG_CUDA(cudaMemcpyAsync(p.obs_kf, t_kf, O * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        }
        SG_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    std::vector<int32_t> hb(N), hc(N);
    std::vector<uint8_t> ha(N);
    for (int j = 0; j < n_new; ++j) {
        hb[j] = new_begin[j];
        hc[j] = p.h_count[src_slot[j]];
        ha[j] = p.h_alive[src_slot[j]];
    }
    p.h_begin.swap(hb);
    p.h_count.swap(hc);
    p.h_alive.swap(ha);
    p.n = n_new;
    p.arena_top = top;
    return SLAMGPU_OK;
}

// room for q more observations at the arena's top: squeeze the garbage out when that is enough, else grow
int reserve_arena(slamgpu_ctx* ctx, PointSetDev& p, int64_t q)
{
    if (p.arena_top + q <= p.cap_obs) return SLAMGPU_OK;
    const int64_t garbage = p.arena_top - p.n_obs;
    if (garbage > 0 && p.n_obs + q <= p.cap_obs - p.cap_obs / 4) SG_TRY(repack(ctx, p, false, nullptr));
    return grow_obs(ctx, p, p.arena_top + q);
}

} // namespace

extern "C" {

int slamgpu_points_create(slamgpu_ctx* ctx, int set_id, int reserve_points, int64_t reserve_observations,
                          int reserve_observers)
{
    SG_CTX(ctx);
    SG_CHECK(reserve_points >= 0 && reserve_observations >= 0 && reserve_observers >= 0, "negative reserve");
    SG_CHECK(reserve_points <= (1 << 30) && reserve_observations <= 0x7FFFFFFF, "reserve too large");
    auto it = ctx->point_sets.find(set_id);
    if (it != ctx->point_sets.end()) {
        SG_CUDA(cudaStreamSynchronize(ctx->stream));
        free_points(it->second);
        ctx->point_sets.erase(it);
    }
    PointSetDev p;
    p.is_mutable = true;
    const int cap_n = std::max(reserve_points, 1024);
    const int64_t cap_obs = std::max<int64_t>(reserve_observations, 4096);
    const int cap_kf = std::max(reserve_observers, 256);
    int rc = dev_alloc(&p.pts_blob, pts_bytes(cap_n), "map allocation failed");
    if (rc == SLAMGPU_OK) rc = dev_alloc(&p.obs_blob, obs_bytes(static_cast<size_t>(cap_obs)), "map allocation failed");
    if (rc == SLAMGPU_OK) rc = dev_alloc(&p.kf_blob, kf_bytes(cap_kf), "map allocation failed");
    if (rc == SLAMGPU_OK && cudaMemsetAsync(p.kf_blob, 0, kf_bytes(cap_kf), ctx->stream) != cudaSuccess)
        rc = fail(SLAMGPU_ERR_CUDA, "map allocation: memset failed");
    if (rc != SLAMGPU_OK) {
        free_points(p);
        return rc;
    }
    carve_pts(p, p.pts_blob, cap_n);
    carve_obs(p, p.obs_blob, static_cast<size_t>(cap_obs));
    carve_kf(p, p.kf_blob, cap_kf);
    p.cap_n = cap_n;
    p.cap_obs = cap_obs;
    p.cap_kf = cap_kf;
    ctx->point_sets[set_id] = p;
    return SLAMGPU_OK;
}

int slamgpu_observer_centers_update(slamgpu_ctx* ctx, int set_id, const int32_t* observer_ids, int n, const float* centers)
{
    SG_CTX(ctx);
    PointSetDev* pp = nullptr;
    SG_TRY(find_map(ctx, set_id, &pp));
    PointSetDev& p = *pp;
    SG_CHECK(n >= 0, "negative observer count");
    if (n == 0) return SLAMGPU_OK;
    SG_CHECK(observer_ids != nullptr && centers != nullptr, "null observer ids / centres");
    int max_id = -1;
    bool moved = false;
    for (int i = 0; i < n; ++i) {
        SG_CHECK(observer_ids[i] >= 0 && observer_ids[i] < (1 << 24), "observer id out of range");
        max_id = std::max(max_id, observer_ids[i]);
        moved = moved || observer_ids[i] < p.n_kf;
    }
    {
        std::vector<int32_t> sorted(observer_ids, observer_ids + n);
        std::sort(sorted.begin(), sorted.end());
        SG_CHECK(std::adjacent_find(sorted.begin(), sorted.end()) == sorted.end(), "an observer may appear only once per call");
    }
    SG_TRY(grow_kf(ctx, p, max_id + 1));
    const size_t b_ids = Carver::padded(sizeof(int32_t) * n), b_c = Carver::padded(sizeof(float) * 3 * n);
    SG_TRY(ensure_host_stage(ctx, b_ids + b_c + 256));
    SG_TRY(ensure_device_io(ctx, b_ids + b_c + 256));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));   // staging is reused
    Carver hv(ctx->h_stage.ptr), dv(ctx->d_io.ptr);
    int32_t* h_ids = hv.take<int32_t>(n);
    float* h_c = hv.take<float>(3 * static_cast<size_t>(n));
    std::memcpy(h_ids, observer_ids, sizeof(int32_t) * n);
    std::memcpy(h_c, centers, sizeof(float) * 3 * n);
    int32_t* d_ids = dv.take<int32_t>(n);
    float* d_c = dv.take<float>(3 * static_cast<size_t>(n));
    SG_CUDA(cudaMemcpyAsync(d_ids, h_ids, b_ids + b_c, cudaMemcpyHostToDevice, ctx->stream));
    map_set_centers_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(dev_of(p), d_ids, n, d_c);
    ctx->launches++;
    p.n_kf = std::max(p.n_kf, max_id + 1);
    // every point that observes a moved key frame gets its normal and range re-derived (src/MapPoint.cpp:24-45);
    // a new observer has no observations yet
    if (moved) launch_stats(ctx, p, nullptr, p.n, true);
    SG_CUDA(cudaMemsetAsync(p.kf_dirty, 0, static_cast<size_t>(p.n_kf), ctx->stream));
    SG_CUDA(cudaGetLastError());
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    return SLAMGPU_OK;
}

// shared by append (idx == null: new slots at the end) and set_observations (idx: existing slots)
static int put_runs(slamgpu_ctx* ctx, PointSetDev& p, const int32_t* idx, int n, const float* pos,
                    const int32_t* obs_offset, const uint8_t* obs_desc, const int32_t* obs_observer)
{
    const int64_t q = obs_offset[n];
    // in place when the new run is no longer than the old one, else at the arena's top
    int64_t at_top = 0;
    for (int i = 0; i < n; ++i) {
        const int cnt = obs_offset[i + 1] - obs_offset[i];
        if (idx == nullptr || cnt > p.h_count[idx[i]]) at_top += cnt;
    }
    SG_TRY(reserve_arena(ctx, p, at_top));
    const size_t N = static_cast<size_t>(n), Q = static_cast<size_t>(q);
    const size_t b_idx = Carver::padded(N * 4), b_pos = pos ? Carver::padded(N * 12) : 0, b_be = 2 * Carver::padded(N * 4),
                 b_desc = Carver::padded(Q * 32), b_kf = Carver::padded(Q * 4), b_dst = Carver::padded(Q * 4);
    const size_t total = b_idx + b_pos + b_be + b_desc + b_kf + b_dst;
    SG_TRY(ensure_host_stage(ctx, total + 256));
    SG_TRY(ensure_device_io(ctx, total + 256));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));   // staging is reused
    Carver hv(ctx->h_stage.ptr), dv(ctx->d_io.ptr);
    int32_t* h_idx = hv.take<int32_t>(N);
    float* h_pos = pos ? hv.take<float>(N * 3) : nullptr;
    int32_t* h_b = hv.take<int32_t>(N);
    int32_t* h_e = hv.take<int32_t>(N);
    uint8_t* h_desc = reinterpret_cast<uint8_t*>(hv.take<uint4>(Q * 2));
    int32_t* h_kf = hv.take<int32_t>(Q);
    int32_t* h_dst = hv.take<int32_t>(Q);
    int32_t* d_idx = dv.take<int32_t>(N);
    float* d_pos = pos ? dv.take<float>(N * 3) : nullptr;
    int32_t* d_b = dv.take<int32_t>(N);
    int32_t* d_e = dv.take<int32_t>(N);
    uint4* d_desc = dv.take<uint4>(Q * 2);
    int32_t* d_kf = dv.take<int32_t>(Q);
    int32_t* d_dst = dv.take<int32_t>(Q);

    const int slot0 = p.n;
    for (int i = 0; i < n; ++i) {
        const int s = idx ? idx[i] : slot0 + i;
        const int cnt = obs_offset[i + 1] - obs_offset[i];
        int32_t b;
        if (idx != nullptr && cnt <= p.h_count[s]) {
            b = p.h_begin[s];
            p.n_obs -= p.h_count[s] - cnt;
        } else {
            b = static_cast<int32_t>(p.arena_top);
            p.arena_top += cnt;
            p.n_obs += cnt - (idx ? p.h_count[s] : 0);
        }
        if (idx == nullptr) {
            p.h_begin.push_back(b);
            p.h_count.push_back(cnt);
            p.h_alive.push_back(1);
        } else {
            p.h_begin[s] = b;
            p.h_count[s] = cnt;
        }
        h_idx[i] = s;
        h_b[i] = b;
        h_e[i] = b + cnt;
        for (int k = 0; k < cnt; ++k) h_dst[obs_offset[i] + k] = b + k;
    }
    if (pos) std::memcpy(h_pos, pos, N * 12);
    if (Q) {
        std::memcpy(h_desc, obs_desc, Q * 32);
        std::memcpy(h_kf, obs_observer, Q * 4);
    }
    if (idx == nullptr) {
        p.n += n;
        p.n_live += n;
    }
    SG_CUDA(cudaMemcpyAsync(ctx->d_io.ptr, ctx->h_stage.ptr, total, cudaMemcpyHostToDevice, ctx->stream));
    const MapDev m = dev_of(p);
    map_scatter_points_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(m, d_idx, n, d_pos, d_b, d_e);
    ctx->launches++;
    if (q > 0) {
        map_scatter_obs_kernel<<<static_cast<unsigned>((2 * q + 127) / 128), 128, 0, ctx->stream>>>(m, d_desc, d_kf, d_dst,
                                                                                                  static_cast<int>(q));
        ctx->launches++;
    }
    launch_stats(ctx, p, d_idx, n, false);
    SG_CUDA(cudaGetLastError());
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    return SLAMGPU_OK;
}

int slamgpu_points_append_observed(slamgpu_ctx* ctx, int set_id, int n_new, const float* pos, const int32_t* obs_offset,
                                   const uint8_t* obs_desc, const int32_t* obs_observer)
{
    SG_CTX(ctx);
    PointSetDev* pp = nullptr;
    SG_TRY(find_map(ctx, set_id, &pp));
    PointSetDev& p = *pp;
    SG_CHECK(n_new >= 0, "negative point count");
    if (n_new == 0) return SLAMGPU_OK;
    SG_CHECK(pos != nullptr, "null positions");
    SG_TRY(check_runs(p, n_new, obs_offset, obs_desc, obs_observer));
    SG_CHECK(static_cast<int64_t>(p.n) + n_new <= (1 << 30), "point set too large");
    SG_TRY(grow_pts(ctx, p, p.n + n_new));
    return put_runs(ctx, p, nullptr, n_new, pos, obs_offset, obs_desc, obs_observer);
}

int slamgpu_points_set_observations(slamgpu_ctx* ctx, int set_id, const int32_t* idx, int n, const int32_t* obs_offset,
                                    const uint8_t* obs_desc, const int32_t* obs_observer)
{
    SG_CTX(ctx);
    PointSetDev* pp = nullptr;
    SG_TRY(find_map(ctx, set_id, &pp));
    PointSetDev& p = *pp;
    SG_CHECK(n >= 0, "negative slot count");
    if (n == 0) return SLAMGPU_OK;
    std::vector<uint8_t> mark;
    SG_TRY(check_slots(p, idx, n, mark));
    SG_TRY(check_runs(p, n, obs_offset, obs_desc, obs_observer));
    return put_runs(ctx, p, idx, n, nullptr, obs_offset, obs_desc, obs_observer);
}

int slamgpu_points_set_position(slamgpu_ctx* ctx, int set_id, const int32_t* idx, int n, const float* pos)
{
    SG_CTX(ctx);
    PointSetDev* pp = nullptr;
    SG_TRY(find_map(ctx, set_id, &pp));
    PointSetDev& p = *pp;
    SG_CHECK(n >= 0, "negative slot count");
    if (n == 0) return SLAMGPU_OK;
    SG_CHECK(pos != nullptr, "null positions");
    std::vector<uint8_t> mark;
    SG_TRY(check_slots(p, idx, n, mark));
    const size_t N = static_cast<size_t>(n);
    const size_t b_idx = Carver::padded(N * 4), b_pos = Carver::padded(N * 12);
    SG_TRY(ensure_host_stage(ctx, b_idx + b_pos + 256));
    SG_TRY(ensure_device_io(ctx, b_idx + b_pos + 256));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));   // staging is reused
    Carver hv(ctx->h_stage.ptr), dv(ctx->d_io.ptr);
    int32_t* h_idx = hv.take<int32_t>(N);
    float* h_pos = hv.take<float>(N * 3);
    std::memcpy(h_idx, idx, N * 4);
    std::memcpy(h_pos, pos, N * 12);
    int32_t* d_idx = dv.take<int32_t>(N);
    float* d_pos = dv.take<float>(N * 3);
    SG_CUDA(cudaMemcpyAsync(ctx->d_io.ptr, ctx->h_stage.ptr, b_idx + b_pos, cudaMemcpyHostToDevice, ctx->stream));
    map_scatter_points_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(dev_of(p), d_idx, n, d_pos, nullptr, nullptr);
    ctx->launches++;
    launch_stats(ctx, p, d_idx, n, false);       // the moved points' normals and ranges (src/MapPoint.cpp:24-45)
    SG_CUDA(cudaGetLastError());
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    return SLAMGPU_OK;
}

int slamgpu_points_remove(slamgpu_ctx* ctx, int set_id, const int32_t* idx, int n)
{
    SG_CTX(ctx);
    PointSetDev* pp = nullptr;
    SG_TRY(find_map(ctx, set_id, &pp));
    PointSetDev& p = *pp;
    SG_CHECK(n >= 0, "negative slot count");
    if (n == 0) return SLAMGPU_OK;
    std::vector<uint8_t> mark;
    SG_TRY(check_slots(p, idx, n, mark));
    const size_t N = static_cast<size_t>(n);
    const size_t b_idx = Carver::padded(N * 4), b_pos = Carver::padded(N * 12), b_be = 2 * Carver::padded(N * 4);
    SG_TRY(ensure_host_stage(ctx, b_idx + b_pos + b_be + 256));
    SG_TRY(ensure_device_io(ctx, b_idx + b_pos + b_be + 256));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));   // staging is reused
    Carver hv(ctx->h_stage.ptr), dv(ctx->d_io.ptr);
    int32_t* h_idx = hv.take<int32_t>(N);
    float* h_pos = hv.take<float>(N * 3);
    int32_t* h_b = hv.take<int32_t>(N);
    int32_t* h_e = hv.take<int32_t>(N);
    const float nan = std::numeric_limits<float>::quiet_NaN();
    for (int i = 0; i < n; ++i) {      // a tombstone: no observations, a position that fails is_in_image (src/Camera.cpp:34-37)
        h_idx[i] = idx[i];
        h_pos[3 * i] = h_pos[3 * i + 1] = h_pos[3 * i + 2] = nan;
        h_b[i] = 0;
        h_e[i] = 0;
        p.n_obs -= p.h_count[idx[i]];
        p.h_count[idx[i]] = 0;
        p.h_begin[idx[i]] = 0;
        p.h_alive[idx[i]] = 0;
    }
    p.n_live -= n;
    int32_t* d_idx = dv.take<int32_t>(N);
    float* d_pos = dv.take<float>(N * 3);
    int32_t* d_b = dv.take<int32_t>(N);
    int32_t* d_e = dv.take<int32_t>(N);
    SG_CUDA(cudaMemcpyAsync(ctx->d_io.ptr, ctx->h_stage.ptr, b_idx + b_pos + b_be, cudaMemcpyHostToDevice, ctx->stream));
    map_scatter_points_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(dev_of(p), d_idx, n, d_pos, d_b, d_e);
    ctx->launches++;
    launch_stats(ctx, p, d_idx, n, false);
    SG_CUDA(cudaGetLastError());
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    return SLAMGPU_OK;
}

int slamgpu_points_compact(slamgpu_ctx* ctx, int set_id, int32_t* out_remap, int* out_n)
{
    SG_CTX(ctx);
    PointSetDev* pp = nullptr;
    SG_TRY(find_map(ctx, set_id, &pp));
    SG_TRY(repack(ctx, *pp, true, out_remap));
    if (out_n) *out_n = pp->n;
    return SLAMGPU_OK;
}

int slamgpu_points_live(slamgpu_ctx* ctx, int set_id, int* live_points, int64_t* arena_used, int64_t* arena_garbage)
{
    SG_CHECK(ctx != nullptr, "null context");
    PointSetDev* pp = nullptr;
    SG_TRY(find_map(ctx, set_id, &pp));
    if (live_points) *live_points = pp->n_live;
    if (arena_used) *arena_used = pp->arena_top;
    if (arena_garbage) *arena_garbage = pp->arena_top - pp->n_obs;
    return SLAMGPU_OK;
}

} // extern "C"
