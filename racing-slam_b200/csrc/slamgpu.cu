// slamgpu.cu -- the C ABI of libslamgpu (include/slamgpu.h): contexts, resident frame / map-point
// state, pinned staging and the stream-ordered match calls.  The kernels live in bf_kernels.cu
// and proj_kernels.cu.  There is no CPU path in this library: every entry point that computes
// launches CUDA kernels on the context's stream, and slamgpu_create() fails without a device.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <thread>

#include <cmath>

#include "slamgpu_internal.cuh"

namespace slamgpu {

static thread_local std::string g_error;

void set_error(const std::string& msg) { g_error = msg; }
int fail(int code, const std::string& msg)
{
    g_error = msg;
    return code;
}

static cudaEvent_t prof_event(slamgpu_ctx* ctx)
{
    if (!ctx->prof_pool.empty()) {
        cudaEvent_t e = ctx->prof_pool.back();
        ctx->prof_pool.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    if (cudaEventCreate(&e) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return e;
}

ProfScope::ProfScope(slamgpu_ctx* c, int k) : ctx(c), kernel(k)
{
    if (!ctx->profiling) return;
    cudaEvent_t start = prof_event(ctx);
    stop = prof_event(ctx);
    if (!start || !stop) {
        stop = nullptr;
        return;
    }
    cudaEventRecord(start, ctx->stream);
    ctx->prof_events[kernel].push_back({start, stop});
}
ProfScope::~ProfScope()
{
    if (stop) cudaEventRecord(stop, ctx->stream);
}

static int grow(slamgpu_ctx* ctx, Arena& a, size_t bytes, bool host)
{
    if (bytes <= a.cap) return SLAMGPU_OK;
    // the old block may still be in use by work queued on the stream
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    if (a.ptr) {
        if (host) cudaFreeHost(a.ptr); else cudaFree(a.ptr);
        a.ptr = nullptr;
        a.cap = 0;
    }
    // small arenas double (a growing map asks for a little more on every frame, and cudaMalloc / cudaMallocHost
    // cost milliseconds); large ones add a quarter
    size_t cap = std::max<size_t>(bytes < (size_t(64) << 20) ? 2 * bytes : bytes + bytes / 4, size_t(1) << 20);
    cap = (cap + 4095) & ~size_t(4095);
    void* p = nullptr;
    const cudaError_t e = host ? cudaMallocHost(&p, cap) : cudaMalloc(&p, cap);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(SLAMGPU_ERR_NOMEM, std::string(host ? "pinned" : "device") + " allocation of " +
                                           std::to_string(cap) + " bytes failed: " + cudaGetErrorString(e));
    }
    a.ptr = p;
    a.cap = cap;
    return SLAMGPU_OK;
}

int wait_stream(slamgpu_ctx* ctx)
{
    SG_CUDA(cudaEventRecord(ctx->sync_event, ctx->stream));
    SG_CUDA(cudaEventSynchronize(ctx->sync_event));
    return SLAMGPU_OK;
}

int ensure_device_ws(slamgpu_ctx* ctx, size_t bytes) { return grow(ctx, ctx->d_ws, bytes, false); }
int ensure_device_io(slamgpu_ctx* ctx, size_t bytes) { return grow(ctx, ctx->d_io, bytes, false); }
int ensure_host_stage(slamgpu_ctx* ctx, size_t bytes) { return grow(ctx, ctx->h_stage, bytes, true); }
int ensure_device_exp(slamgpu_ctx* ctx, size_t bytes) { return grow(ctx, ctx->d_exp, bytes, false); }

// Host float helpers in the oracle's / Eigen's coefficient order (see proj_kernels.cu).  This
// translation unit is compiled with -ffp-contract=off for the host, so a*b+c stays two roundings.
static inline float hsum3(float a0, float a1, float a2)
{
    volatile float t = a1 + a2;
    return a0 + t;
}
static inline float pose_at(const float* pose, int r, int c) { return pose[c * 4 + r]; }

// src/Camera.cpp:27, first product: M = K * pose[0:3, 0:4]
void host_projection_matrix(const float* pose, float fx, float fy, float cx, float cy, float* M)
{
    const float K[3][3] = {{fx, 0.0f, cx}, {0.0f, fy, cy}, {0.0f, 0.0f, 1.0f}};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 4; ++j) {
            volatile float p0 = K[i][0] * pose_at(pose, 0, j);
            volatile float p1 = K[i][1] * pose_at(pose, 1, j);
            volatile float p2 = K[i][2] * pose_at(pose, 2, j);
            M[i * 4 + j] = hsum3(p0, p1, p2);
        }
}

// src/Frame.cpp:39-42: C = -R^T * t
void host_camera_center(const float* pose, float* C)
{
    for (int i = 0; i < 3; ++i) {
        volatile float p0 = -pose_at(pose, 0, i) * pose_at(pose, 0, 3);
        volatile float p1 = -pose_at(pose, 1, i) * pose_at(pose, 1, 3);
        volatile float p2 = -pose_at(pose, 2, i) * pose_at(pose, 2, 3);
        C[i] = hsum3(p0, p1, p2);
    }
}

// KDTree2D::build (src/KDTree.cpp:8-43) as a permutation: the node of range [s, e) is
// perm[(s+e)/2] after std::nth_element on axis depth%2.  Ties between equal coordinates are
// decided by libstdc++'s nth_element, so the same call, comparison results and ranges are used.
// Two things make it faster than the reference's form without changing a single comparison outcome:
//   * the elements are (x, y, index) records instead of indices into a point array -- nth_element
//     only sees comparator results and moves elements, so the arrangement is the same permutation,
//     without the indirection;
//   * after a range is partitioned its two halves are disjoint, so the four quarter-ranges below the
//     first two levels are built on separate threads (the order in which sub-ranges are processed
//     cannot change the result).  This is the Frame constructor's cost (src/Frame.cpp:11-15): 1.3 ms
//     single-threaded for 8000 keypoints, the largest item of a tracked frame
//     (profiles/r01_c3_sequence_2000_frames_v1.json).
// Measured and dropped: a restatement of nth_element with branch-free partition scans (no faster than
// libstdc++'s: ~4 ns per element visit at every level, not branch misses) and a persistent worker pool
// (condition-variable hand-offs cost more than the two thread forks they save).
struct KdItem {
    float x, y;
    uint32_t idx;
    uint32_t pad;
};

static void kd_build_range(KdItem* it, int s, int e, int depth, int fork_min)
{
    struct Range { int s, e, depth; };
    std::vector<Range> todo;
    std::vector<std::thread> forks;
    todo.push_back({s, e, depth});
    while (!todo.empty()) {
        const Range r = todo.back();
        todo.pop_back();
        if (r.s >= r.e) continue;
        const int mid = (r.s + r.e) / 2;
        if (r.depth % 2 == 0)
            std::nth_element(it + r.s, it + mid, it + r.e, [](const KdItem& a, const KdItem& b) { return a.x < b.x; });
        else
            std::nth_element(it + r.s, it + mid, it + r.e, [](const KdItem& a, const KdItem& b) { return a.y < b.y; });
        if (r.e - r.s >= fork_min) {
            forks.emplace_back(kd_build_range, it, mid + 1, r.e, r.depth + 1, fork_min);
        } else {
            todo.push_back({mid + 1, r.e, r.depth + 1});
        }
        todo.push_back({r.s, mid, r.depth + 1});
    }
    for (auto& t : forks) t.join();
}

static void kd_build_perm(const float* xy, int n, uint32_t* out_perm)
{
    std::vector<KdItem> items(static_cast<size_t>(n));
    for (int i = 0; i < n; ++i) items[i] = {xy[2 * i], xy[2 * i + 1], static_cast<uint32_t>(i), 0u};
    // SLAMGPU_KD_THREADS = 1, 2, 4 (default), 8 ...: ranges of at least 2n / threads elements fork
    static const int max_threads = [] {
        const char* e = std::getenv("SLAMGPU_KD_THREADS");
        const int hw = static_cast<int>(std::max(1u, std::thread::hardware_concurrency()));
        return std::max(1, std::min(e ? std::atoi(e) : 4, hw));
    }();
    int fork_min = n + 1;
    if (max_threads > 1 && n >= 2048) fork_min = std::max(512, 2 * n / max_threads - 16);
    kd_build_range(items.data(), 0, n, 0, fork_min);
    for (int i = 0; i < n; ++i) out_perm[i] = items[i].idx;
}


#define SG_CTX(ctx)                                                       \
    do {                                                                  \
        SG_CHECK((ctx) != nullptr, "null context");                       \
        SG_CUDA(cudaSetDevice((ctx)->device));                            \
    } while (0)

void free_points(PointSetDev& p)
{
    if (p.blob) cudaFree(p.blob);
    if (p.pts_blob) cudaFree(p.pts_blob);
    if (p.obs_blob) cudaFree(p.obs_blob);
    if (p.kf_blob) cudaFree(p.kf_blob);
    p = PointSetDev();
}

} // namespace slamgpu

using namespace slamgpu;

// =============================================================================================
// context
// =============================================================================================
extern "C" {

int slamgpu_abi_version(void) { return SLAMGPU_ABI_VERSION; }
const char* slamgpu_last_error(void) { return g_error.c_str(); }

int slamgpu_device_count(int* count)
{
    SG_CHECK(count != nullptr, "null count");
    int n = 0;
    const cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        *count = 0;
        return fail(SLAMGPU_ERR_CUDA, std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e));
    }
    *count = n;
    return SLAMGPU_OK;
}

int slamgpu_create(int device, slamgpu_ctx** out)
{
    SG_CHECK(out != nullptr, "null out pointer");
    *out = nullptr;
    int n = 0;
    SG_TRY(slamgpu_device_count(&n));
    if (n <= 0) return fail(SLAMGPU_ERR_CUDA, "no CUDA device: libslamgpu has no CPU fallback");
    SG_CHECK(device >= 0 && device < n, "device index out of range");
    SG_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    SG_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        return fail(SLAMGPU_ERR_CUDA, std::string("libslamgpu is built for sm_100a only; device is sm_") +
                                          std::to_string(prop.major) + std::to_string(prop.minor));
    }
    slamgpu_ctx* ctx = new slamgpu_ctx();
    ctx->device = device;
    ctx->n_sm = prop.multiProcessorCount;
    const char* walk = std::getenv("SLAMGPU_PROJ_WALK");
    ctx->proj_walk = walk != nullptr && walk[0] == '1';
    const char* bs = std::getenv("SLAMGPU_BATCH_SLICE");
    if (bs != nullptr) ctx->batch_slice = std::max(std::atoi(bs), 0);
    const char* pq = std::getenv("SLAMGPU_PROJ_QUEUE");
    if (pq != nullptr) ctx->proj_queue = std::atoi(pq) != 0;
    const char* tp = std::getenv("SLAMGPU_PROJ_TWOPASS");
    if (tp != nullptr) ctx->proj_two_pass = std::atoi(tp);
    const char* mr = std::getenv("SLAMGPU_MAP_RESERVE");
    if (mr != nullptr) ctx->map_reserve = std::min(std::max(std::atoi(mr), 0), 1 << 28);
    const char* trb = std::getenv("SLAMGPU_TRACE_BATCH");
    ctx->trace_batch = trb != nullptr && trb[0] == '1';
    const char* cos = std::getenv("SLAMGPU_TC4_CROSS_ONE_SCAN");
    if (cos != nullptr) ctx->tc4_cross_one_scan = cos[0] != '0';
    const char* kce = std::getenv("SLAMGPU_KFDB_CACHE");
    if (kce != nullptr) ctx->kfdb_cache_exp = kce[0] != '0';
    const char* tff = std::getenv("SLAMGPU_TC4_FUSE_FINISH");
    if (tff != nullptr) ctx->tc4_fuse_finish = tff[0] != '0';
    const char* bpu = std::getenv("SLAMGPU_BF_PUSH");
    if (bpu != nullptr) ctx->bf_push = bpu[0] != '0';
    const char* bzc = std::getenv("SLAMGPU_BF_ZEROCOPY");
    if (bzc != nullptr) ctx->bf_zero_copy = bzc[0] != '0';
    const char* pzc = std::getenv("SLAMGPU_PROJ_ZEROCOPY");
    if (pzc != nullptr) ctx->proj_zero_copy = pzc[0] != '0';
    const char* god = std::getenv("SLAMGPU_GRID_ON_DEVICE");
    if (god != nullptr) ctx->grid_on_device = god[0] != '0';
    const char* kdl = std::getenv("SLAMGPU_KD_LAZY");
    if (kdl != nullptr) ctx->kd_lazy = kdl[0] != '0';
    const char* kdd = std::getenv("SLAMGPU_KD_DELAY_MS");
    if (kdd != nullptr) ctx->kd_delay_ms = std::min(std::max(std::atoi(kdd), 0), 10000);
    const char* kds = std::getenv("SLAMGPU_KD_SPECULATE");
    if (kds != nullptr) ctx->kd_speculate = kds[0] != '0';
    const char* kda = std::getenv("SLAMGPU_KD_ASYNC");
    if (kda != nullptr) ctx->kd_async = kda[0] != '0';
    const char* v2 = std::getenv("SLAMGPU_PROJ_V2");
    ctx->proj_v2 = v2 != nullptr && v2[0] == '1';
    const char* gcell = std::getenv("SLAMGPU_GRID_CELL");
    if (gcell != nullptr && std::atof(gcell) >= 1.0) ctx->grid_cell = static_cast<float>(std::atof(gcell));
    const char* lanes = std::getenv("SLAMGPU_PROJ_LANES");
    if (lanes != nullptr) {
        const int v = std::atoi(lanes);
        ctx->proj_lanes = (v == 8 || v == 32) ? v : 0;
    }
    const char* engine = std::getenv("SLAMGPU_BF_ENGINE");
    if (engine != nullptr) ctx->bf_engine = std::strcmp(engine, "int") == 0 ? 1 : std::strcmp(engine, "tc") == 0 ? 2 : 0;
    const char* teams = std::getenv("SLAMGPU_TC4_TEAMS");
    if (teams != nullptr) ctx->tc4_teams = teams[0] != '0';
    const char* narrow = std::getenv("SLAMGPU_TC4_NARROW");
    if (narrow != nullptr) ctx->tc4_narrow_last = narrow[0] != '0';
#ifdef SLAMGPU_DEBUG_VARIANTS       // timing experiments that can corrupt results: never in the release library
    const char* variant = std::getenv("SLAMGPU_TC_" "VARIANT");
    if (variant != nullptr) ctx->tc_variant = std::atoi(variant);
#endif
    const cudaError_t e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete ctx;
        return fail(SLAMGPU_ERR_CUDA, std::string("cudaStreamCreate: ") + cudaGetErrorString(e));
    }
    if (cudaEventCreateWithFlags(&ctx->sync_event, cudaEventDisableTiming) != cudaSuccess) {
        const std::string msg = cudaGetErrorString(cudaGetLastError());
        slamgpu_destroy(ctx);
        return fail(SLAMGPU_ERR_CUDA, "context resources: " + msg);
    }
    // set-up for slamgpu_knn2_dev_noalloc: the address of its fixed scratch array on this device, so that the entry
    // point itself never has to ask (a failure here is reported by that entry point, not by slamgpu_create)
    if (bf_noalloc_setup(device) != SLAMGPU_OK) cudaGetLastError();
    *out = ctx;
    return SLAMGPU_OK;
}

static void free_frame(FrameDev& f)
{
    if (f.pending && f.pending->worker.joinable()) f.pending->worker.join();   // a KD build still running
    if (f.blob) cudaFree(f.blob);
    f = FrameDev();
}

int slamgpu_destroy(slamgpu_ctx* ctx)
{
    if (ctx == nullptr) return SLAMGPU_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto& kv : ctx->frames) free_frame(kv.second);
    for (auto& kv : ctx->point_sets) free_points(kv.second);
    free_all_vocabs(ctx);
    free_all_kfdbs(ctx);
    if (ctx->d_ws.ptr) cudaFree(ctx->d_ws.ptr);
    if (ctx->d_io.ptr) cudaFree(ctx->d_io.ptr);
    if (ctx->d_exp.ptr) cudaFree(ctx->d_exp.ptr);
    if (ctx->h_stage.ptr) cudaFreeHost(ctx->h_stage.ptr);
    for (auto& v : ctx->prof_events)
        for (auto& pr : v) {
            cudaEventDestroy(pr.first);
            cudaEventDestroy(pr.second);
        }
    for (cudaEvent_t e : ctx->prof_pool) cudaEventDestroy(e);
    for (cudaEvent_t e : ctx->slice_events) cudaEventDestroy(e);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->sync_event) cudaEventDestroy(ctx->sync_event);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return SLAMGPU_OK;
}

int slamgpu_sync(slamgpu_ctx* ctx)
{
    SG_CTX(ctx);
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    return SLAMGPU_OK;
}

void* slamgpu_stream(slamgpu_ctx* ctx) { return ctx ? static_cast<void*>(ctx->stream) : nullptr; }
int64_t slamgpu_launch_count(slamgpu_ctx* ctx) { return ctx ? ctx->launches : 0; }

int slamgpu_host_alloc(void** ptr, size_t bytes)
{
    SG_CHECK(ptr != nullptr, "null ptr");
    *ptr = nullptr;
    if (bytes == 0) return SLAMGPU_OK;
    const cudaError_t e = cudaMallocHost(ptr, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(SLAMGPU_ERR_NOMEM, std::string("cudaMallocHost: ") + cudaGetErrorString(e));
    }
    return SLAMGPU_OK;
}

int slamgpu_host_free(void* ptr)
{
    if (ptr) SG_CUDA(cudaFreeHost(ptr));
    return SLAMGPU_OK;
}

int slamgpu_kd_build_host(const float* xy, int n, uint32_t* out_perm)
{
    SG_CHECK(n >= 0, "negative size");
    SG_CHECK(n == 0 || (xy != nullptr && out_perm != nullptr), "null pointer");
    kd_build_perm(xy, n, out_perm);
    return SLAMGPU_OK;
}

int slamgpu_profile_enable(slamgpu_ctx* ctx, int on)
{
    SG_CHECK(ctx != nullptr, "null context");
    ctx->profiling = on != 0;
    return SLAMGPU_OK;
}

int slamgpu_profile_read(slamgpu_ctx* ctx, int kernel, double* total_ms, int64_t* launches)
{
    SG_CTX(ctx);
    SG_CHECK(kernel >= SLAMGPU_KERNEL_BF_SCAN && kernel <= SLAMGPU_KERNEL_PNP_SCORE, "unknown kernel id");
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    double sum = 0.0;
    auto& v = ctx->prof_events[kernel];
    for (auto& pr : v) {
        float ms = 0.0f;
        SG_CUDA(cudaEventElapsedTime(&ms, pr.first, pr.second));
        sum += ms;
        ctx->prof_pool.push_back(pr.first);
        ctx->prof_pool.push_back(pr.second);
    }
    if (total_ms) *total_ms = sum;
    if (launches) *launches = static_cast<int64_t>(v.size());
    v.clear();
    return SLAMGPU_OK;
}

// ---- host restatements of the float expressions of the projection path (no device needed) --------------
// The same operation order as the kernels (proj_kernels.cu, map_api.cu) and the oracle, in plain host floats
// (-ffp-contract=off): what MapMatcherGpu::self_check / audit and tools/eigen_order_check.cpp compare with the
// reference's Eigen expressions, bit for bit.
int slamgpu_host_project(const float* pose_colmajor16, const slamgpu_proj_params* params, const float* xyz, float* out_uv)
{
    SG_CHECK(pose_colmajor16 && params && xyz && out_uv, "null pointer");
    float M[12];
    host_projection_matrix(pose_colmajor16, params->fx, params->fy, params->cx, params->cy, M);
    float h[3];
    for (int i = 0; i < 3; ++i) {      // M[:, :3] * X summed a0 + (a1 + a2), then + M[:, 3] (src/Camera.cpp:27)
        volatile float p0 = M[4 * i] * xyz[0], p1 = M[4 * i + 1] * xyz[1], p2 = M[4 * i + 2] * xyz[2];
        volatile float s = hsum3(p0, p1, p2);
        h[i] = s + M[4 * i + 3];
    }
    if (h[2] < 0.0f) {                 // src/Camera.cpp:28-30
        out_uv[0] = -1.0f;
        out_uv[1] = -1.0f;
    } else {
        out_uv[0] = h[0] / h[2];
        out_uv[1] = h[1] / h[2];
    }
    return SLAMGPU_OK;
}

int slamgpu_host_camera_center(const float* pose_colmajor16, float* out_center)
{
    SG_CHECK(pose_colmajor16 && out_center, "null pointer");
    host_camera_center(pose_colmajor16, out_center);
    return SLAMGPU_OK;
}

static void host_normalized3(const float v[3], float out[3])      // Eigen normalized(): z > 0 ? v / sqrt(z) : v
{
    volatile float a = v[0] * v[0], b = v[1] * v[1], c = v[2] * v[2];
    const float z = hsum3(a, b, c);
    if (z > 0.0f) {
        const float n = std::sqrt(z);
        out[0] = v[0] / n;
        out[1] = v[1] / n;
        out[2] = v[2] / n;
    } else {
        out[0] = v[0]; out[1] = v[1]; out[2] = v[2];
    }
}

// MapPoint::avg_viewing_normal / observed_distance_range (src/MapPoint.cpp:24-45) of one point, its observers'
// camera centres in the host's observation order
int slamgpu_host_point_stats(const float* xyz, int n_obs, const float* obs_center, float* out_normal, float* out_range)
{
    SG_CHECK(xyz && out_normal && out_range && n_obs >= 0 && (n_obs == 0 || obs_center), "bad argument");
    float acc[3] = {0.0f, 0.0f, 0.0f};
    float nearest = 3.402823466e+38f, furthest = 0.0f;
    for (int o = 0; o < n_obs; ++o) {
        const float* C = obs_center + 3 * static_cast<size_t>(o);
        const float d[3] = {xyz[0] - C[0], xyz[1] - C[1], xyz[2] - C[2]};
        float v[3], vv[3];
        host_normalized3(d, v);
        host_normalized3(v, vv);
        for (int k = 0; k < 3; ++k) {
            volatile float t = acc[k] + vv[k];
            acc[k] = t;
        }
        volatile float a = d[0] * d[0], b = d[1] * d[1], c = d[2] * d[2];
        const float dist = std::sqrt(hsum3(a, b, c));
        nearest = dist < nearest ? dist : nearest;
        furthest = furthest < dist ? dist : furthest;
    }
    host_normalized3(acc, out_normal);
    out_range[0] = nearest;
    out_range[1] = furthest;
    return SLAMGPU_OK;
}

int slamgpu_trace_read(slamgpu_ctx* ctx, double* out_us4, int64_t* out_calls)
{
    SG_CHECK(ctx != nullptr && out_us4 != nullptr, "null pointer");
    for (int i = 0; i < 4; ++i) {
        out_us4[i] = ctx->trace_us[i];
        ctx->trace_us[i] = 0.0;
    }
    if (out_calls) *out_calls = ctx->trace_calls;
    ctx->trace_calls = 0;
    return SLAMGPU_OK;
}

int slamgpu_bench_popc(slamgpu_ctx* ctx, double* popc_per_s, int* n_sm, double* sm_mhz_est)
{
    SG_CTX(ctx);
    return popc_microbench(ctx, popc_per_s, n_sm, sm_mhz_est);
}

} // extern "C"
