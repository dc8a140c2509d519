// Internal declarations shared by the libslamgpu translation units (not installed).
#pragma once

#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "slamgpu.h"

// SLAMGPU_TC_VARIANT (timing experiments: skip the MMAs, the TMEM reads, the rare path ...) exists only in builds
// made with -DSLAMGPU_DEBUG_VARIANTS.  In the release library the switches are compiled out: the bit tests below
// are constant-false and slamgpu_create does not read the variable, so no environment can corrupt a result.
#ifdef SLAMGPU_DEBUG_VARIANTS
#define SG_TC_VARIANT(value, bits) (((value) & (bits)) != 0)
#else
#define SG_TC_VARIANT(value, bits) false
#endif


namespace slamgpu {

void set_error(const std::string& msg);
int fail(int code, const std::string& msg);

#define SG_CUDA(expr)                                                                                   \
    do {                                                                                                \
        cudaError_t _e = (expr);                                                                        \
        if (_e != cudaSuccess) {                                                                        \
            return ::slamgpu::fail(SLAMGPU_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
        }                                                                                               \
    } while (0)

#define SG_CHECK(cond, msg)                                                        \
    do {                                                                           \
        if (!(cond)) return ::slamgpu::fail(SLAMGPU_ERR_INVALID, std::string(msg)); \
    } while (0)

#define SG_TRY(expr)              \
    do {                          \
        int _rc = (expr);         \
        if (_rc != SLAMGPU_OK) return _rc; \
    } while (0)

// ---------------------------------------------------------------------------------------------
// Packed (distance, index) key: a plain unsigned min gives the reference's rule "lowest distance,
// then lowest train index" (SURVEY 8a B3).  Distance <= 256 needs 9 bits -> 22 index bits + 1 spare.
// ---------------------------------------------------------------------------------------------
constexpr int kIdxBits = 22;
constexpr uint32_t kIdxMask = (1u << kIdxBits) - 1u;
constexpr uint32_t kKeyNone = 0xFFFFFFFFu;

// The KD-tree of a frame being built on a host thread (slamgpu_frame_upload) until a call needs it.
struct FrameKdPending {
    std::thread worker;
    std::atomic<int> done{0};         // set by the worker when the image is complete
    bool lazy = false;                // no worker: the tree is built by the first call that needs it (finalize_frame)
    std::vector<float> xy;            // private copy of the keypoints
    std::vector<uint32_t> perm;       // the tree as its permutation
    std::vector<uint8_t> image;       // host image of the KD-ordered device arrays
    float gx0 = 0.f, gy0 = 0.f, inv_cell = 1.f;
    int gw = 1, gh = 1;
};

// Device-resident frame (what match_via_reproject reads from `const Frame&`).
struct FrameDev {
    int nkp = 0;
    void* blob = nullptr;         // one allocation; the pointers below are carved out of it
    size_t blob_bytes = 0;
    float2* kd_xy = nullptr;      // [nkp] keypoint coordinates in KD position order
    int32_t* kd_idx = nullptr;    // [nkp] keypoint index at KD position m (= the permutation)
    uint4* desc = nullptr;        // [nkp*2] descriptors in keypoint order
    uint8_t* matched = nullptr;   // [nkp]
    float2* xy = nullptr;         // [nkp] keypoint order (diagnostics)
    unsigned long long* proposals = nullptr;  // [nkp] packed (dist << 32 | visit rank)
    // uniform grid over the keypoints (candidate lookup): cell c holds cell_items[cell_start[c] .. cell_start[c+1])
    int32_t* cell_start = nullptr;            // [gw*gh + 1]
    int32_t* cell_items = nullptr;            // [nkp] KD positions, cell-major
    int4* cell_rec = nullptr;                 // [nkp] cell-major (x bits, y bits, keypoint index, KD position)
    int32_t* cell_start_spec = nullptr;       // [gw*gh + 1] uploaded with the frame (same values as cell_start)
    int4* cell_rec_spec = nullptr;            // [nkp] the same without the tree: (x bits, y bits, keypoint index, keypoint index);
                                              // resident with the frame itself, so a match call need not wait for the KD build
    float gx0 = 0.f, gy0 = 0.f, inv_cell = 1.f;
    int gw = 1, gh = 1;
    std::vector<uint32_t> perm_host;
    uint8_t* kd_image = nullptr;              // first byte of [kd_xy][kd_idx][cell_start][cell_items][cell_rec]
    size_t kd_image_bytes = 0;
    std::shared_ptr<FrameKdPending> pending;  // set while the KD image is not yet resident
};

// Device-resident map-point set (what match_via_reproject reads from `MapPoint&`).
struct PointSetDev {
    int n = 0;
    void* blob = nullptr;         // one allocation; the pointers below are carved out of it
    int64_t n_obs = 0;
    float* pos = nullptr;         // [n*3]
    float* normal = nullptr;      // [n*3]
    float* range = nullptr;       // [n*2]
    int32_t* obs_offset = nullptr;// [n+1]
    uint4* obs_desc = nullptr;    // [n_obs*2]
    uint8_t* pt_matched = nullptr;// [cap_n] scratch of slamgpu_track_match: frame.is_matched(point)
    int cap_n = 0;                // the arrays above are carved for cap_n points / cap_obs observations,
    int64_t cap_obs = 0;          // so that slamgpu_points_append can grow the set in place
    // Observations of point i are obs_desc[obs_offset[i] .. obs_end[i]).  A set made by slamgpu_points_upload /
    // _append is CSR (obs_end = obs_offset + 1, one array).  A MUTABLE set (slamgpu_points_create, map_api.cu)
    // has two arrays and an observation arena with slack: a point whose observations change gets a new run at
    // the arena's top, removed points are tombstones (no observations, NaN position: they fail is_in_image),
    // and the arena / the slots are repacked in Map order when the garbage outweighs the live part.
    const int32_t* obs_end = nullptr;
    bool is_mutable = false;
    void* pts_blob = nullptr;     // mutable: [pos][normal][range][obs_offset][obs_end][pt_matched] for cap_n slots
    void* obs_blob = nullptr;     // mutable: [obs_desc][obs_kf] for cap_obs observations
    void* kf_blob = nullptr;      // mutable: [kf_center][kf_dirty] for cap_kf observers
    int32_t* obs_end_w = nullptr; // mutable: the writable obs_end
    int32_t* obs_kf = nullptr;    // [cap_obs] observer (key frame) id of every observation
    float* kf_center = nullptr;   // [cap_kf * 3] Frame::camera_center() of every observer, by id
    uint8_t* kf_dirty = nullptr;  // [cap_kf] set while an observer's centre has moved and its points are stale
    int n_kf = 0, cap_kf = 0;
    int n_live = 0;
    int64_t arena_top = 0;        // first free observation of the arena (n_obs = live observations)
    std::vector<int32_t> h_begin, h_count;   // host mirror of the runs, by slot
    std::vector<uint8_t> h_alive;
};
void free_points(PointSetDev& p);   // slamgpu.cu

// Device-resident DBoW2 vocabulary, re-laid out by CHILD POSITION: position c of the child array is one
// node, its descriptor sits at cdesc[c] (the children of a node are contiguous, so one level of the
// descent is one contiguous read), rec[c] = (first child position, child count, word id) of that node.
struct VocabDev {
    int n_nodes = 0, n_children = 0, n_words = 0;
    int root_begin = 0, root_count = 0, max_children = 0;
    void* blob = nullptr;
    int4* rec = nullptr;          // [n_children] (child_begin, n_children, word_id, 0)
    uint4* cdesc = nullptr;       // [n_children * 2]
    double* cweight = nullptr;    // [n_children]
    double* word_weight = nullptr;// [n_words] weight by word id
    std::vector<double> word_weight_host;   // the same on the host: the bag-of-words value fold runs there
};

// Resident bag-of-words database: concatenated (ids, values) with offsets, grown geometrically.
struct BowDbDev {
    int n_vectors = 0, cap_vectors = 0;
    int64_t n_entries = 0, cap_entries = 0;
    int32_t* ids = nullptr;
    double* vals = nullptr;
    int32_t* offset = nullptr;    // [cap_vectors + 1]
};

// Resident key-frame descriptor database of the loop detector (LoopDetector::Impl keeps every key frame,
// src/LoopDetector.cpp:314-330): concatenated rows + frame offsets, grown geometrically.
struct KfDbDev {
    uint8_t* desc = nullptr;            // [cap_rows x 32]
    int32_t* offset = nullptr;          // device copy of h_offset, [cap_frames + 1]
    int64_t cap_rows = 0;
    int cap_frames = 0;
    int max_rows = 0;                   // largest frame
    std::vector<int32_t> h_offset;      // [n_frames + 1], h_offset[0] = 0
    // the FP4-expanded copy the tensor-core scan reads (128 B per row, `exp4_pitch` rows per frame), kept with the
    // database: built on the first query that needs it and extended frame by frame afterwards, so a query expands
    // only the key frames appended since the last one (a larger frame than any before changes the pitch: rebuilt)
    uint8_t* exp4 = nullptr;
    size_t exp4_bytes = 0;
    int exp4_pitch = 0, exp4_frames = 0;
};

struct Arena {
    void* ptr = nullptr;
    size_t cap = 0;
};

} // namespace slamgpu

struct slamgpu_ctx {
    int device = 0;
    int n_sm = 0;
    cudaStream_t stream = nullptr;
    slamgpu::Arena d_ws;      // device scratch of the kernels (partial top-2 keys, frame argument tables)
    slamgpu::Arena d_io;      // device copies of the host-buffer API's inputs and outputs
    slamgpu::Arena h_stage;   // pinned staging
    int64_t launches = 0;
    bool profiling = false;
    int proj_lanes = 0;       // SLAMGPU_PROJ_LANES=8|32 overrides the lanes-per-point choice (0 = by size)
    bool proj_v2 = false;     // SLAMGPU_PROJ_V2=1: the previous group kernel (row-by-row candidate loop)
    float grid_cell = 16.0f;  // SLAMGPU_GRID_CELL: keypoint grid cell size in pixels (16: fewest slots per search at radius 15-20)
    bool proj_queue = true;   // SLAMGPU_PROJ_QUEUE=0: score candidates batch by batch instead of through the per-group queue
    int proj_two_pass = 0;    // SLAMGPU_PROJ_TWOPASS=1|2: force / forbid the gate + search form of K4 (0 = by visit count)
    int map_reserve = 1 << 18; // SLAMGPU_MAP_RESERVE: points reserved by a point set's first growth (4 observations each)
    bool kd_async = true;     // SLAMGPU_KD_ASYNC=0: build the KD-tree inside slamgpu_frame_upload instead of on a thread
    bool kd_speculate = true; // SLAMGPU_KD_SPECULATE=0: a match call waits for the KD-tree instead of running on the tree-free grid
    bool kfdb_cache_exp = true;   // SLAMGPU_KFDB_CACHE=0: the resident key-frame database is expanded anew by every query
    bool tc4_cross_one_scan = true;  // SLAMGPU_TC4_CROSS_ONE_SCAN=0: the tensor-core cross-check runs a row scan and a column scan
    bool tc4_fuse_finish = true;  // SLAMGPU_TC4_FUSE_FINISH=0: batches always run the finish kernel after the FP4 scan
    bool bf_push = true;          // SLAMGPU_BF_PUSH=0: batch calls into pinned result lists bring their lists back by host-queued copies
    bool bf_zero_copy = true;     // SLAMGPU_BF_ZEROCOPY=0: single-pair brute-force calls download their results with a copy
    bool proj_zero_copy = true;   // SLAMGPU_PROJ_ZEROCOPY=0: small projection calls copy their lists / results instead of
                                  // letting the kernels read and write the pinned staging buffer directly
    bool kd_lazy = true;      // SLAMGPU_KD_LAZY=0: always start the KD builder thread at upload (round 1's behaviour + speculation)
    int spec_tie_streak = 0;  // > 0 after a speculative call hit a tie: the next uploads build their tree eagerly again
    int kd_delay_ms = 0;      // SLAMGPU_KD_DELAY_MS: the KD builder sleeps first (tests: makes match calls find the tree missing)
    int64_t spec_calls = 0, spec_repeats = 0;   // speculative match calls, and how many had to be repeated with the tree
    bool proj_walk = false;   // SLAMGPU_PROJ_WALK=1: use the thread-per-point KD-walk kernel
    int bf_engine = 0;        // SLAMGPU_BF_ENGINE=int|tc overrides the brute-force engine choice (0 = by size, 1 = int, 2 = tc)
    int tc_variant = 0;       // SLAMGPU_TC_VARIANT: debug switches of the tensor-core scans; read only in -DSLAMGPU_DEBUG_VARIANTS builds
    bool tc4_narrow_last = false;  // SLAMGPU_TC4_NARROW=1: size the last train tile of the FP4 scan to the rows left (same results;
                                   // measured no faster: an MMA costs the same at N = 80 as at N = 240 and the narrow tile issues 9 instead of 5)
    bool tc4_teams = false;        // SLAMGPU_TC4_TEAMS=1: the FP4 scan's epilogue warps as two teams on alternate tiles (measured slower than
                                   // every warp reading its half of every tile first and folding afterwards; same results)
    bool grid_attr_set = false;                       // dynamic-smem opt-in of grid_build_kernel done on this context's device
    bool tc_attr_set = false, tc4_attr_set = false;   // dynamic-smem opt-in of the tensor-core scans done on this context's device
    slamgpu::Arena d_exp;     // expanded (int8) descriptors of the tensor-core scan
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events[4];   // per SLAMGPU_KERNEL_*
    std::vector<cudaEvent_t> prof_pool;
    std::map<int, slamgpu::FrameDev> frames;
    std::map<int, slamgpu::PointSetDev> point_sets;
    std::map<int, slamgpu::VocabDev> vocabs;
    std::map<int, slamgpu::BowDbDev> bow_dbs;
    std::map<int, slamgpu::KfDbDev> kf_dbs;
    bool bow_smem_set = false;
    bool trace_batch = false;                  // SLAMGPU_TRACE_BATCH=1
    double trace_us[4] = {0, 0, 0, 0};
    int64_t trace_calls = 0;
    int batch_slice = 0;                       // SLAMGPU_BATCH_SLICE: pairs per slice of the batch calls (0 = by scratch size)
    cudaStream_t copy_stream = nullptr;        // result copies of the batch call, overlapped with later slices
    bool grid_on_device = true;                // SLAMGPU_GRID_ON_DEVICE=0: slamgpu_frame_upload builds the tree-free grid on the host
    std::vector<int32_t> grid_scratch;         // slamgpu_frame_upload's counting sort (kept: no allocation per frame)
    cudaEvent_t sync_event = nullptr;          // wait_stream(): record + cudaEventSynchronize returns 0.5-1.3 us sooner than
                                               // cudaStreamSynchronize on this driver (tools/lat_probe.cu)
    std::vector<cudaEvent_t> slice_events;     // 3 per slice of slamgpu_match_descriptors_batch (created on demand)
};

namespace slamgpu {

// Block until everything queued on the context's stream has completed (the end of every host-buffer call).
int wait_stream(slamgpu_ctx* ctx);
int ensure_device_ws(slamgpu_ctx* ctx, size_t bytes);
int ensure_device_io(slamgpu_ctx* ctx, size_t bytes);
int ensure_host_stage(slamgpu_ctx* ctx, size_t bytes);

// carve 256-byte aligned pieces out of an arena
struct Carver {
    uint8_t* base;
    size_t off = 0;
    explicit Carver(void* p) : base(static_cast<uint8_t*>(p)) {}
    template <class T>
    T* take(size_t count)
    {
        T* r = reinterpret_cast<T*>(base + off);
        off += (count * sizeof(T) + 255) & ~size_t(255);
        return r;
    }
    static size_t padded(size_t bytes) { return (bytes + 255) & ~size_t(255); }
};

// ---- brute force (bf_kernels.cu) ------------------------------------------------------------
struct BfPlan {
    int r = 1;          // queries per thread
    int tile = 256;     // queries per CTA
    int tiles = 1;      // query tiles per pair
    int splits = 1;     // train splits per pair
    int t_per_split = 0;
};
BfPlan bf_make_plan(int n_sm, int nq_max, int nt_max, int64_t n_pairs);
size_t bf_partial_elems(const BfPlan& p, int64_t n_pairs, int max_rows);

// Scan: partial top-2 keys[(pair*splits + split)*max_rows + q] for every query row.
int bf_launch_scan(slamgpu_ctx* ctx, const BfPlan& plan, const uint8_t* d_query, int nq, const uint8_t* d_train,
                   int nt, const int32_t* d_frame_offset, const int32_t* d_pairs, int n_pairs, int max_rows,
                   bool swap_roles, uint2* d_keys, uint32_t* d_col_keys = nullptr);
// Finish, raw: merged top-2 -> idx/dist [nq x 2].
int bf_launch_finish_raw(slamgpu_ctx* ctx, const BfPlan& plan, const uint2* d_keys, int nq, int nt, int max_rows,
                         int32_t* d_idx, int32_t* d_dist);
// Finish, filtered (match_descriptors): out_train[pair][q] / count[pair].
int bf_launch_finish_filter(slamgpu_ctx* ctx, const BfPlan& plan, const uint2* d_keys, int nq, int nt,
                            const int32_t* d_frame_offset, const int32_t* d_pairs, int n_pairs, int max_rows,
                            float max_dist, float ratio, int32_t* d_out_train, int32_t* d_out_count,
                            cudaStream_t stream = nullptr /* default: the context's stream */);
// Cross-check mutuality from row partials and column partials.
int bf_launch_mutual(slamgpu_ctx* ctx, const BfPlan& row_plan, const uint2* d_row_keys, int row_stride,
                     const BfPlan& col_plan, const uint2* d_col_keys, int col_stride, int nq, int nt,
                     float strict_max_dist, int32_t* d_idx, int32_t* d_dist);
// Cross-check mutuality with the column minima of the fused scan (d_col_keys[nt]).
int bf_launch_mutual_direct(slamgpu_ctx* ctx, const BfPlan& row_plan, const uint2* d_row_keys, int row_stride,
                            const uint32_t* d_col_keys, int nq, float strict_max_dist, int32_t* d_idx, int32_t* d_dist);
// Ordered compaction of out_train[pair][q] into (query, train) lists.
// Short batches into host-visible lists: filter + per-pair compaction (one CTA per pair), then offsets + gather.
int bf_launch_finish_compact(slamgpu_ctx* ctx, const BfPlan& plan, const uint2* d_keys, const int32_t* d_frame_offset,
                             const int32_t* d_pairs, int n_pairs, int max_rows, float max_dist, float ratio,
                             int32_t* d_local /* 2 x n_pairs x max_rows */, int32_t* d_count, int32_t* d_offset,
                             int32_t* h_offset_mirror, int32_t* query_list, int32_t* train_list, int64_t capacity);
// slamgpu_knn2_dev_noalloc (bf_kernels.cu): resolve the fixed scratch array's device address (device current)
int bf_noalloc_setup(int device);
int bf_launch_scan_counts(slamgpu_ctx* ctx, const int32_t* d_count, int n, int32_t* d_offset /* n+1 */,
                          int32_t* h_offset_mirror = nullptr /* n+1, device-addressable host memory */,
                          cudaStream_t stream = nullptr);
int bf_launch_compact(slamgpu_ctx* ctx, const int32_t* d_out_train, const int32_t* d_frame_offset,
                      const int32_t* d_pairs, int nq_single, int n_pairs, int max_rows, const int32_t* d_offset,
                      int32_t* d_query_list, int32_t* d_train_list, int64_t capacity, cudaStream_t stream = nullptr);
// Entries [offset[0], offset[n_pairs]) of the two device lists -> the same positions of two device-addressable HOST
// lists (pinned memory), copied by SM stores: the range comes from the device's own offsets, so no host round trip
// sits between a slice's compaction and its way home.
int bf_launch_push_lists(slamgpu_ctx* ctx, const int32_t* d_offset, int n_pairs, const int32_t* d_query_list,
                         const int32_t* d_train_list, int32_t* h_query_list, int32_t* h_train_list, int64_t capacity,
                         cudaStream_t stream);

// ---- brute force on the tensor cores (bf_tc_kernels.cu) ---------------------------------------
// Expanded descriptors: 256 int8 per row in the UMMA canonical layout, frames at a pitch of
// bf_tc_pitch_rows(max_rows) rows.  The scan writes 2 * plan.splits partials per query row.
int bf_tc_pitch_rows(int max_rows);
int bf_tc_partials(const slamgpu_ctx* ctx);   // partial top-2 records per (pair, split, row) the scan writes
size_t bf_tc_expanded_bytes(int n_frames, int max_rows);
BfPlan bf_tc_make_plan(const slamgpu_ctx* ctx, int nq_max, int nt_max, int64_t n_pairs);
int bf_tc_launch_expand(slamgpu_ctx* ctx, const uint8_t* d_src, const int32_t* d_frame_offset, int n_frames, int n_rows,
                        int max_rows, int pitch_rows, uint8_t* d_dst);
int bf_tc_launch_expand_list(slamgpu_ctx* ctx, const uint8_t* d_src, const int32_t* d_frame_offset, const int32_t* d_frame_list,
                             int n_list, int max_rows, int pitch_rows, uint8_t* d_dst);
int bf_tc_launch_expand_pair(slamgpu_ctx* ctx, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int pitch_rows,
                             uint8_t* d_dst);
int bf_tc_launch_scan(slamgpu_ctx* ctx, const BfPlan& plan, const uint8_t* d_exp_q, int nq, const uint8_t* d_exp_t,
                      int nt, int pitch_rows, const int32_t* d_frame_offset, const int32_t* d_pairs, int n_pairs,
                      int max_rows, int d_cut, uint2* d_keys);
int ensure_device_exp(slamgpu_ctx* ctx, size_t bytes);

// ---- filtered brute force on the block-scaled FP4 tensor cores (bf_tc4_kernels.cu) --------------
// Expanded descriptors: 256 E2M1 values (128 bytes) per row, frames at bf_tc4_pitch_rows(max_rows) rows.
int bf_tc4_pitch_rows(int max_rows);
size_t bf_tc4_expanded_bytes(int n_frames, int max_rows);
BfPlan bf_tc4_make_plan(const slamgpu_ctx* ctx, int nq_max, int nt_max, int64_t n_pairs);
int bf_tc4_launch_expand(slamgpu_ctx* ctx, const uint8_t* d_src, const int32_t* d_frame_offset, int n_frames, int n_rows,
                         int pitch_rows, uint8_t* d_dst);
int bf_tc4_launch_expand_list(slamgpu_ctx* ctx, const uint8_t* d_src, const int32_t* d_frame_offset, const int32_t* d_frame_list,
                              int n_list, int pitch_rows, uint8_t* d_dst);
int bf_tc4_launch_expand_pair(slamgpu_ctx* ctx, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int pitch_rows,
                              uint8_t* d_dst);
// fuse != null (plan.splits == 1 only): the scan applies the match_descriptors filter itself and writes
// d_out_train [pair][max_rows] / d_out_count [pair] -- no key records, no finish launch
struct Tc4Fuse {
    float max_dist, ratio;
    int32_t* d_out_train;
    int32_t* d_out_count;
};
int bf_tc4_launch_scan(slamgpu_ctx* ctx, const BfPlan& plan, const uint8_t* d_exp_q, int nq, const uint8_t* d_exp_t,
                       int nt, int pitch_rows, const int32_t* d_frame_offset, const int32_t* d_pairs, int n_pairs,
                       int max_rows, int d_cut, uint2* d_keys, const Tc4Fuse* fuse = nullptr,
                       // single pair: every hit also competes for its train row's column minimum (distance << 22 | query
                       // row) in d_col_keys [nt], which the caller has set to all-ones: the cross-check in one scan
                       uint32_t* d_col_keys = nullptr);

// ---- projection (proj_kernels.cu) -----------------------------------------------------------
struct ProjFrameArgs {
    int nkp;
    const float2* kd_xy;
    const int32_t* kd_idx;
    const uint4* desc;
    const uint8_t* matched;
    const float2* xy;
    unsigned long long* proposals;
    const int32_t* cell_start;
    const int32_t* cell_items;
    const int4* cell_rec;
    float gx0, gy0, inv_cell;
    int gw, gh;
    int kd_ready;            // 0: speculative call before the KD-tree is built -- cell_rec holds (x, y, keypoint, keypoint),
    unsigned int* ties;      //    kd_xy / kd_idx are not read and ties at the minimum are counted here instead of settled
    unsigned int* ties_out;  // zero-copy call: where the emit kernel leaves the tie count (mapped host memory) or null
    const uint8_t* skip;     // [n_points] for this frame or null
    int32_t* out_point;      // [nkp]
    int32_t* out_dist;       // [nkp]
    float M[12];             // K * pose[0:3,0:4], row-major
    float C[3];              // camera centre
};
struct ProjCommonArgs {
    int n_points;
    const float* pos;
    const float* normal;
    const float* range;
    const int32_t* obs_offset;    // first observation of a point ...
    const int32_t* obs_end;       // ... and one past its last (obs_offset + 1 for a CSR set)
    const uint4* obs_desc;
    const int32_t* order;
    int n_visit;
    float width_f, height_f, radius, r2, min_cos, nearer, further;
    int thr;        // accept Hamming d iff d < thr  (thr = ceil(max_dist))
    int replace;
    float* dbg_uv;
    int32_t* dbg_kp;
    int32_t* dbg_d;
    unsigned long long* stats;
};
int proj_launch(slamgpu_ctx* ctx, const ProjCommonArgs& common, const ProjFrameArgs* h_frames,
                const ProjFrameArgs* d_frames, int n_frames, int max_nkp);
// Tracker::match_with_last_key_frame + match_with_map chained on the stream (src/Tracker.cpp:232-248):
// stage 1 visits kf_order (points observed by the last key frame), its matches are applied on the
// device (Frame::add_map_match, src/Frame.cpp:80-102), stage 2 visits the whole set.
int proj_launch_track(slamgpu_ctx* ctx, const ProjCommonArgs& common, const ProjFrameArgs& frame, uint8_t* d_pt_matched,
                      uint8_t* d_kp_matched, const int32_t* d_matched_points, int n_matched, const int32_t* d_kf_order,
                      int n_kf, int32_t* d_out /* [4 x nkp]: kf point, kf dist, map point, map dist */);
int proj_launch_flags(slamgpu_ctx* ctx, const int32_t* d_idx, int n, uint8_t* d_flags, int n_flags);
int proj_launch_point_stats(slamgpu_ctx* ctx, const float* d_pos, int n, const int32_t* d_obs_offset,
                            const float* d_obs_center, float* d_normal, float* d_range);
int proj_launch_grid_build(slamgpu_ctx* ctx, const float2* d_xy, int n, float gx0, float gy0, float inv_cell, int gw, int gh,
                           int32_t* d_cell_start, int4* d_cell_rec, bool* done);
int proj_launch_kd_gather(slamgpu_ctx* ctx, const float2* d_xy, const int32_t* d_perm, int n, float2* d_kd_xy);

// host-side float helpers that follow the oracle's / Eigen's operation order
void host_projection_matrix(const float* pose_colmajor, float fx, float fy, float cx, float cy, float* M12);
void host_camera_center(const float* pose_colmajor, float* C3);

// RAII bracket: records an event pair around a launch when ctx->profiling is on
struct ProfScope {
    slamgpu_ctx* ctx;
    int kernel;
    cudaEvent_t stop = nullptr;
    ProfScope(slamgpu_ctx* c, int k);
    ~ProfScope();
};

void free_all_vocabs(slamgpu_ctx* ctx);   // voc_api.cu
// ---- bag of words (voc_kernels.cu) -----------------------------------------------------------
int voc_launch_transform(slamgpu_ctx* ctx, const VocabDev& v, const uint8_t* d_desc, int n, int32_t* d_word,
                         double* d_weight);
// one CTA per image: sort the image's words, run-length encode -> (id, count) per distinct word, ascending
int voc_launch_bow_count(slamgpu_ctx* ctx, const int32_t* d_word, const int32_t* d_image_offset, int n_images,
                         int32_t* d_ids, int32_t* d_counts, int32_t* d_n_distinct);
int voc_launch_score_l1(slamgpu_ctx* ctx, const int32_t* d_q_ids, const double* d_q_vals, int nq,
                        const int32_t* d_db_ids, const double* d_db_vals, const int32_t* d_db_offset, int n_db,
                        double* d_score);

// ---- PnP verification support (pnp_kernels.cu) ------------------------------------------------
void free_all_kfdbs(slamgpu_ctx* ctx);
int pnp_launch_score(slamgpu_ctx* ctx, const float* d_object, const float* d_image, const int32_t* d_corr_begin,
                     const int32_t* d_corr_count, const double* d_hyp_rt, const int32_t* d_hyp_problem, int n_hyp,
                     const double* K4, double reproj_error, int32_t* d_out_count);
int pnp_launch_mask(slamgpu_ctx* ctx, const float* d_object, const float* d_image, const int32_t* d_corr_begin,
                    const int32_t* d_corr_count, int n_problems, const double* d_hyp_rt, const int32_t* d_sel_hyp,
                    const double* K4, double reproj_error, uint8_t* d_out_mask, float* d_out_err,
                    int32_t* d_out_inliers);
int pnp_launch_assemble(slamgpu_ctx* ctx, const int32_t* d_match_train, const int32_t* d_frame_offset,
                        const int32_t* d_pairs, int n_pairs, int max_rows, const float* d_row_pos,
                        const float* d_row_xy, float* d_out_object, float* d_out_image, int32_t* d_out_query,
                        int32_t* d_out_count);

int popc_microbench(slamgpu_ctx* ctx, double* popc_per_s, int* n_sm, double* sm_mhz_est);

} // namespace slamgpu
