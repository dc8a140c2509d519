/* slamgpu.h -- C ABI of libslamgpu: B200-native (sm_100a) ORB descriptor matching for Racing-SLAM.
 *
 * This is the drop-in boundary for the ONE hot path this library replaces: binary descriptor
 * matching.  Every entry point names the reference interface it stands in for (paths relative
 * to the reference tree).  Plain pointers and sizes only; no C++/torch types.
 *
 * Conventions
 *   - Descriptors are 32-byte ORB rows (cv::Mat N x 32 CV_8U, src/features/FeatureExtractor.h:15-18).
 *   - All functions return SLAMGPU_OK (0) or an error code; slamgpu_last_error() gives the text
 *     (thread-local).  The reference has no error channel (an empty vector means "no matches",
 *     src/MapMatcher.cpp:139-141); the adapter maps errors to exceptions.
 *   - One context = one CUDA device + one stream + pinned staging.  Calls on one context are
 *     stream-ordered; the host-buffer entry points return after their results are in the
 *     caller's buffers.  The *_dev entry points take DEVICE pointers (16-byte aligned), enqueue on
 *     the context's stream and return without synchronising.
 *   - There is no CPU fallback: without a CUDA device slamgpu_create() fails.
 *   - "missing" indices / distances are -1.
 */
#ifndef SLAMGPU_H
#define SLAMGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SLAMGPU_ABI_VERSION 2

#if defined(__GNUC__)
#define SLAMGPU_API __attribute__((visibility("default")))
#else
#define SLAMGPU_API
#endif

enum {
    SLAMGPU_OK = 0,
    SLAMGPU_ERR_INVALID = 1,  /* bad argument (null pointer, negative size, size over the limits below) */
    SLAMGPU_ERR_CUDA = 2,     /* a CUDA runtime call failed; text in slamgpu_last_error() */
    SLAMGPU_ERR_NOMEM = 3,
    SLAMGPU_ERR_NOTFOUND = 4  /* unknown frame_id / set_id */
};

/* Largest train set per brute-force call: packed (distance, index) keys keep 22 index bits. */
#define SLAMGPU_MAX_TRAIN (1 << 22)

typedef struct slamgpu_ctx slamgpu_ctx;

SLAMGPU_API int slamgpu_abi_version(void);
SLAMGPU_API const char* slamgpu_last_error(void);
SLAMGPU_API int slamgpu_device_count(int* count);

/* Context: replaces nothing in the reference (which is stateless, src/MapMatcher.h:16-38);
 * it is what the three MapMatcher owners (src/Tracker.cpp:42, src/Mapper.cpp:87,
 * src/LoopDetector.cpp:319) would each hold. */
SLAMGPU_API int slamgpu_create(int device, slamgpu_ctx** ctx);
SLAMGPU_API int slamgpu_destroy(slamgpu_ctx* ctx);
SLAMGPU_API int slamgpu_sync(slamgpu_ctx* ctx);
SLAMGPU_API void* slamgpu_stream(slamgpu_ctx* ctx); /* the context's cudaStream_t, for interop and timing */
/* Kernels launched on this context since creation (bench.py's gpu_launches). */
SLAMGPU_API int64_t slamgpu_launch_count(slamgpu_ctx* ctx);

/* Page-locked host memory for callers that want zero-staging uploads (optional). */
SLAMGPU_API int slamgpu_host_alloc(void** ptr, size_t bytes);
SLAMGPU_API int slamgpu_host_free(void* ptr);

/* ------------------------------------------------------------------------------------------
 * Brute force
 * ---------------------------------------------------------------------------------------- */

/* cv::BFMatcher(NORM_HAMMING).knnMatch(query, train, knn, k=2)  -- src/MapMatcher.cpp:145-148.
 * out_idx / out_dist are nq x 2: the first two of a stable sort by (distance, trainIdx).
 * With nt == 1 the second column is -1 (the reference asks for k = 1 then, :145). */
SLAMGPU_API int slamgpu_knn2(slamgpu_ctx* ctx, const uint8_t* query, int nq, const uint8_t* train, int nt,
                 int32_t* out_idx, int32_t* out_dist);

/* cv::BFMatcher::create(NORM_HAMMING, crossCheck=true)->knnMatch(query, train, m, 1)
 * -- src/features/FeatureExtractor.cpp:20-21.  out_idx[q] = t iff t is q's nearest train row and
 * q is t's nearest query row (lowest index wins ties both ways), else -1. */
SLAMGPU_API int slamgpu_crosscheck(slamgpu_ctx* ctx, const uint8_t* query, int nq, const uint8_t* train, int nt,
                       int32_t* out_idx, int32_t* out_dist);

/* MapMatcher::match_descriptors  -- src/MapMatcher.cpp:129-163.
 * query = frame.features().descriptors; train = the key frame's MATCHED keypoints' descriptors in
 * ascending keypoint order (:135-138).  Keeps a row unless d0 > max_dist (:152) or, when nt >= 2,
 * d0 > ratio * d1 in float (:156).  Emits (queryIdx, trainIdx) in ascending query order; the
 * adapter turns trainIdx into points[trainIdx].  Buffers hold nq entries. */
SLAMGPU_API int slamgpu_match_descriptors(slamgpu_ctx* ctx, const uint8_t* query, int nq, const uint8_t* train, int nt,
                              float max_dist, float ratio, int32_t* out_query, int32_t* out_train,
                              int* out_count);

/* BaseFeatureExtractor::match_features(prev, cur)  -- src/features/FeatureExtractor.cpp:15-30.
 * Cross-check with query = cur, train = prev; keeps distance < max_dist (strict, :25); emits
 * FeatureMatch(train_index, query_index) in ascending query order.  Buffers hold ncur entries. */
SLAMGPU_API int slamgpu_match_features(slamgpu_ctx* ctx, const uint8_t* prev, int nprev, const uint8_t* cur, int ncur,
                           float max_dist, int32_t* out_train, int32_t* out_query, int* out_count);

/* Device-resident forms of the above (inputs and outputs already in HBM). */
SLAMGPU_API int slamgpu_knn2_dev(slamgpu_ctx* ctx, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                     int32_t* d_out_idx, int32_t* d_out_dist);
SLAMGPU_API int slamgpu_crosscheck_dev(slamgpu_ctx* ctx, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                           int32_t* d_out_idx, int32_t* d_out_dist);

/* slamgpu_knn2_dev that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same results as slamgpu_knn2_dev (integer XOR + POPC scan; the outputs are
 * bit-identical to every other engine's).  The scan's partial keys live in a fixed 8 MiB device array inside the
 * library instead of the context's growable workspace, hence the limit: nq <= 8192 and nt <= 8192, anything larger
 * is refused with SLAMGPU_ERR_INVALID and a message (use slamgpu_knn2_dev).  Every GPU that loads the library holds
 * the array until exit, and every call on that GPU shares it: calls on one context are stream-ordered and safe,
 * calls that overlap on the GPU (different contexts / streams) RACE on the scratch -- serialise them.  The call
 * queues two kernels on the context's stream (slamgpu_stream) and returns; it calls no cudaMalloc, cudaFree or any
 * allocating routine.  The scratch address is resolved once per device inside slamgpu_create. */
SLAMGPU_API int slamgpu_knn2_dev_noalloc(slamgpu_ctx* ctx, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                             int32_t* d_out_idx, int32_t* d_out_dist);
/* d_out_train[q] = trainIdx of the surviving match or -1; *d_out_count (device int) = survivors. */
SLAMGPU_API int slamgpu_match_descriptors_dev(slamgpu_ctx* ctx, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                                  float max_dist, float ratio, int32_t* d_out_train, int32_t* d_out_count);

/* Batched match_descriptors over a key-frame database -- the LoopDetector all-pairs verification
 * (src/LoopDetector.cpp:176-229 calls match_descriptors once per (query, candidate); BASELINE C5).
 *   d_desc          all frames' descriptors, concatenated
 *   d_frame_offset  [n_frames + 1] row offsets into d_desc
 *   d_pairs         [n_pairs x 2] (query_frame, train_frame)
 *   max_rows        upper bound of rows per frame (sizes the outputs)
 *   d_out_train     [n_pairs x max_rows] trainIdx per query row or -1
 *   d_out_count     [n_pairs] surviving matches per pair
 * Pairs are independent: shard them across GPUs by giving each context its own slice. */
SLAMGPU_API int slamgpu_match_descriptors_batch_dev(slamgpu_ctx* ctx, const uint8_t* d_desc, const int32_t* d_frame_offset,
                                        int n_frames, const int32_t* d_pairs, int n_pairs, int max_rows,
                                        float max_dist, float ratio, int32_t* d_out_train, int32_t* d_out_count);
/* Host-buffer form: uploads desc / offsets / pairs, returns per-pair counts and a compacted
 * (pair-major, ascending query) list of (queryIdx, trainIdx); out_offset is [n_pairs + 1].
 * out_query / out_train hold `capacity` entries; *out_total is the number produced (if it exceeds
 * capacity the call fails with SLAMGPU_ERR_NOMEM and *out_total tells the size needed). */
SLAMGPU_API int slamgpu_match_descriptors_batch(slamgpu_ctx* ctx, const uint8_t* desc, const int32_t* frame_offset,
                                    int n_frames, const int32_t* pairs, int n_pairs, float max_dist, float ratio,
                                    int32_t* out_offset, int32_t* out_query, int32_t* out_train, int64_t capacity,
                                    int64_t* out_total);

/* Resident key-frame database: LoopDetector keeps every key frame it has seen and verifies the newest one
 * against retrieved candidates (src/LoopDetector.cpp:314-330 the store, :176-229 verify_pnp ->
 * match_descriptors once per (query, candidate)).  A key frame's descriptors are uploaded ONCE, when it is
 * appended; the match call then sends only the pair list.  Frames are numbered 0, 1, ... in append order
 * (*out_frame, may be NULL); same results, limits and errors as slamgpu_match_descriptors_batch on the
 * concatenation of the appended frames. */
SLAMGPU_API int slamgpu_kfdb_append(slamgpu_ctx* ctx, int db_id, const uint8_t* desc, int n_rows, int* out_frame);
SLAMGPU_API int slamgpu_kfdb_count(slamgpu_ctx* ctx, int db_id, int* n_frames, int64_t* n_rows);
SLAMGPU_API int slamgpu_kfdb_match_pairs(slamgpu_ctx* ctx, int db_id, const int32_t* pairs, int n_pairs, float max_dist,
                             float ratio, int32_t* out_offset, int32_t* out_query, int32_t* out_train,
                             int64_t capacity, int64_t* out_total);
SLAMGPU_API int slamgpu_kfdb_release(slamgpu_ctx* ctx, int db_id);

/* ------------------------------------------------------------------------------------------
 * Projection-guided search
 * ---------------------------------------------------------------------------------------- */

/* Constants of src/MapMatcher.cpp:12-18 and the camera of src/Camera.h:8-25, as runtime values. */
typedef struct slamgpu_proj_params {
    float fx, fy, cx, cy;   /* Camera::get_intrinsic_matrix(), src/Camera.cpp:5-13 */
    int32_t width, height;  /* Camera::get_width()/get_height() */
    float radius;           /* SEARCH_RADIUS = 20            src/MapMatcher.cpp:12 */
    float max_dist;         /* m_max_descriptor_distance = 64 src/features/OrbFeatureExtractor.h:14-17 */
    float min_cos;          /* MIN_VIEWING_ANGLE_COSINE = 0.5 :13 */
    float nearer;           /* MAX_NEARER_RATIO = 2.0         :16 */
    float further;          /* MAX_FURTHER_RATIO = 1.25       :17 */
    int32_t replace;        /* 0 = match_map / match_key_frame (:173), 1 = match_for_fuse (:124) */
} slamgpu_proj_params;

/* Frame state kept resident on the device: what match_via_reproject reads from `const Frame&`
 * (src/MapMatcher.cpp:45-98): keypoints (src/Frame.h:43), descriptors (:40), the KD-tree
 * (src/Frame.cpp:11-15) and is_matched(size_t) (src/Frame.cpp:143-146).
 *   kd_perm  the KD-tree as its permutation: perm[(s+e)/2] is the node of range [s,e)
 *            (src/KDTree.cpp:18-43).  Pass NULL to have the library build it with the same
 *            std::nth_element call; pass your own when the host's libstdc++ must decide ties.
 *   matched  [nkp] bytes, non-zero = keypoint already has a map point; NULL = none matched.
 * The call returns once keypoints, descriptors and flags are resident (the arrays may be reused).  When the
 * library builds the tree itself it does so on a host thread of its own and the first call that reads the tree
 * (any match_projection / track_match / get_kd_perm on this frame) waits for it: upload a frame as soon as its
 * keypoints exist and the build overlaps whatever comes before matching.  SLAMGPU_KD_ASYNC=0 builds it inside
 * the call; SLAMGPU_KD_THREADS (default 4) is the number of host threads one build forks.
 * Uploading to a frame_id that is in use replaces it (and reuses its device allocation when large enough). */
SLAMGPU_API int slamgpu_frame_upload(slamgpu_ctx* ctx, int frame_id, int nkp, const float* xy, const uint8_t* desc,
                         const uint32_t* kd_perm, const uint8_t* matched);
SLAMGPU_API int slamgpu_frame_set_matched(slamgpu_ctx* ctx, int frame_id, const uint8_t* matched);
SLAMGPU_API int slamgpu_frame_release(slamgpu_ctx* ctx, int frame_id);
/* The permutation the library built / was given (for tests and for hosts that want to reuse it). */
SLAMGPU_API int slamgpu_frame_get_kd_perm(slamgpu_ctx* ctx, int frame_id, uint32_t* out_perm);
/* Build the permutation on the host exactly as KDTree2D::build does (src/KDTree.cpp:8-43). */
SLAMGPU_API int slamgpu_kd_build_host(const float* xy, int n, uint32_t* out_perm);

/* Map-point state kept resident: what match_via_reproject reads from `MapPoint&`:
 * position() (src/MapPoint.h:16), avg_viewing_normal() (src/MapPoint.cpp:24-33),
 * observed_distance_range() (:35-45) and one descriptor per observation, flattened in the host's
 * observations() iteration order (src/MapPoint.h:38, src/MapMatcher.cpp:86-87).
 *   normal / range may be NULL when obs_center is given: the library then derives them on the
 *   device from the observers' camera centres (src/Frame.cpp:39-42) in the same order. */
SLAMGPU_API int slamgpu_points_upload(slamgpu_ctx* ctx, int set_id, int n, const float* pos, const float* normal,
                          const float* range, const int32_t* obs_offset, const uint8_t* obs_desc,
                          const float* obs_center);
/* Map growth (Mapper::insert -> Map::create_point, src/Mapper.cpp:152-174, src/Map.cpp:20-40): append
 * n_new points at the end of the set, i.e. in Map's own vector order (src/Map.cpp:164-190).  Same
 * arguments as slamgpu_points_upload for the appended block alone; obs_offset is relative to it
 * (obs_offset[0] = 0).  Existing indices stay valid; storage grows geometrically on the device. */
SLAMGPU_API int slamgpu_points_append(slamgpu_ctx* ctx, int set_id, int n_new, const float* pos, const float* normal,
                          const float* range, const int32_t* obs_offset, const uint8_t* obs_desc,
                          const float* obs_center);
SLAMGPU_API int slamgpu_points_count(slamgpu_ctx* ctx, int set_id, int* n_points, int64_t* n_observations);
SLAMGPU_API int slamgpu_points_release(slamgpu_ctx* ctx, int set_id);
/* Read back the per-point normal [n x 3] and range [n x 2] held on the device. */
SLAMGPU_API int slamgpu_points_get_stats(slamgpu_ctx* ctx, int set_id, float* out_normal, float* out_range);

/* ---- Mutable map (SURVEY 8f N1 / N2 "with incremental updates") ----------------------------------------
 * The reference mutates existing points between matcher calls; a set made by slamgpu_points_create follows it
 * with DELTAS instead of a re-upload of the whole map:
 *   Map::associate / disassociate (src/Map.cpp:97-124, src/MapPoint.cpp:47-55)  -> slamgpu_points_set_observations
 *   Map::fuse (src/Map.cpp:78-95; Mapper::fuse_match src/Mapper.cpp:176-196)    -> _set_observations (kept) + _remove (discarded)
 *   Map::remove_point (src/Map.cpp:63-76; Mapper::cull_points src/Mapper.cpp:396-430) -> slamgpu_points_remove
 *   MapPoint::set_position (src/MapPoint.cpp:19-22; src/Optimization.cpp:371,534, src/Mapper.cpp:391) -> _set_position
 *   a key frame's pose moved by bundle adjustment / pose graph                  -> slamgpu_observer_centers_update
 *   Map::create_point (src/Map.cpp:20-61)                                       -> slamgpu_points_append_observed
 * SLOTS: a point's index in every match call.  Slots are handed out in append order = Map's vector order
 * (src/Map.cpp:164-190), which is the visit rank that decides proposal ties (src/MapMatcher.cpp:95-97).  A removed
 * point's slot becomes a tombstone that matches nothing; slamgpu_points_compact squeezes tombstones out WITHOUT
 * reordering (exactly what vector::erase from the middle does to the indices of the later points) and reports
 * the renumbering.  OBSERVERS: every key frame that observes a point has a small id chosen by the caller and a
 * camera centre (Frame::camera_center(), src/Frame.cpp:39-42); normals and ranges (src/MapPoint.cpp:24-45) are
 * re-derived on the device, in the observation order the caller gives (the host's observations() iteration
 * order), for exactly the points an update touches.  A mutable set_id works with every match call above.
 * slamgpu_points_upload / _append do not apply to it. */
SLAMGPU_API int slamgpu_points_create(slamgpu_ctx* ctx, int set_id, int reserve_points, int64_t reserve_observations,
                          int reserve_observers);
/* New observers, or observers whose pose changed: centers [n x 3].  Every point that observes a moved key frame
 * gets its normal and range re-derived.  Ids are distinct within a call; an id must have a centre before an
 * observation may name it. */
SLAMGPU_API int slamgpu_observer_centers_update(slamgpu_ctx* ctx, int set_id, const int32_t* observer_ids, int n,
                                    const float* centers);
/* Append n_new points (slots n .. n + n_new - 1): pos [n_new x 3], obs_offset [n_new + 1] relative to the block,
 * obs_desc [x 32] and obs_observer [one id per observation] in the host's observations() order. */
SLAMGPU_API int slamgpu_points_append_observed(slamgpu_ctx* ctx, int set_id, int n_new, const float* pos,
                                   const int32_t* obs_offset, const uint8_t* obs_desc, const int32_t* obs_observer);
/* Replace the WHOLE observation run of each listed slot (the host's current iteration order; an
 * unordered_map may reorder on insert / erase, so the run is restated, not patched). */
SLAMGPU_API int slamgpu_points_set_observations(slamgpu_ctx* ctx, int set_id, const int32_t* idx, int n,
                                    const int32_t* obs_offset, const uint8_t* obs_desc, const int32_t* obs_observer);
SLAMGPU_API int slamgpu_points_set_position(slamgpu_ctx* ctx, int set_id, const int32_t* idx, int n, const float* pos);
/* Tombstone the listed slots (distinct, live). */
SLAMGPU_API int slamgpu_points_remove(slamgpu_ctx* ctx, int set_id, const int32_t* idx, int n);
/* Drop the tombstones, keep the order: out_remap [slot count before] = new slot or -1 (may be NULL), *out_n = slots
 * after.  The caller renumbers its own point -> slot table the same way (or calls this and reads out_remap). */
SLAMGPU_API int slamgpu_points_compact(slamgpu_ctx* ctx, int set_id, int32_t* out_remap, int* out_n);
/* Live points, observation-arena entries in use and how many of them are garbage (superseded runs). */
SLAMGPU_API int slamgpu_points_live(slamgpu_ctx* ctx, int set_id, int* live_points, int64_t* arena_used,
                        int64_t* arena_garbage);

/* MapMatcher::match_map / match_key_frame / match_for_fuse -- src/MapMatcher.cpp:107-127,165-175.
 *   skip     [n] bytes or NULL: non-zero drops the point up front -- frame.is_matched(point)
 *            (:53) and, for match_key_frame, !point.is_observed_by(kf) (:169)
 *   order    [n_order] indices into the point set or NULL (= every point, set order): the visit
 *            order that decides proposal ties (:95, src/Map.cpp:164-190); negative entries model
 *            match_for_fuse's nullptr (:119-121)
 *   out_point[kp] = point-set index matched to keypoint kp or -1; out_dist[kp] = Hamming distance.
 *   accepted_matches (:34-43) is the list of kp with out_point[kp] >= 0, ascending. */
SLAMGPU_API int slamgpu_match_projection(slamgpu_ctx* ctx, int frame_id, int set_id, const float* pose_colmajor16,
                             const slamgpu_proj_params* params, const uint8_t* skip, const int32_t* order,
                             int n_order, int32_t* out_point, int32_t* out_dist);

/* Same with `skip` as a LIST: matched_points = the indices of the points with frame.is_matched(point)
 * (src/MapMatcher.cpp:53) -- at most one per keypoint, so a call against a large resident set uploads
 * O(keypoints + n_order) instead of a mask over the whole set.  match_key_frame's !is_observed_by(kf) (:169)
 * is expressed through `order` = the key frame's points, ascending. */
SLAMGPU_API int slamgpu_match_projection_lists(slamgpu_ctx* ctx, int frame_id, int set_id, const float* pose_colmajor16,
                                   const slamgpu_proj_params* params, const int32_t* matched_points, int n_matched,
                                   const int32_t* order, int n_order, int32_t* out_point, int32_t* out_dist);

/* Same, with diagnostics: per visited point its projection dbg_uv [n_visit x 2] (NaN if skipped up
 * front), chosen keypoint dbg_kp and distance dbg_d (-1 = proposes nothing); stats[0] = projections
 * within 1e-4 px of an image border, stats[1] = (point, keypoint) pairs within 1e-4 px of the
 * search radius -- the candidates the north-star tolerance exempts from bit-exactness.  Any of
 * the four may be NULL. */
SLAMGPU_API int slamgpu_match_projection_ex(slamgpu_ctx* ctx, int frame_id, int set_id, const float* pose_colmajor16,
                                const slamgpu_proj_params* params, const uint8_t* skip, const int32_t* order,
                                int n_order, int32_t* out_point, int32_t* out_dist, float* dbg_uv,
                                int32_t* dbg_kp, int32_t* dbg_d, int64_t* stats);

/* The per-frame tracking pair, chained on the device -- Tracker::match_with_last_key_frame then
 * Tracker::match_with_map (src/Tracker.cpp:232-248):
 *   1. match_key_frame(frame, map, last_key_frame): visits kf_points = the indices of the points with
 *      is_observed_by(last_key_frame) (src/MapMatcher.cpp:169), strictly ascending = map order;
 *   2. every match is applied as Frame::add_map_match does (src/Frame.cpp:80-102): the keypoint and the
 *      point become matched in the frame;
 *   3. match_map(frame, map): visits every point of the set against the updated frame, and its matches
 *      are applied too, so the resident frame ends up as the host's frame after both loops.
 * matched_points = indices of the points with frame.is_matched(point) before the call (from
 * track_from_last_frame, src/Tracker.cpp:197-230); the frame's keypoint flags are the resident ones
 * (slamgpu_frame_upload / slamgpu_frame_set_matched).  replace in params is ignored (0).  One upload of
 * the two index lists, one download of the four [nkp] result arrays; results are those of the two
 * reference calls made one after the other. */
SLAMGPU_API int slamgpu_track_match(slamgpu_ctx* ctx, int frame_id, int set_id, const float* pose_colmajor16,
                        const slamgpu_proj_params* params, const int32_t* matched_points, int n_matched,
                        const int32_t* kf_points, int n_kf, int32_t* out_kf_point, int32_t* out_kf_dist,
                        int32_t* out_map_point, int32_t* out_map_dist);

/* Batched over key frames on a frozen snapshot -- Mapper::fuse_loop's window loop
 * (src/Mapper.cpp:208-219; BASELINE C4).  n_frames resident frames, one pose each; one shared
 * point set / order / params; skip is [n_frames x n] or NULL.  Outputs are concatenated per
 * frame in the order given (sum of the frames' nkp entries).  Key frames are independent:
 * shard them across GPUs by giving each context its own slice. */
SLAMGPU_API int slamgpu_match_projection_batch(slamgpu_ctx* ctx, const int* frame_ids, int n_frames, int set_id,
                                   const float* poses_colmajor16, const slamgpu_proj_params* params,
                                   const uint8_t* skip, const int32_t* order, int n_order, int32_t* out_point,
                                   int32_t* out_dist);

/* ------------------------------------------------------------------------------------------
 * Bag of words: the step before the brute-force verification in LoopDetector --
 * bow_of (src/LoopDetector.cpp:85-92) -> DBoW2 TemplatedVocabulary::transform
 * (thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1064-1120,1218-1257, FORB::distance FORB.cpp:81-101)
 * and vocabulary->score (src/LoopDetector.cpp:369) -> L1Scoring::score (ScoringObject.cpp:23-68).
 * ---------------------------------------------------------------------------------------- */

/* DBoW2 enums, same values: WeightingType (BowVector.h:36-42) and the norm the scoring object asks for
 * (ScoringObject mustNormalize: L1_NORM scoring -> SLAMGPU_BOW_NORM_L1). */
#define SLAMGPU_BOW_TF_IDF 0
#define SLAMGPU_BOW_TF 1
#define SLAMGPU_BOW_IDF 2
#define SLAMGPU_BOW_BINARY 3
#define SLAMGPU_BOW_NORM_NONE 0
#define SLAMGPU_BOW_NORM_L1 1
#define SLAMGPU_BOW_NORM_L2 2
#define SLAMGPU_BOW_MAX_FEATURES 16384 /* features per image in slamgpu_vocab_bow */

/* Vocabulary tree kept resident (TemplatedVocabulary::m_nodes): node 0 is the root;
 *   child_offset [n_nodes + 1], child_ids [child_offset[n_nodes]]  children of node i in DBoW2's own order
 *                (Node::children; transform() takes the FIRST child with the strictly smallest distance)
 *   desc [n_nodes x 32]   Node::descriptor (the root's is ignored)
 *   weight [n_nodes]      Node::weight (the word weight, e.g. idf; read at the leaves)
 *   word_id [n_nodes]     Node::word_id at the leaves, ignored for inner nodes */
SLAMGPU_API int slamgpu_vocab_upload(slamgpu_ctx* ctx, int vocab_id, int n_nodes, const int32_t* child_offset,
                         const int32_t* child_ids, const uint8_t* desc, const double* weight, const int32_t* word_id);
SLAMGPU_API int slamgpu_vocab_release(slamgpu_ctx* ctx, int vocab_id);

/* transform(feature, word_id, weight) for n descriptors (TemplatedVocabulary.h:1218-1257): out_word [n],
 * out_weight [n] (may be NULL). */
SLAMGPU_API int slamgpu_vocab_transform(slamgpu_ctx* ctx, int vocab_id, const uint8_t* desc, int n, int32_t* out_word,
                            double* out_weight);
/* Device-pointer form: enqueues on the context's stream, no synchronisation. */
SLAMGPU_API int slamgpu_vocab_transform_dev(slamgpu_ctx* ctx, int vocab_id, const uint8_t* d_desc, int n,
                                int32_t* d_out_word, double* d_out_weight);

/* transform(features, bow) for a batch of images (TemplatedVocabulary.h:1064-1120, BowVector.cpp:33-84):
 * image i owns descriptors [image_offset[i], image_offset[i+1]) (at most SLAMGPU_BOW_MAX_FEATURES each).
 * Its bag-of-words vector comes back as out_ids / out_vals [out_offset[i], out_offset[i+1]) ascending by
 * word id, exactly the std::map DBoW2 builds; out_ids / out_vals need image_offset[n_images] entries.
 * Words are found, sorted and counted on the device; the per-word value fold and the normalisation are
 * double-precision sums whose order DBoW2 fixes (feature order, ascending word id) and are done in that
 * order on the host from the counts. */
SLAMGPU_API int slamgpu_vocab_bow(slamgpu_ctx* ctx, int vocab_id, const uint8_t* desc, const int32_t* image_offset,
                      int n_images, int weighting, int norm, int32_t* out_ids, double* out_vals,
                      int32_t* out_offset);

/* L1Scoring::score(query, db[i]) for every vector of a database (LoopDetector::Impl::score_candidates,
 * src/LoopDetector.cpp:346-373): vectors are (ids ascending, values); the database is concatenated with
 * db_offset [n_db + 1].  out_score [n_db].  The terms of a score are summed in ascending word order as
 * DBoW2 does, so the doubles are bit-identical. */
SLAMGPU_API int slamgpu_vocab_score_l1(slamgpu_ctx* ctx, const int32_t* q_ids, const double* q_vals, int nq,
                           const int32_t* db_ids, const double* db_vals, const int32_t* db_offset, int n_db,
                           double* out_score);

/* Resident database of bag-of-words vectors -- LoopDetector::Impl::bows, one vector per key frame
 * (src/LoopDetector.cpp:346-373): append a key frame's vector when it is inserted, score a query against
 * every resident vector with one small upload and one download of n_vectors doubles.  Vector indices are
 * the append order.  db_id is created by the first append. */
SLAMGPU_API int slamgpu_bowdb_append(slamgpu_ctx* ctx, int db_id, const int32_t* ids, const double* vals, int n);
SLAMGPU_API int slamgpu_bowdb_count(slamgpu_ctx* ctx, int db_id, int* n_vectors);
SLAMGPU_API int slamgpu_bowdb_score_l1(slamgpu_ctx* ctx, int db_id, const int32_t* q_ids, const double* q_vals, int nq,
                           double* out_score);
SLAMGPU_API int slamgpu_bowdb_release(slamgpu_ctx* ctx, int db_id);

/* ------------------------------------------------------------------------------------------
 * PnP verification support: the data-parallel part of verify_pnp
 * (src/LoopDetector.cpp:176-229) on either side of cv::solvePnPRansac's model solver.
 *
 * A PROBLEM is one (query, candidate) pair: a run of 3-D / 2-D correspondences.  A MODEL is one pose
 * hypothesis of a problem as 12 doubles: R row-major (cv::Rodrigues of the solver's rvec), then t.
 * K4 = fx, fy, cx, cy as doubles (the reference converts its intrinsics to CV_64F,
 * src/LoopDetector.cpp:199-200); no distortion (cv::Mat() at :207).
 * What stays on the host: the minimal-set sampling and the EPnP solve of each RANSAC iteration (OpenCV's
 * RNG and solver; the sample sequence does not depend on the scores, so an adapter can draw every
 * iteration's model first and score them in one call) and the final refinement.
 * ---------------------------------------------------------------------------------------- */

/* Input assembly, src/LoopDetector.cpp:188-197, for a batch of pairs straight from
 * slamgpu_match_descriptors_batch_dev's output (same d_frame_offset / d_pairs / max_rows / d_match_train):
 *   d_row_pos [rows x 3]  position of the map point behind each descriptor row (match.point.position())
 *   d_row_xy  [rows x 2]  key point of each descriptor row (query.keypoint(i).pt)
 * Pair p's correspondences, ascending query row, land at [p*max_rows, p*max_rows + d_out_count[p]) of
 * d_out_object (x3) / d_out_image (x2) / d_out_query (the query row; may be NULL). */
SLAMGPU_API int slamgpu_pnp_assemble_batch_dev(slamgpu_ctx* ctx, const int32_t* d_match_train, const int32_t* d_frame_offset,
                                   int n_frames, const int32_t* d_pairs, int n_pairs, int max_rows,
                                   const float* d_row_pos, const float* d_row_xy, float* d_out_object,
                                   float* d_out_image, int32_t* d_out_query, int32_t* d_out_count);

/* Inlier count of every model: PnPRansacCallback::computeError (projectPoints, float squared error) +
 * RANSACPointSetRegistrator::findInliers (err <= (float)(reproj_error^2)) of cv::solvePnPRansac
 * (src/LoopDetector.cpp:203-214: reproj_error 4.0).  Problem p owns correspondences
 * [d_corr_begin[p], d_corr_begin[p] + d_corr_count[p]); model h belongs to problem d_hyp_problem[h].
 * d_out_count [n_hyp].  Enqueues on the context's stream, no synchronisation. */
SLAMGPU_API int slamgpu_pnp_score_dev(slamgpu_ctx* ctx, const float* d_object, const float* d_image,
                          const int32_t* d_corr_begin, const int32_t* d_corr_count, int n_problems,
                          const double* d_hyp_rt, const int32_t* d_hyp_problem, int n_hyp, const double* K4,
                          double reproj_error, int32_t* d_out_count);
/* Inlier mask of one chosen model per problem (d_sel_hyp[p] = model index or -1 for "none": zeros):
 * d_out_mask [same indexing as d_object rows], d_out_err (float squared errors; may be NULL),
 * d_out_inliers [n_problems] (may be NULL). */
SLAMGPU_API int slamgpu_pnp_mask_dev(slamgpu_ctx* ctx, const float* d_object, const float* d_image,
                         const int32_t* d_corr_begin, const int32_t* d_corr_count, int n_problems,
                         const double* d_hyp_rt, const int32_t* d_sel_hyp, const double* K4, double reproj_error,
                         uint8_t* d_out_mask, float* d_out_err, int32_t* d_out_inliers);

/* RANSACPointSetRegistrator::run's choice over models given in iteration order with their inlier counts
 * (host arithmetic, no device work): a model replaces the best when good > max(best, model_points - 1); the
 * iteration bound shrinks by RANSACUpdateNumIters after every replacement.  *out_best = kept model or -1,
 * *out_iters = models looked at (may be NULL).  cv::solvePnPRansac with SOLVEPNP_EPNP uses
 * model_points = 5, confidence 0.99, max_iters 200 at src/LoopDetector.cpp:203-214. */
SLAMGPU_API int slamgpu_ransac_select(const int32_t* good_count, int n_models, int n_points, int model_points,
                          double confidence, int max_iters, int* out_best, int* out_iters);

/* Host-buffer form of the whole scoring step for a batch of problems: upload, count every model, choose
 * per problem as RANSAC would, compute the kept model's mask, download.
 *   object [n_corr x 3], image [n_corr x 2], corr_offset [n_problems + 1]
 *   hyp_rt [n_hyp x 12], hyp_offset [n_problems + 1] (models of a problem in iteration order)
 *   out_count [n_hyp]; out_best [n_problems] (index within the problem, or -1); out_inliers [n_problems];
 *   out_mask [n_corr] (the `inliers` output of cv::solvePnPRansac as a mask; may be NULL). */
SLAMGPU_API int slamgpu_pnp_ransac_score(slamgpu_ctx* ctx, const float* object, const float* image, const int32_t* corr_offset,
                             int n_problems, const double* hyp_rt, const int32_t* hyp_offset, const double* K4,
                             double reproj_error, double confidence, int model_points, int max_iters,
                             int32_t* out_count, int32_t* out_best, int32_t* out_inliers, uint8_t* out_mask);

/* ------------------------------------------------------------------------------------------
 * Host restatements (no device): the float expressions of the projection path in the operation order the
 * kernels use, so that a host WITH Eigen can compare them bit for bit with the reference's own code
 * (MapMatcherGpu::self_check / audit, tools/eigen_order_check.cpp).  At u >= 1024 px one float ulp is
 * 1.2e-4 px, more than the 1e-4 px tolerance, so the order matters.
 *   slamgpu_host_project        Camera::project (src/Camera.cpp:25-32): out_uv [2]
 *   slamgpu_host_camera_center  Frame::camera_center (src/Frame.cpp:39-42): out_center [3]
 *   slamgpu_host_point_stats    MapPoint::avg_viewing_normal / observed_distance_range (src/MapPoint.cpp:24-45)
 *                               of one point: obs_center [n_obs x 3] in observation order, out_normal [3], out_range [2]
 * ---------------------------------------------------------------------------------------- */
SLAMGPU_API int slamgpu_host_project(const float* pose_colmajor16, const slamgpu_proj_params* params, const float* xyz,
                         float* out_uv);
SLAMGPU_API int slamgpu_host_camera_center(const float* pose_colmajor16, float* out_center);
SLAMGPU_API int slamgpu_host_point_stats(const float* xyz, int n_obs, const float* obs_center, float* out_normal,
                             float* out_range);

/* ------------------------------------------------------------------------------------------
 * Diagnostics
 * ---------------------------------------------------------------------------------------- */
/* Per-kernel device timing for roofline reports: while enabled, every launch of the two dominant
 * kernels is bracketed by a CUDA-event pair on the context's stream.  slamgpu_profile_read
 * synchronises, returns the summed duration and launch count since the last read, and resets.
 *   kernel  SLAMGPU_KERNEL_BF_SCAN (K1 brute-force scan), SLAMGPU_KERNEL_PROJ_MATCH (K4) or
 *           SLAMGPU_KERNEL_VOC_TRANSFORM (vocabulary descent), SLAMGPU_KERNEL_PNP_SCORE (RANSAC inlier scoring). */
#define SLAMGPU_KERNEL_BF_SCAN 0
#define SLAMGPU_KERNEL_PROJ_MATCH 1
#define SLAMGPU_KERNEL_VOC_TRANSFORM 2
#define SLAMGPU_KERNEL_PNP_SCORE 3
/* A projection call that arrives before the frame's KD-tree has been built (slamgpu_frame_upload builds it on a
 * host thread) does not wait for it: it runs on the tree-free candidate grid and COUNTS the points whose best
 * distance is shared by two keypoints -- the only place where the reference's KD visit order matters
 * (src/KDTree.cpp:52-82, src/MapMatcher.cpp:88).  A zero count makes the results final; otherwise the call waits
 * for the tree and repeats itself.  calls = speculative calls so far, repeats = how many had to be repeated.
 * SLAMGPU_KD_SPECULATE=0 switches the speculation off. */
SLAMGPU_API int slamgpu_speculation_count(slamgpu_ctx* ctx, int64_t* calls, int64_t* repeats);
/* With SLAMGPU_TRACE_BATCH=1 the host-buffer batch calls (slamgpu_match_descriptors_batch, slamgpu_kfdb_match_pairs)
 * add up host-clock microseconds per phase: out_us4 = {queueing the call's work, waiting for the slices while their
 * lists are copied out, the tail after the last slice, the whole call}; the sums reset on read. */
SLAMGPU_API int slamgpu_trace_read(slamgpu_ctx* ctx, double* out_us4, int64_t* out_calls);
SLAMGPU_API int slamgpu_profile_enable(slamgpu_ctx* ctx, int on);
SLAMGPU_API int slamgpu_profile_read(slamgpu_ctx* ctx, int kernel, double* total_ms, int64_t* launches);
/* POPC-pipe microbenchmark ("confirm with a POPC microbenchmark on the box"):
 * returns 32-bit POPC results per second over the whole device and the SM count. */
SLAMGPU_API int slamgpu_bench_popc(slamgpu_ctx* ctx, double* popc_per_s, int* n_sm, double* sm_mhz_est);

#ifdef __cplusplus
}
#endif
#endif /* SLAMGPU_H */
