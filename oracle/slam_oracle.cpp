// slam_oracle.cpp -- CPU ORACLE for the descriptor-matching hot path.  TEST INFRASTRUCTURE ONLY.
//
// This file is the checker, not the product.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load it.  Nothing under
// racing-slam_b200/ links, imports or executes it; the product path fails loudly when
// libslamgpu.so (the CUDA library) is missing.
//
// What it restates (all citations relative to /root/reference):
//   * brute-force kNN / cross-check semantics of cv::BFMatcher as the reference calls it
//     (src/MapMatcher.cpp:129-163, src/features/FeatureExtractor.cpp:15-30).  The arithmetic
//     lives in OpenCV (vcpkg port `opencv`, baseline 4b6c50d9..., vcpkg-configuration.json:4),
//     which is not vendored; the published algorithm is restated here and PINNED against
//     cv2 4.13.0 outputs committed under tests/golden/ (oracle/gen_golden.py).
//   * projection-guided search: src/MapMatcher.cpp:12-18,25-43,45-98,117-127,165-175;
//     src/Camera.cpp:25-37; src/KDTree.cpp:8-82; src/Frame.cpp:39-42,138-151;
//     src/MapPoint.cpp:24-45.
//     The KD-tree part is PINNED against the reference's own src/KDTree.cpp compiled from
//     where it lies (oracle/Makefile -> oracle/_ref/).  The floating-point part (Eigen
//     expression order of src/Camera.cpp:27, src/Frame.cpp:41, src/MapPoint.cpp:27-32) is
//     PARITY UNPINNED: Eigen is absent from this image, the order below is derived from
//     Eigen's documented coefficient-based product / redux evaluation and cannot be run
//     against the real thing.  The north_star's 1e-4 px tolerance covers exactly this.
//
// Build: g++ -O2 -std=c++17 -ffp-contract=off -fopenmp -fPIC -shared (no -march: the reference
// builds for baseline x86-64, so no FMA contraction -- CMakeLists.txt:1-7).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#define ORC_API extern "C" __attribute__((visibility("default")))

namespace {

// ---------------------------------------------------------------------------------------------
// B4: cv::norm(a, b, NORM_HAMMING) on two 1x32 CV_8U rows (src/MapMatcher.cpp:87)
// = popcount of the byte-wise XOR.
// ---------------------------------------------------------------------------------------------
inline int hamming256(const uint8_t* a, const uint8_t* b)
{
    uint64_t x[4], y[4];
    std::memcpy(x, a, 32);
    std::memcpy(y, b, 32);
    return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) +
           __builtin_popcountll(x[2] ^ y[2]) + __builtin_popcountll(x[3] ^ y[3]);
}

struct Top2 {
    int d0 = std::numeric_limits<int>::max(), i0 = -1;
    int d1 = std::numeric_limits<int>::max(), i1 = -1;
    // B3: kNN = first k of a stable sort by (distance, trainIdx).  Train indices arrive in
    // ascending order, so a strict '<' keeps the lowest index on equal distances.
    inline void push(int d, int i)
    {
        if (d < d0) {
            d1 = d0; i1 = i0; d0 = d; i0 = i;
        } else if (d < d1) {
            d1 = d; i1 = i;
        }
    }
};

// ---------------------------------------------------------------------------------------------
// P5: KDTree2D (src/KDTree.cpp:8-82).  The tree shape is implicit: the node of range [s,e) is
// indices[(s+e)/2] after std::nth_element on axis depth%2, so the tree IS the permutation.
// ---------------------------------------------------------------------------------------------
void kd_build_rec(const float* xy, std::vector<size_t>& indices, int depth, int start, int end)
{
    if (start >= end) return;                                   // src/KDTree.cpp:21-22
    const int axis = depth % 2;                                 // :24
    const int mid = (start + end) / 2;                          // :25
    std::nth_element(indices.begin() + start, indices.begin() + mid, indices.begin() + end,
                     [&](size_t a, size_t b) {                  // :28-37, same comparator
                         if (axis == 0) return xy[2 * a] < xy[2 * b];
                         return xy[2 * a + 1] < xy[2 * b + 1];
                     });
    kd_build_rec(xy, indices, depth + 1, start, mid);           // :40
    kd_build_rec(xy, indices, depth + 1, mid + 1, end);         // :41
}

struct KdView {
    const float* xy;
    const uint32_t* perm;
    int n;
};

// src/KDTree.cpp:52-82 on the implicit tree.  `visit` receives keypoint indices in the
// reference's push_back order (node, near subtree, far subtree).
template <class F>
void kd_radius_rec(const KdView& kd, int start, int end, float tx, float ty, float r2, int depth, F& visit)
{
    if (start >= end) return;                                   // :58-59 (null node)
    const int mid = (start + end) / 2;
    const uint32_t idx = kd.perm[mid];
    const float dx = kd.xy[2 * idx] - tx;                       // :61
    const float dy = kd.xy[2 * idx + 1] - ty;                   // :62
    const float d2 = dx * dx + dy * dy;                         // :63 (no FMA: -ffp-contract=off)
    if (d2 <= r2) visit(static_cast<int>(idx));                 // :65-67
    const int axis = depth % 2;                                 // :69
    const float delta = (axis == 0) ? dx : dy;                  // :70
    // :73-76  delta > 0 -> near = left
    const bool left_near = delta > 0;
    const int ns = left_near ? start : mid + 1, ne = left_near ? mid : end;
    const int fs = left_near ? mid + 1 : start, fe = left_near ? end : mid;
    kd_radius_rec(kd, ns, ne, tx, ty, r2, depth + 1, visit);    // :76
    if (delta * delta <= r2) {                                  // :79
        kd_radius_rec(kd, fs, fe, tx, ty, r2, depth + 1, visit);
    }
}

// ---------------------------------------------------------------------------------------------
// Eigen expression order (derived; see header).  A fixed 3-term coefficient-based product /
// redux evaluates a0 + (a1 + a2).
// ---------------------------------------------------------------------------------------------
inline float sum3(float a0, float a1, float a2) { return a0 + (a1 + a2); }

// pose is an Eigen::Matrix4f: COLUMN-major, element (r,c) at pose[c*4 + r].
inline float P(const float* pose, int r, int c) { return pose[c * 4 + r]; }

// src/Frame.cpp:39-42   C = -R^T * t
void camera_center(const float* pose, float C[3])
{
    for (int i = 0; i < 3; i++) {
        C[i] = sum3(-P(pose, 0, i) * P(pose, 0, 3), -P(pose, 1, i) * P(pose, 1, 3), -P(pose, 2, i) * P(pose, 2, 3));
    }
}

// M = K * pose[0:3,0:4] (src/Camera.cpp:27, first product), K = [[fx,0,cx],[0,fy,cy],[0,0,1]].
void projection_matrix(const float* pose, float fx, float fy, float cx, float cy, float M[12] /* row-major 3x4 */)
{
    const float K[3][3] = {{fx, 0.0F, cx}, {0.0F, fy, cy}, {0.0F, 0.0F, 1.0F}};
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 4; j++)
            M[i * 4 + j] = sum3(K[i][0] * P(pose, 0, j), K[i][1] * P(pose, 1, j), K[i][2] * P(pose, 2, j));
}

// src/Camera.cpp:25-32.  Eigen's homogeneous product: M[:, :3] * X, then += M[:, 3].
inline void project_with(const float M[12], const float X[3], float uv[2])
{
    float h[3];
    for (int i = 0; i < 3; i++) {
        h[i] = sum3(M[i * 4 + 0] * X[0], M[i * 4 + 1] * X[1], M[i * 4 + 2] * X[2]);
        h[i] = h[i] + M[i * 4 + 3];
    }
    if (h[2] < 0) {                                             // :28-30
        uv[0] = -1.0F; uv[1] = -1.0F;
        return;
    }
    uv[0] = h[0] / h[2];                                        // :31 hnormalized(): true division
    uv[1] = h[1] / h[2];
}

// Eigen normalized(): z = squaredNorm(); z > 0 ? v / sqrt(z) : v
inline void normalized3(const float v[3], float out[3])
{
    const float z = sum3(v[0] * v[0], v[1] * v[1], v[2] * v[2]);
    if (z > 0.0F) {
        const float n = std::sqrt(z);
        out[0] = v[0] / n; out[1] = v[1] / n; out[2] = v[2] / n;
    } else {
        out[0] = v[0]; out[1] = v[1]; out[2] = v[2];
    }
}
inline float norm3(const float v[3]) { return std::sqrt(sum3(v[0] * v[0], v[1] * v[1], v[2] * v[2])); }

} // namespace

// =============================================================================================
// exported API (ctypes)
// =============================================================================================

ORC_API int orc_hamming256(const uint8_t* a, const uint8_t* b) { return hamming256(a, b); }

// B3 kNN-2.  idx/dist are nq x 2; missing neighbours (nt < 2) are -1.
ORC_API void orc_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < nq; i++) {
        Top2 best;
        for (int j = 0; j < nt; j++) best.push(hamming256(q + 32 * (size_t)i, t + 32 * (size_t)j), j);
        idx[2 * i] = best.i0;  dist[2 * i] = best.i0 < 0 ? -1 : best.d0;
        idx[2 * i + 1] = best.i1;  dist[2 * i + 1] = best.i1 < 0 ? -1 : best.d1;
    }
}

// B3 cross-check (k = 1): keep (q,t) iff t = argmin_t d(q,.) and q = argmin_q d(.,t), lowest
// index winning ties in both directions.  idx[q] = t or -1.
ORC_API void orc_crosscheck(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist)
{
    std::vector<int> row_best(nq, -1), row_d(nq, 0), col_best(nt, -1);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < nq; i++) {
        int bd = std::numeric_limits<int>::max(), bi = -1;
        for (int j = 0; j < nt; j++) {
            const int d = hamming256(q + 32 * (size_t)i, t + 32 * (size_t)j);
            if (d < bd) { bd = d; bi = j; }
        }
        row_best[i] = bi; row_d[i] = bd;
    }
#pragma omp parallel for schedule(static)
    for (int j = 0; j < nt; j++) {
        int bd = std::numeric_limits<int>::max(), bi = -1;
        for (int i = 0; i < nq; i++) {
            const int d = hamming256(q + 32 * (size_t)i, t + 32 * (size_t)j);
            if (d < bd) { bd = d; bi = i; }
        }
        col_best[j] = bi;
    }
    for (int i = 0; i < nq; i++) {
        const int tj = row_best[i];
        const bool ok = tj >= 0 && col_best[tj] == i;
        idx[i] = ok ? tj : -1;
        dist[i] = ok ? row_d[i] : -1;
    }
}

// B1: filter of MapMatcher::match_descriptors (src/MapMatcher.cpp:145-161) applied to kNN rows.
// Emits (query, train) in ascending query order; returns the count.
ORC_API int orc_match_descriptors(const uint8_t* q, int nq, const uint8_t* t, int nt, float max_dist, float ratio,
                                  int32_t* out_query, int32_t* out_train)
{
    if (nt == 0 || nq == 0) return 0;                           // :139-141
    std::vector<int32_t> idx(2 * (size_t)nq), dist(2 * (size_t)nq);
    orc_knn2(q, nq, t, nt, idx.data(), dist.data());
    const int k = nt >= 2 ? 2 : 1;                              // :145
    int count = 0;
    for (int i = 0; i < nq; i++) {
        const float d0 = static_cast<float>(dist[2 * i]);
        if (d0 > max_dist) continue;                            // :152 (non-strict: d0 == max passes)
        if (k > 1) {
            const float d1 = static_cast<float>(dist[2 * i + 1]);
            if (d0 > ratio * d1) continue;                      // :156 float multiply
        }
        out_query[count] = i;
        out_train[count] = idx[2 * i];
        count++;
    }
    return count;
}

// B2: BaseFeatureExtractor::match_features (src/features/FeatureExtractor.cpp:15-30).
// query = cur, train = prev; keep distance < max_distance (strict).  FeatureMatch(train, query).
ORC_API int orc_match_features(const uint8_t* prev, int nprev, const uint8_t* cur, int ncur, float max_dist,
                               int32_t* out_train, int32_t* out_query)
{
    std::vector<int32_t> idx(ncur > 0 ? ncur : 1), dist(ncur > 0 ? ncur : 1);
    orc_crosscheck(cur, ncur, prev, nprev, idx.data(), dist.data());
    int count = 0;
    for (int i = 0; i < ncur; i++) {
        if (idx[i] < 0) continue;                               // match.size() > 0
        if (!(static_cast<float>(dist[i]) < max_dist)) continue; // :25
        out_train[count] = idx[i];
        out_query[count] = i;
        count++;
    }
    return count;
}

// P5 build: perm[m] = keypoint index stored at tree position m.
ORC_API void orc_kd_build(const float* xy, int n, uint32_t* perm)
{
    std::vector<size_t> indices(n);
    for (int i = 0; i < n; i++) indices[i] = i;                 // src/KDTree.cpp:11-14
    kd_build_rec(xy, indices, 0, 0, n);
    for (int i = 0; i < n; i++) perm[i] = static_cast<uint32_t>(indices[i]);
}

// P5 search: keypoint indices in reference visit order; returns the count (may exceed cap;
// only the first cap are written).
ORC_API int orc_kd_radius_search(const float* xy, const uint32_t* perm, int n, float tx, float ty, float radius,
                                 int32_t* out, int cap)
{
    KdView kd{xy, perm, n};
    int count = 0;
    auto visit = [&](int idx) {
        if (count < cap) out[count] = idx;
        count++;
    };
    kd_radius_rec(kd, 0, n, tx, ty, radius * radius, 0, visit); // src/KDTree.cpp:48
    return count;
}

ORC_API void orc_camera_center(const float* pose_colmajor, float* C) { camera_center(pose_colmajor, C); }

ORC_API void orc_projection_matrix(const float* pose_colmajor, float fx, float fy, float cx, float cy, float* M12)
{
    projection_matrix(pose_colmajor, fx, fy, cx, cy, M12);
}

// P1 for n points; uv is n x 2.
ORC_API void orc_project(const float* pose_colmajor, float fx, float fy, float cx, float cy, const float* pos, int n,
                         float* uv)
{
    float M[12];
    projection_matrix(pose_colmajor, fx, fy, cx, cy, M);
    for (int i = 0; i < n; i++) project_with(M, pos + 3 * (size_t)i, uv + 2 * (size_t)i);
}

// src/MapPoint.cpp:24-45 for n points.  Observations are given flattened in the HOST's
// unordered_map iteration order: obs_offset[n+1] into obs_center[sum_obs x 3] (the observing
// key frames' camera centres, src/Frame.cpp:39-42).
ORC_API void orc_point_stats(const float* pos, int n, const int32_t* obs_offset, const float* obs_center,
                             float* normal /* n x 3 */, float* range /* n x 2: nearest, furthest */)
{
    for (int i = 0; i < n; i++) {
        const float* X = pos + 3 * (size_t)i;
        float acc[3] = {0.0F, 0.0F, 0.0F};                      // :26
        float nearest = std::numeric_limits<float>::max();      // :37
        float furthest = 0.0F;                                  // :38
        for (int o = obs_offset[i]; o < obs_offset[i + 1]; o++) {
            const float* C = obs_center + 3 * (size_t)o;
            const float d[3] = {X[0] - C[0], X[1] - C[1], X[2] - C[2]};
            float v[3], vv[3];
            normalized3(d, v);                                  // :29
            normalized3(v, vv);                                 // :30 (normalised twice)
            acc[0] = acc[0] + vv[0]; acc[1] = acc[1] + vv[1]; acc[2] = acc[2] + vv[2];
            const float dist = norm3(d);                        // :40
            nearest = std::min(nearest, dist);                  // :41
            furthest = std::max(furthest, dist);                // :42
        }
        normalized3(acc, normal + 3 * (size_t)i);               // :32
        range[2 * i] = nearest;
        range[2 * i + 1] = furthest;
    }
}

struct OrcProjParams {
    float fx, fy, cx, cy;
    int32_t width, height;
    float radius;        // SEARCH_RADIUS            src/MapMatcher.cpp:12
    float max_dist;      // m_max_descriptor_distance
    float min_cos;       // MIN_VIEWING_ANGLE_COSINE :13
    float nearer;        // MAX_NEARER_RATIO         :16
    float further;       // MAX_FURTHER_RATIO        :17
    int32_t replace;     // match_for_fuse -> 1      :124
};

// P0-P8.  Points are visited in `order` (n_order entries indexing the point set) or, when
// order == nullptr, 0..n_points-1.  skip[i] != 0 drops point i up front (P0 `is_matched(point)`
// and P9 `!is_observed_by(kf)`; negative entries of `order` model match_for_fuse's nullptr).
// out_point[kp] = point-set index or -1; out_dist[kp] = Hamming distance or -1.
// Optional dbg_uv (n_visit x 2), dbg_kp / dbg_d (n_visit): per visited point, projection, chosen
// keypoint and distance (-1 when the point proposes nothing).  stats[0] = projections within
// 1e-4 px of an image border, stats[1] = (point, keypoint) pairs within 1e-4 px of the radius.
ORC_API void orc_match_projection(int nkp, const float* kp_xy, const uint8_t* kp_desc, const uint32_t* kd_perm,
                                  const uint8_t* kp_matched, int n_points, const float* pos, const float* normal,
                                  const float* range, const int32_t* obs_offset, const uint8_t* obs_desc,
                                  const uint8_t* skip, const int32_t* order, int n_order,
                                  const float* pose_colmajor, const OrcProjParams* prm, int32_t* out_point,
                                  int32_t* out_dist, float* dbg_uv, int32_t* dbg_kp, int32_t* dbg_d, int64_t* stats)
{
    // src/MapMatcher.cpp:25-32 empty_proposals
    std::vector<float> prop_dist(nkp > 0 ? nkp : 1, prm->max_dist);
    std::vector<int32_t> prop_point(nkp > 0 ? nkp : 1, -1);
    float M[12], C[3];
    projection_matrix(pose_colmajor, prm->fx, prm->fy, prm->cx, prm->cy, M);
    camera_center(pose_colmajor, C);
    KdView kd{kp_xy, kd_perm, nkp};
    const float r2 = prm->radius * prm->radius;                 // src/KDTree.cpp:48
    const int n_visit = order ? n_order : n_points;
    int64_t near_border = 0, near_radius = 0;

    for (int v = 0; v < n_visit; v++) {
        const int pi = order ? order[v] : v;
        if (dbg_uv) { dbg_uv[2 * v] = NAN; dbg_uv[2 * v + 1] = NAN; }
        if (dbg_kp) dbg_kp[v] = -1;
        if (dbg_d) dbg_d[v] = -1;
        if (pi < 0) continue;                                   // :119-121 nullptr
        if (skip && skip[pi]) continue;                         // :53 / :169
        const float* X = pos + 3 * (size_t)pi;
        float uv[2];
        project_with(M, X, uv);                                 // :57
        if (dbg_uv) { dbg_uv[2 * v] = uv[0]; dbg_uv[2 * v + 1] = uv[1]; }
        // src/Camera.cpp:34-37
        if (!(uv[0] >= 0 && uv[0] < (float)prm->width && uv[1] >= 0 && uv[1] < (float)prm->height)) continue;
        if (std::fabs(uv[0]) < 1e-4F || std::fabs(uv[0] - (float)prm->width) < 1e-4F || std::fabs(uv[1]) < 1e-4F ||
            std::fabs(uv[1] - (float)prm->height) < 1e-4F)
            near_border++;

        const float ray[3] = {X[0] - C[0], X[1] - C[1], X[2] - C[2]};                     // :62
        float rn[3];
        normalized3(ray, rn);
        const float* nrm = normal + 3 * (size_t)pi;                                       // :63
        if (sum3(nrm[0] * rn[0], nrm[1] * rn[1], nrm[2] * rn[2]) < prm->min_cos) continue; // :64
        const float nearest = range[2 * pi], furthest = range[2 * pi + 1];                // :69
        const float distance = norm3(ray);                                                // :70
        if (distance < nearest / prm->nearer || distance > furthest * prm->further) continue; // :71

        if (stats) {   // diagnostic only: every keypoint whose distance is within 1e-4 px of the radius
            for (int k = 0; k < nkp; k++) {
                const float dx = kp_xy[2 * k] - uv[0], dy = kp_xy[2 * k + 1] - uv[1];
                if (std::fabs(std::sqrt(dx * dx + dy * dy) - prm->radius) < 1e-4F) near_radius++;
            }
        }

        size_t best_index = 0;                                  // :77
        float best = prm->max_dist;                             // :78
        auto visit = [&](int idx) {
            if (!prm->replace && kp_matched && kp_matched[idx]) return;                   // :81
            const uint8_t* d = kp_desc + 32 * (size_t)idx;
            for (int o = obs_offset[pi]; o < obs_offset[pi + 1]; o++) {                   // :86
                const double orb = (double)hamming256(d, obs_desc + 32 * (size_t)o);      // :87
                if (orb < best) {                                                         // :88 strict
                    best = (float)orb;
                    best_index = idx;
                }
            }
        };
        kd_radius_rec(kd, 0, nkp, uv[0], uv[1], r2, 0, visit);  // :75

        if (nkp > 0 && best < prop_dist[best_index]) {          // :95
            prop_dist[best_index] = best;
            prop_point[best_index] = pi;
            // dbg_* record the point's own choice, whether or not it survives later points
        }
        if (best < prm->max_dist) {
            if (dbg_kp) dbg_kp[v] = (int32_t)best_index;
            if (dbg_d) dbg_d[v] = (int32_t)best;
        }
    }
    for (int i = 0; i < nkp; i++) {                             // :34-43 accepted_matches
        out_point[i] = prop_point[i];
        out_dist[i] = prop_point[i] >= 0 ? (int32_t)prop_dist[i] : -1;
    }
    if (stats) { stats[0] = near_border; stats[1] = near_radius; }
}

ORC_API int orc_abi_version(void) { return 1; }

// =============================================================================================
// Bag-of-words transform and scoring: what LoopDetector does before the brute-force
// verification -- bow_of (src/LoopDetector.cpp:85-92) -> DBoW2 TemplatedVocabulary::transform, and
// vocabulary->score (:369) -> L1Scoring::score.  DBoW2 is vendored in the reference
// (thirdparty/DBoW2); this restatement is PINNED against that code compiled unmodified into
// oracle/_ref/libref_dbow2.so (oracle/ref_dbow2_wrap.cpp), fixtures under tests/golden/voc_*.npz.
//
// The vocabulary is a tree given as arrays: child_offset [n_nodes + 1] / child_ids in DBoW2's own
// child order, desc [n_nodes x 32], weight [n_nodes] (word weight, e.g. idf), word_id [n_nodes]
// (-1 for inner nodes).  Node 0 is the root.
// =============================================================================================

// thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1257: descend from the root, at each level take the
// FIRST child with the strictly smallest FORB::distance (FORB.cpp:81-101 = 256-bit Hamming).
ORC_API void orc_voc_transform_each(const int32_t* child_offset, const int32_t* child_ids, const uint8_t* node_desc,
                                    const double* weight, const int32_t* word_id, const uint8_t* desc, int n,
                                    int32_t* out_word, double* out_weight)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; i++) {
        const uint8_t* f = desc + 32 * (size_t)i;
        int node = 0;
        do {
            const int b = child_offset[node], e = child_offset[node + 1];
            int best = child_ids[b];
            int best_d = hamming256(f, node_desc + 32 * (size_t)best);
            for (int c = b + 1; c < e; c++) {
                const int id = child_ids[c];
                const int d = hamming256(f, node_desc + 32 * (size_t)id);
                if (d < best_d) {
                    best_d = d;
                    best = id;
                }
            }
            node = best;
        } while (child_offset[node + 1] > child_offset[node]);
        out_word[i] = word_id[node];
        out_weight[i] = weight[node];
    }
}

// TemplatedVocabulary.h:1064-1120 + BowVector.cpp:33-84.  weighting: 0 TF_IDF, 1 TF, 2 IDF, 3 BINARY;
// norm: 0 = none (scoring without normalisation), 1 = L1, 2 = L2 (ScoringObject mustNormalize).
// Words arrive per feature in feature order; values accumulate in that order; the map is ascending by
// word id; the norm is summed in ascending order.  Returns the number of entries.
ORC_API int orc_voc_bow(const int32_t* word, const double* w, int n, int weighting, int norm, int32_t* out_ids,
                        double* out_vals)
{
    std::vector<std::pair<int32_t, int>> order(n);        // (word, feature index): stable order by word
    int m = 0;
    for (int i = 0; i < n; i++)
        if (w[i] > 0) order[m++] = {word[i], i};          // "not stopped"
    order.resize(m);
    std::stable_sort(order.begin(), order.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    int cnt = 0;
    for (int i = 0; i < m;) {
        int j = i;
        double v = w[order[i].second];                    // insert(id, w)
        for (j = i + 1; j < m && order[j].first == order[i].first; j++)
            if (weighting == 0 || weighting == 1) v += w[order[j].second];   // addWeight; addIfNotExist keeps the first
        out_ids[cnt] = order[i].first;
        out_vals[cnt] = v;
        cnt++;
        i = j;
    }
    const bool must = norm != 0;
    if ((weighting == 0 || weighting == 1) && cnt > 0 && !must) {
        const double nd = cnt;
        for (int i = 0; i < cnt; i++) out_vals[i] /= nd;
    }
    if (must) {
        double s = 0.0;
        if (norm == 1) {
            for (int i = 0; i < cnt; i++) s += std::fabs(out_vals[i]);
        } else {
            for (int i = 0; i < cnt; i++) s += out_vals[i] * out_vals[i];
            s = std::sqrt(s);
        }
        if (s > 0.0)
            for (int i = 0; i < cnt; i++) out_vals[i] /= s;
    }
    return cnt;
}

// ScoringObject.cpp:23-68 (L1Scoring::score): merge of two ascending bow vectors.
ORC_API double orc_voc_score_l1(const int32_t* ids_a, const double* vals_a, int na, const int32_t* ids_b,
                                const double* vals_b, int nb)
{
    int i = 0, j = 0;
    double score = 0;
    while (i < na && j < nb) {
        if (ids_a[i] == ids_b[j]) {
            const double vi = vals_a[i], wi = vals_b[j];
            score += std::fabs(vi - wi) - std::fabs(vi) - std::fabs(wi);
            i++;
            j++;
        } else if (ids_a[i] < ids_b[j]) {
            i = (int)(std::lower_bound(ids_a + i, ids_a + na, ids_b[j]) - ids_a);
        } else {
            j = (int)(std::lower_bound(ids_b + j, ids_b + nb, ids_a[i]) - ids_b);
        }
    }
    return -score / 2.0;
}

// ---------------------------------------------------------------------------------------------
// the data-parallel part of verify_pnp (src/LoopDetector.cpp:176-229) around
// cv::solvePnPRansac (called at :203-214 with 200 iterations, reprojection error 4.0, confidence 0.99,
// SOLVEPNP_EPNP; also src/Initialization.cpp:211-222).  The arithmetic lives in OpenCV (vcpkg port
// `opencv`, baseline 4b6c50d9..., not vendored).  Restated from its published source:
//   * projection: calib3d cvProjectPoints2Internal without distortion -- doubles, rotation applied as
//     R[0]*X + R[1]*Y + R[2]*Z + t[0] left to right, z = z ? 1./z : 1, x *= z, u = x*fx + cx, stored as
//     float.  PINNED: bit-identical to cv2 4.13.0 projectPoints on the fixtures of oracle/gen_golden_pnp.py
//     (tests/golden/pnp_*.npz; R = cv2.Rodrigues(rvec), which projectPoints applies itself).
//   * error: PnPRansacCallback::computeError (solvepnp.cpp) -- err = norm(Matx21f(ip - pp), NORM_L2SQR)
//     = float dx*dx + dy*dy with dx, dy float differences.
//   * inliers: RANSACPointSetRegistrator::findInliers (ptsetreg.cpp) -- float t = (float)(thr*thr);
//     inlier iff err <= t.
//   * model choice: RANSACPointSetRegistrator::run -- a model replaces the best when
//     goodCount > max(maxGoodCount, modelPoints - 1); niters = RANSACUpdateNumIters(...) after each
//     replacement; models past niters are never scored.
// The last three are PARITY UNPINNED (cv2 exposes neither the per-model errors nor the iteration
// count); only the projection they consume is pinned.
// ---------------------------------------------------------------------------------------------
namespace {
inline void pnp_project_one(const double* rt, const double* K4, const float* X3, float* uv)
{
    const double X = X3[0], Y = X3[1], Z = X3[2];
    double x = rt[0] * X + rt[1] * Y + rt[2] * Z + rt[9];
    double y = rt[3] * X + rt[4] * Y + rt[5] * Z + rt[10];
    double z = rt[6] * X + rt[7] * Y + rt[8] * Z + rt[11];
    z = z ? 1. / z : 1;
    x *= z;
    y *= z;
    uv[0] = (float)(x * K4[0] + K4[2]);
    uv[1] = (float)(y * K4[1] + K4[3]);
}
} // namespace

// rt = R row-major [9] then t [3]; K4 = fx, fy, cx, cy.
ORC_API void orc_pnp_project(const double* rt, const double* K4, const float* object, int n, float* out_uv)
{
    for (int i = 0; i < n; i++) pnp_project_one(rt, K4, object + 3 * i, out_uv + 2 * i);
}

// -> inlier count; out_err [n] (may be null), out_mask [n] (may be null)
ORC_API int orc_pnp_score(const double* rt, const double* K4, const float* object, const float* image, int n,
                          double reproj_error, float* out_err, uint8_t* out_mask)
{
    const float t = (float)(reproj_error * reproj_error);
    int nz = 0;
    for (int i = 0; i < n; i++) {
        float uv[2];
        pnp_project_one(rt, K4, object + 3 * i, uv);
        const float dx = image[2 * i] - uv[0], dy = image[2 * i + 1] - uv[1];
        float s = 0;
        s += dx * dx;
        s += dy * dy;
        const int f = s <= t;
        if (out_err) out_err[i] = s;
        if (out_mask) out_mask[i] = (uint8_t)f;
        nz += f;
    }
    return nz;
}

// ptsetreg.cpp RANSACUpdateNumIters
ORC_API int orc_ransac_update_num_iters(double p, double ep, int model_points, int max_iters)
{
    p = std::max(p, 0.);
    p = std::min(p, 1.);
    ep = std::max(ep, 0.);
    ep = std::min(ep, 1.);
    double num = std::max(1. - p, std::numeric_limits<double>::min());
    double denom = 1. - std::pow(1. - ep, model_points);
    if (denom < std::numeric_limits<double>::min()) return 0;
    num = std::log(num);
    denom = std::log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)std::lrint(num / denom);
}

// RANSACPointSetRegistrator::run over models given in iteration order with their inlier counts:
// -> index of the kept model or -1; *out_iters = models the loop looked at.
ORC_API int orc_ransac_select(const int32_t* good_count, int n_models, int n_points, int model_points,
                              double confidence, int max_iters, int* out_iters)
{
    int niters = std::max(max_iters, 1), best = -1, max_good = 0, iter = 0;
    for (; iter < niters && iter < n_models; iter++) {
        const int good = good_count[iter];
        if (good > std::max(max_good, model_points - 1)) {
            best = iter;
            max_good = good;
            niters = orc_ransac_update_num_iters(confidence, (double)(n_points - good) / n_points, model_points, niters);
        }
    }
    if (out_iters) *out_iters = iter;
    return best;
}

// verify_pnp's input assembly, src/LoopDetector.cpp:188-197: correspondences (ascending query row, as
// match_descriptors emits them) -> object_points / image_points.  match_train [nq] = train row or -1.
ORC_API int orc_pnp_assemble(const int32_t* match_train, int nq, const float* train_pos, const float* query_xy,
                             float* out_object, float* out_image, int32_t* out_query)
{
    int n = 0;
    for (int q = 0; q < nq; q++) {
        const int t = match_train[q];
        if (t < 0) continue;
        std::memcpy(out_object + 3 * n, train_pos + 3 * t, 12);
        std::memcpy(out_image + 2 * n, query_xy + 2 * q, 8);
        if (out_query) out_query[n] = q;
        n++;
    }
    return n;
}

// Counts of many models over ragged problems (the CPU leg of bench.py's N3 measurement): model h belongs to
// problem hyp_problem[h]; OpenMP over models when threads > 1.
ORC_API void orc_pnp_score_batch(const float* object, const float* image, const int32_t* corr_begin,
                                 const int32_t* corr_count, const double* hyp_rt, const int32_t* hyp_problem, int n_hyp,
                                 const double* K4, double reproj_error, int threads, int32_t* out_count)
{
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads > 0 ? threads : 1)
    for (int h = 0; h < n_hyp; h++) {
        const int p = hyp_problem[h];
        out_count[h] = orc_pnp_score(hyp_rt + 12 * (size_t)h, K4, object + 3 * (size_t)corr_begin[p],
                                     image + 2 * (size_t)corr_begin[p], corr_count[p], reproj_error, nullptr, nullptr);
    }
}
