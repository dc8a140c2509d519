// MapMatcherGpu.cpp -- see MapMatcherGpu.h.  Citations are relative to the reference tree.
#include "MapMatcherGpu.h"

#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <unordered_map>

#include "Camera.h"
#include "Frame.h"
#include "Map.h"
#include "MapPoint.h"

namespace slam {

namespace {

void check(int rc, const char* what)
{
    // The reference has no error channel (an empty vector means "no matches"); a GPU failure is not
    // "no matches", so it is raised instead of being swallowed.
    if (rc != SLAMGPU_OK) {
        throw std::runtime_error(std::string("libslamgpu: ") + what + ": " + slamgpu_last_error());
    }
}

// cv::Mat N x 32 CV_8U -> contiguous bytes (descriptor rows are contiguous; the Mat may not be)
void append_rows(const cv::Mat& m, std::vector<uint8_t>& out)
{
    for (int r = 0; r < m.rows; r++) {
        const uint8_t* p = m.ptr<uint8_t>(r);
        out.insert(out.end(), p, p + 32);
    }
}

constexpr int FRAME_SLOT = 0;
constexpr int POINT_SLOT = 0;

} // namespace

MapMatcherGpu::MapMatcherGpu(const Camera& camera, float max_descriptor_distance, cv::NormTypes norm_type, int device)
    : m_camera(camera), m_max_descriptor_distance(max_descriptor_distance)
{
    if (norm_type != cv::NORM_HAMMING) {
        throw std::invalid_argument("MapMatcherGpu implements NORM_HAMMING (ORB) matching only");
    }
    check(slamgpu_create(device, &m_ctx), "slamgpu_create");
}

MapMatcherGpu::~MapMatcherGpu() { slamgpu_destroy(m_ctx); }

std::vector<MapPointMatch> MapMatcherGpu::match_map(const Frame& frame, Map& map) const
{
    std::vector<MapPoint*> visit;
    for (auto& point : map) {                                   // src/MapMatcher.cpp:168, map order
        visit.push_back(&point);
    }
    return project(frame, visit, nullptr, false);
}

std::vector<MapPointMatch> MapMatcherGpu::match_key_frame(const Frame& frame, Map& map, KeyFrame* key_frame) const
{
    std::vector<MapPoint*> visit;
    for (auto& point : map) {
        visit.push_back(&point);
    }
    return project(frame, visit, key_frame, false);
}

std::vector<MapPointMatch> MapMatcherGpu::match_for_fuse(const Frame& frame, const std::vector<MapPoint*>& points) const
{
    return project(frame, points, nullptr, true);               // nullptr entries are kept: :119-121
}

// Frame state: keypoints, descriptors, is_matched(size_t) (src/Frame.cpp:143-146) -> resident frame slot.
int MapMatcherGpu::upload_frame(const Frame& frame) const
{
    const auto& features = frame.features();
    const int nkp = static_cast<int>(features.keypoints.size());
    if (nkp == 0) {
        return 0;
    }
    std::vector<float> xy(2 * static_cast<size_t>(nkp));
    std::vector<uint8_t> matched(static_cast<size_t>(nkp));
    for (int i = 0; i < nkp; i++) {
        xy[2 * i] = features.keypoints[i].pt.x;                 // the KD-tree is built over these, src/Frame.cpp:11-15
        xy[2 * i + 1] = features.keypoints[i].pt.y;
        matched[i] = frame.is_matched(static_cast<size_t>(i)) ? 1 : 0;
    }
    std::vector<uint8_t> desc;
    desc.reserve(32 * static_cast<size_t>(nkp));
    append_rows(features.descriptors, desc);
    // kd_perm = nullptr: the library rebuilds the tree with the same std::nth_element call
    // (Frame::m_kd_tree is private, src/Frame.h:69)
    check(slamgpu_frame_upload(m_ctx, FRAME_SLOT, nkp, xy.data(), desc.data(), nullptr, matched.data()),
          "slamgpu_frame_upload");
    return nkp;
}

// Point state: unique points in first-visit order, observations in the host's iteration order -> resident set.
void MapMatcherGpu::upload_points(const std::vector<MapPoint*>& visit, std::vector<MapPoint*>& uniq,
                                  std::vector<int32_t>& order) const
{
    std::unordered_map<const MapPoint*, int> slot;
    uniq.clear();
    order.assign(visit.size(), -1);
    for (size_t v = 0; v < visit.size(); v++) {
        if (visit[v] == nullptr) {
            continue;                                           // match_for_fuse nullptr entry
        }
        auto it = slot.find(visit[v]);
        if (it == slot.end()) {
            it = slot.emplace(visit[v], static_cast<int>(uniq.size())).first;
            uniq.push_back(visit[v]);
        }
        order[v] = it->second;
    }
    const int n = static_cast<int>(uniq.size());
    std::vector<float> pos(3 * static_cast<size_t>(n));
    std::vector<int32_t> obs_offset(static_cast<size_t>(n) + 1, 0);
    std::vector<uint8_t> obs_desc;
    std::vector<float> obs_center;
    std::unordered_map<const KeyFrame*, Eigen::Vector3f> centers;
    for (int i = 0; i < n; i++) {
        const MapPoint& p = *uniq[i];
        pos[3 * i] = p.position().x();
        pos[3 * i + 1] = p.position().y();
        pos[3 * i + 2] = p.position().z();
        for (const auto& [obs_keyframe, obs_index] : p.observations()) {        // src/MapMatcher.cpp:86, same iteration order
            const cv::Mat d = obs_keyframe->descriptor(obs_index);
            const uint8_t* dp = d.ptr<uint8_t>(0);
            obs_desc.insert(obs_desc.end(), dp, dp + 32);
            auto c = centers.find(obs_keyframe);
            if (c == centers.end()) {
                c = centers.emplace(obs_keyframe, obs_keyframe->camera_center()).first;   // src/Frame.cpp:39-42
            }
            obs_center.push_back(c->second.x());
            obs_center.push_back(c->second.y());
            obs_center.push_back(c->second.z());
        }
        obs_offset[i + 1] = static_cast<int32_t>(obs_desc.size() / 32);
    }
    // normal / range = nullptr: derived on the device from the observers' camera centres in the
    // same order as MapPoint::avg_viewing_normal / observed_distance_range (src/MapPoint.cpp:24-45)
    check(slamgpu_points_upload(m_ctx, POINT_SLOT, n, pos.data(), nullptr, nullptr, obs_offset.data(), obs_desc.data(),
                                obs_center.data()),
          "slamgpu_points_upload");
}

slamgpu_proj_params MapMatcherGpu::params(bool replace) const
{
    slamgpu_proj_params prm;
    const Eigen::Matrix3f& K = m_camera.get_intrinsic_matrix();
    prm.fx = K(0, 0);
    prm.fy = K(1, 1);
    prm.cx = K(0, 2);
    prm.cy = K(1, 2);
    prm.width = m_camera.get_width();
    prm.height = m_camera.get_height();
    prm.radius = search_radius;
    prm.max_dist = m_max_descriptor_distance;
    prm.min_cos = min_viewing_angle_cosine;
    prm.nearer = max_nearer_ratio;
    prm.further = max_further_ratio;
    prm.replace = replace ? 1 : 0;
    return prm;
}

std::vector<MapPointMatch> MapMatcherGpu::project(const Frame& frame, const std::vector<MapPoint*>& visit,
                                                  KeyFrame* required_observer, bool replace) const
{
    const int nkp = upload_frame(frame);
    if (nkp == 0) {
        return {};
    }
    std::vector<MapPoint*> uniq;
    std::vector<int32_t> order;
    upload_points(visit, uniq, order);
    std::vector<uint8_t> skip(uniq.size());
    for (size_t i = 0; i < uniq.size(); i++) {
        const MapPoint& p = *uniq[i];
        // P0 is_matched(point) (src/MapMatcher.cpp:53) and P9 !is_observed_by(kf) (:169)
        skip[i] = (frame.is_matched(p) || (required_observer != nullptr && !p.is_observed_by(required_observer))) ? 1 : 0;
    }
    const slamgpu_proj_params prm = params(replace);
    std::vector<int32_t> out_point(static_cast<size_t>(nkp)), out_dist(static_cast<size_t>(nkp));
    const Eigen::Matrix4f& pose = frame.pose();                 // column-major, as the ABI expects
    check(slamgpu_match_projection(m_ctx, FRAME_SLOT, POINT_SLOT, pose.data(), &prm, skip.data(), order.data(),
                                   static_cast<int>(order.size()), out_point.data(), out_dist.data()),
          "slamgpu_match_projection");

    std::vector<MapPointMatch> matches;                         // accepted_matches, :34-43
    for (int i = 0; i < nkp; i++) {
        if (out_point[i] >= 0) {
            matches.push_back(MapPointMatch{*uniq[static_cast<size_t>(out_point[i])], static_cast<size_t>(i)});
        }
    }
    return matches;
}

std::pair<std::vector<MapPointMatch>, std::vector<MapPointMatch>>
MapMatcherGpu::track(Frame& frame, Map& map, KeyFrame& last_key_frame) const
{
    std::pair<std::vector<MapPointMatch>, std::vector<MapPointMatch>> result;
    const int nkp = upload_frame(frame);
    if (nkp == 0) {
        return result;
    }
    std::vector<MapPoint*> visit, uniq;
    for (auto& point : map) {                                   // map order, src/MapMatcher.cpp:168
        visit.push_back(&point);
    }
    std::vector<int32_t> order;
    upload_points(visit, uniq, order);                          // a Map holds every point once: uniq == visit
    std::vector<int32_t> matched_points, kf_points;
    for (size_t i = 0; i < uniq.size(); i++) {
        if (frame.is_matched(*uniq[i])) {
            matched_points.push_back(static_cast<int32_t>(i));  // :53
        }
        if (uniq[i]->is_observed_by(&last_key_frame)) {
            kf_points.push_back(static_cast<int32_t>(i));       // :169, ascending = map order
        }
    }
    const slamgpu_proj_params prm = params(false);
    const size_t n = static_cast<size_t>(nkp);
    std::vector<int32_t> kf_point(n), kf_dist(n), map_point(n), map_dist(n);
    const Eigen::Matrix4f& pose = frame.pose();
    check(slamgpu_track_match(m_ctx, FRAME_SLOT, POINT_SLOT, pose.data(), &prm, matched_points.data(),
                              static_cast<int>(matched_points.size()), kf_points.data(),
                              static_cast<int>(kf_points.size()), kf_point.data(), kf_dist.data(), map_point.data(),
                              map_dist.data()),
          "slamgpu_track_match");
    for (int i = 0; i < nkp; i++) {
        if (kf_point[i] >= 0) {
            result.first.push_back(MapPointMatch{*uniq[static_cast<size_t>(kf_point[i])], static_cast<size_t>(i)});
        }
        if (map_point[i] >= 0) {
            result.second.push_back(MapPointMatch{*uniq[static_cast<size_t>(map_point[i])], static_cast<size_t>(i)});
        }
    }
    for (const auto& match : result.first) {                    // src/Tracker.cpp:235-237
        frame.add_map_match(match);
    }
    for (const auto& match : result.second) {                   // :244-246
        frame.add_map_match(match);
    }
    return result;
}

std::vector<MapPointMatch> MapMatcherGpu::match_descriptors(const Frame& frame, const KeyFrame& key_frame) const
{
    std::vector<MapPoint*> points;
    std::vector<uint8_t> train;
    points.reserve(key_frame.num_map_matches());
    for (const auto& match : key_frame.map_matches()) {         // ascending keypoint index, :135-138
        points.push_back(&match.point);
        const cv::Mat d = key_frame.descriptor(match.keypoint_index);
        const uint8_t* dp = d.ptr<uint8_t>(0);
        train.insert(train.end(), dp, dp + 32);
    }
    const cv::Mat& query_mat = frame.features().descriptors;
    if (points.empty() || query_mat.empty()) {
        return {};                                              // :139-141
    }
    std::vector<uint8_t> query;
    append_rows(query_mat, query);
    const int nq = query_mat.rows;
    std::vector<int32_t> out_query(static_cast<size_t>(nq)), out_train(static_cast<size_t>(nq));
    int count = 0;
    check(slamgpu_match_descriptors(m_ctx, query.data(), nq, train.data(), static_cast<int>(points.size()),
                                    m_max_descriptor_distance, match_ratio, out_query.data(), out_train.data(), &count),
          "slamgpu_match_descriptors");
    std::vector<MapPointMatch> matches;
    for (int i = 0; i < count; i++) {                           // :159-160
        matches.push_back({*points[static_cast<size_t>(out_train[i])], static_cast<size_t>(out_query[i])});
    }
    return matches;
}

std::vector<std::vector<MapPointMatch>>
MapMatcherGpu::match_descriptors_batch(const Frame& frame, const std::vector<const KeyFrame*>& candidates) const
{
    std::vector<std::vector<MapPointMatch>> result(candidates.size());
    const cv::Mat& query_mat = frame.features().descriptors;
    if (candidates.empty() || query_mat.empty()) {
        return result;                                          // :139-141 for every candidate
    }
    // descriptor frame 0 = the query's rows; frame i + 1 = candidate i's matched rows, ascending keypoint index
    std::vector<uint8_t> desc;
    append_rows(query_mat, desc);
    const int nq = query_mat.rows;
    std::vector<int32_t> frame_offset{0, nq};
    std::vector<std::vector<MapPoint*>> points(candidates.size());
    std::vector<int32_t> pairs;
    for (size_t c = 0; c < candidates.size(); c++) {
        const KeyFrame& key_frame = *candidates[c];
        points[c].reserve(key_frame.num_map_matches());
        for (const auto& match : key_frame.map_matches()) {     // :135-138
            points[c].push_back(&match.point);
            const cv::Mat d = key_frame.descriptor(match.keypoint_index);
            const uint8_t* dp = d.ptr<uint8_t>(0);
            desc.insert(desc.end(), dp, dp + 32);
        }
        frame_offset.push_back(frame_offset.back() + static_cast<int32_t>(points[c].size()));
        pairs.push_back(0);
        pairs.push_back(static_cast<int32_t>(c) + 1);
    }
    const int n_pairs = static_cast<int>(candidates.size());
    const int64_t capacity = static_cast<int64_t>(nq) * n_pairs;   // a query row matches at most once per candidate
    std::vector<int32_t> out_offset(static_cast<size_t>(n_pairs) + 1), out_query(static_cast<size_t>(capacity)),
        out_train(static_cast<size_t>(capacity));
    int64_t total = 0;
    check(slamgpu_match_descriptors_batch(m_ctx, desc.data(), frame_offset.data(), n_pairs + 1, pairs.data(), n_pairs,
                                          m_max_descriptor_distance, match_ratio, out_offset.data(), out_query.data(),
                                          out_train.data(), capacity, &total),
          "slamgpu_match_descriptors_batch");
    for (int c = 0; c < n_pairs; c++) {
        for (int32_t i = out_offset[c]; i < out_offset[c + 1]; i++) {          // :159-160
            result[static_cast<size_t>(c)].push_back(
                {*points[static_cast<size_t>(c)][static_cast<size_t>(out_train[i])], static_cast<size_t>(out_query[i])});
        }
    }
    return result;
}

std::vector<FeatureMatch> MapMatcherGpu::match_features(const ExtractedFeatures& prev_features,
                                                        const ExtractedFeatures& features, float max_distance) const
{
    std::vector<uint8_t> prev, cur;
    append_rows(prev_features.descriptors, prev);
    append_rows(features.descriptors, cur);
    const int ncur = features.descriptors.rows;
    std::vector<int32_t> out_train(static_cast<size_t>(ncur) + 1), out_query(static_cast<size_t>(ncur) + 1);
    int count = 0;
    check(slamgpu_match_features(m_ctx, prev.data(), prev_features.descriptors.rows, cur.data(), ncur, max_distance,
                                 out_train.data(), out_query.data(), &count),
          "slamgpu_match_features");
    std::vector<FeatureMatch> feature_matches;
    for (int i = 0; i < count; i++) {
        feature_matches.push_back(FeatureMatch(out_train[i], out_query[i]));   // src/features/FeatureExtractor.cpp:26
    }
    return feature_matches;
}

} // namespace slam
