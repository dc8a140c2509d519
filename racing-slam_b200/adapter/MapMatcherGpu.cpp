// MapMatcherGpu.cpp -- see MapMatcherGpu.h.  Citations are relative to the reference tree.
#include "MapMatcherGpu.h"

#include <algorithm>
#include <chrono>
#include <cmath>
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
constexpr int POINT_SLOT = 0;      // per-call uploads (not attached)
constexpr int MIRROR_SLOT = 1;     // the resident mirror of the attached map

uint64_t mix(uint64_t h, uint64_t v)
{
    h ^= v + 0x9E3779B97F4A7C15ULL + (h << 6) + (h >> 2);
    return h;
}

uint32_t float_bits(float f)
{
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return u;
}

// KDTree2D::build as a permutation (src/KDTree.cpp:8-43): the node of range [s, e) is element (s + e) / 2 after
// std::nth_element on axis depth % 2 -- the same call on the same comparator, compiled by the reference's toolchain.
void kd_build_range(std::vector<uint32_t>& idx, const std::vector<float>& xy, size_t s, size_t e, int depth)
{
    if (s >= e) {
        return;
    }
    const size_t mid = (s + e) / 2;
    const int axis = depth % 2;
    std::nth_element(idx.begin() + static_cast<long>(s), idx.begin() + static_cast<long>(mid),
                     idx.begin() + static_cast<long>(e),
                     [&](uint32_t a, uint32_t b) { return xy[2 * a + axis] < xy[2 * b + axis]; });
    kd_build_range(idx, xy, s, mid, depth + 1);
    kd_build_range(idx, xy, mid + 1, e, depth + 1);
}

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
    if (m_mirror.map == &map) {
        return project_resident(frame, nullptr, false);         // every slot, in map order
    }
    std::vector<MapPoint*> visit;
    for (auto& point : map) {                                   // src/MapMatcher.cpp:168, map order
        visit.push_back(&point);
    }
    return project(frame, visit, nullptr, false);
}

std::vector<MapPointMatch> MapMatcherGpu::match_key_frame(const Frame& frame, Map& map, KeyFrame* key_frame) const
{
    if (m_mirror.map == &map) {
        if (key_frame == nullptr) {
            return project_resident(frame, nullptr, false);     // match(frame, map, nullptr) visits every point
        }
        sync();
        const std::vector<int32_t> order = observed_slots(key_frame);          // :169, ascending = map order
        return project_resident(frame, &order, false);
    }
    std::vector<MapPoint*> visit;
    for (auto& point : map) {
        visit.push_back(&point);
    }
    return project(frame, visit, key_frame, false);
}

std::vector<MapPointMatch> MapMatcherGpu::match_for_fuse(const Frame& frame, const std::vector<MapPoint*>& points) const
{
    if (attached()) {
        sync();
        std::vector<int32_t> order(points.size(), -1);          // nullptr entries stay -1: :119-121
        bool all_resident = true;
        for (size_t i = 0; i < points.size() && all_resident; i++) {
            if (points[i] == nullptr) {
                continue;
            }
            auto it = m_mirror.slot.find(points[i]);
            all_resident = it != m_mirror.slot.end();
            if (all_resident) {
                order[i] = it->second;
            }
        }
        if (all_resident) {
            return project_resident(frame, &order, true);
        }
    }
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
    // kd_perm = nullptr: the library rebuilds the tree with the same std::nth_element call on a thread of its
    // own (Frame::m_kd_tree is private, src/Frame.h:69); kd_from_host_toolchain builds it here instead
    std::vector<uint32_t> perm;
    if (kd_from_host_toolchain) {
        perm.resize(static_cast<size_t>(nkp));
        for (int i = 0; i < nkp; i++) {
            perm[i] = static_cast<uint32_t>(i);
        }
        kd_build_range(perm, xy, 0, static_cast<size_t>(nkp), 0);
    }
    check(slamgpu_frame_upload(m_ctx, FRAME_SLOT, nkp, xy.data(), desc.data(), perm.empty() ? nullptr : perm.data(),
                               matched.data()),
          "slamgpu_frame_upload");
    return nkp;
}

// Point state: unique points in first-visit order, observations in the host's iteration order -> resident set.
void MapMatcherGpu::upload_points(const std::vector<MapPoint*>& visit, std::vector<MapPoint*>& uniq,
                                  std::vector<int32_t>& order) const
{
    m_stats.legacy_uploads++;
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

// ---- the resident path: the frame goes up, two short index lists go up, nothing else ------------------

// slots of the points with frame.is_matched(point) (src/MapMatcher.cpp:53): at most one per keypoint
std::vector<int32_t> MapMatcherGpu::matched_slots(const Frame& frame) const
{
    std::vector<int32_t> slots;
    for (const auto& match : frame.map_matches()) {
        auto it = m_mirror.slot.find(&match.point);
        if (it != m_mirror.slot.end()) {
            slots.push_back(it->second);
        }
    }
    std::sort(slots.begin(), slots.end());
    slots.erase(std::unique(slots.begin(), slots.end()), slots.end());
    return slots;
}

// slots of the points with is_observed_by(key_frame) (src/MapMatcher.cpp:169), ascending = map order.  A key
// frame observes a point exactly where it holds it as a map match (Map::associate sets both sides,
// src/Map.cpp:97-113), so its match list names them all; is_observed_by filters a list that is ahead of the map.
std::vector<int32_t> MapMatcherGpu::observed_slots(KeyFrame* key_frame) const
{
    std::vector<int32_t> slots;
    for (const auto& match : key_frame->map_matches()) {
        if (!match.point.is_observed_by(key_frame)) {
            continue;
        }
        auto it = m_mirror.slot.find(&match.point);
        if (it != m_mirror.slot.end()) {
            slots.push_back(it->second);
        }
    }
    std::sort(slots.begin(), slots.end());
    slots.erase(std::unique(slots.begin(), slots.end()), slots.end());
    return slots;
}

std::vector<MapPointMatch> MapMatcherGpu::project_resident(const Frame& frame, const std::vector<int32_t>* order,
                                                           bool replace) const
{
    sync();
    if (order != nullptr && order->empty()) {
        return {};                                              // nothing to visit
    }
    const int nkp = upload_frame(frame);
    if (nkp == 0) {
        return {};
    }
    const std::vector<int32_t> matched = matched_slots(frame);
    const slamgpu_proj_params prm = params(replace);
    std::vector<int32_t> out_point(static_cast<size_t>(nkp)), out_dist(static_cast<size_t>(nkp));
    const Eigen::Matrix4f& pose = frame.pose();
    check(slamgpu_match_projection_lists(m_ctx, FRAME_SLOT, MIRROR_SLOT, pose.data(), &prm, matched.data(),
                                         static_cast<int>(matched.size()), order == nullptr ? nullptr : order->data(),
                                         order == nullptr ? 0 : static_cast<int>(order->size()), out_point.data(),
                                         out_dist.data()),
          "slamgpu_match_projection_lists");
    std::vector<MapPointMatch> matches;                         // accepted_matches, :34-43
    for (int i = 0; i < nkp; i++) {
        if (out_point[i] >= 0) {
            matches.push_back(MapPointMatch{*m_mirror.slot_point[static_cast<size_t>(out_point[i])], static_cast<size_t>(i)});
        }
    }
    return matches;
}

std::pair<std::vector<MapPointMatch>, std::vector<MapPointMatch>>
MapMatcherGpu::track(Frame& frame, Map& map, KeyFrame& last_key_frame) const
{
    std::pair<std::vector<MapPointMatch>, std::vector<MapPointMatch>> result;
    const bool resident = m_mirror.map == &map;
    if (resident) {
        sync();
    }
    const int nkp = upload_frame(frame);
    if (nkp == 0) {
        return result;
    }
    std::vector<MapPoint*> uniq;
    std::vector<int32_t> matched_points, kf_points;
    int set_id = POINT_SLOT;
    if (resident) {
        set_id = MIRROR_SLOT;
        matched_points = matched_slots(frame);                  // :53
        kf_points = observed_slots(&last_key_frame);     // :169, ascending = map order
    } else {
        std::vector<MapPoint*> visit;
        for (auto& point : map) {                               // map order, src/MapMatcher.cpp:168
            visit.push_back(&point);
        }
        std::vector<int32_t> order;
        upload_points(visit, uniq, order);                      // a Map holds every point once: uniq == visit
        for (size_t i = 0; i < uniq.size(); i++) {
            if (frame.is_matched(*uniq[i])) {
                matched_points.push_back(static_cast<int32_t>(i));  // :53
            }
            if (uniq[i]->is_observed_by(&last_key_frame)) {
                kf_points.push_back(static_cast<int32_t>(i));       // :169, ascending = map order
            }
        }
    }
    const std::vector<MapPoint*>& by_index = resident ? m_mirror.slot_point : uniq;
    const slamgpu_proj_params prm = params(false);
    const size_t n = static_cast<size_t>(nkp);
    std::vector<int32_t> kf_point(n), kf_dist(n), map_point(n), map_dist(n);
    const Eigen::Matrix4f& pose = frame.pose();
    check(slamgpu_track_match(m_ctx, FRAME_SLOT, set_id, pose.data(), &prm, matched_points.data(),
                              static_cast<int>(matched_points.size()), kf_points.data(),
                              static_cast<int>(kf_points.size()), kf_point.data(), kf_dist.data(), map_point.data(),
                              map_dist.data()),
          "slamgpu_track_match");
    for (int i = 0; i < nkp; i++) {
        if (kf_point[i] >= 0) {
            result.first.push_back(MapPointMatch{*by_index[static_cast<size_t>(kf_point[i])], static_cast<size_t>(i)});
        }
        if (map_point[i] >= 0) {
            result.second.push_back(MapPointMatch{*by_index[static_cast<size_t>(map_point[i])], static_cast<size_t>(i)});
        }
    }
    // The reference's own bookkeeping, unchanged: Frame::add_map_match sweeps all keypoints per match
    // (src/Frame.cpp:90-98), so this is O(matches x keypoints) host work whichever matcher produced the lists.
    const auto t0 = std::chrono::steady_clock::now();
    for (const auto& match : result.first) {                    // src/Tracker.cpp:235-237
        frame.add_map_match(match);
    }
    for (const auto& match : result.second) {                   // :244-246
        frame.add_map_match(match);
    }
    m_stats.last_apply_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return result;
}

// ---- the mirror --------------------------------------------------------------------------------------

void MapMatcherGpu::on_point_created(MapPoint& point)
{
    if (attached() && m_mirror.mode == MirrorMode::Hooks) {
        m_mirror.created.push_back(&point);
    }
}

void MapMatcherGpu::on_point_removed(MapPoint& point)
{
    if (!attached() || m_mirror.mode != MirrorMode::Hooks) {
        return;
    }
    auto it = m_mirror.slot.find(&point);
    if (it != m_mirror.slot.end()) {
        m_mirror.removed_slots.push_back(it->second);
        m_mirror.slot_point[static_cast<size_t>(it->second)] = nullptr;
        m_mirror.slot.erase(it);                                // the address may be handed out again
    } else {                                                    // created and removed between two syncs
        auto c = std::find(m_mirror.created.begin(), m_mirror.created.end(), &point);
        if (c != m_mirror.created.end()) {
            m_mirror.created.erase(c);
        }
    }
    m_mirror.obs_dirty.erase(&point);
    m_mirror.pos_dirty.erase(&point);
}

void MapMatcherGpu::on_observations_changed(MapPoint& point)
{
    if (attached() && m_mirror.mode == MirrorMode::Hooks) {
        m_mirror.obs_dirty.insert(&point);
    }
}

void MapMatcherGpu::on_position_changed(MapPoint& point)
{
    if (attached() && m_mirror.mode == MirrorMode::Hooks) {
        m_mirror.pos_dirty.insert(&point);
    }
}

void MapMatcherGpu::on_pose_changed(KeyFrame& key_frame)
{
    if (attached() && m_mirror.mode == MirrorMode::Hooks) {
        m_mirror.pose_dirty.insert(&key_frame);
    }
}

void MapMatcherGpu::attach(Map& map, MirrorMode mode)
{
    detach();
    m_mirror.map = &map;
    m_mirror.mode = mode;
    const size_t n = map.size();
    // room for a run of BASELINE C3's size without a reallocation on the device (45 MB of the 180 GB)
    check(slamgpu_points_create(m_ctx, MIRROR_SLOT, static_cast<int>(std::max<size_t>(2 * n, size_t(1) << 18)),
                                static_cast<int64_t>(std::max<size_t>(8 * n, size_t(1) << 20)), 4096),
          "slamgpu_points_create");
    for (auto& point : map) {
        m_mirror.created.push_back(&point);
    }
    m_stats.full_uploads++;
    flush();
}

void MapMatcherGpu::detach()
{
    if (attached()) {
        slamgpu_points_release(m_ctx, MIRROR_SLOT);
    }
    m_mirror = Mirror();
}

int32_t MapMatcherGpu::observer_id(const KeyFrame* key_frame, std::vector<int32_t>& new_ids,
                                   std::vector<float>& new_centers) const
{
    auto it = m_mirror.observer.find(key_frame);
    if (it != m_mirror.observer.end()) {
        return it->second;
    }
    const int32_t id = static_cast<int32_t>(m_mirror.observer_frame.size());
    m_mirror.observer.emplace(key_frame, id);
    m_mirror.observer_frame.push_back(key_frame);
    const Eigen::Vector3f c = key_frame->camera_center();       // src/Frame.cpp:39-42, the host's own arithmetic
    const float cf[3] = {c.x(), c.y(), c.z()};
    m_mirror.observer_center.insert(m_mirror.observer_center.end(), cf, cf + 3);
    new_ids.push_back(id);
    new_centers.insert(new_centers.end(), cf, cf + 3);
    return id;
}

// observations of each point in the host's iteration order (src/MapMatcher.cpp:86), observers by id
void MapMatcherGpu::flatten_runs(const std::vector<MapPoint*>& points, std::vector<int32_t>& obs_offset,
                                 std::vector<uint8_t>& obs_desc, std::vector<int32_t>& obs_observer,
                                 std::vector<int32_t>& new_ids, std::vector<float>& new_centers,
                                 std::vector<uint64_t>* hashes) const
{
    obs_offset.assign(points.size() + 1, 0);
    for (size_t i = 0; i < points.size(); i++) {
        uint64_t h = 0x243F6A8885A308D3ULL;
        for (const auto& [obs_keyframe, obs_index] : points[i]->observations()) {
            const cv::Mat d = obs_keyframe->descriptor(obs_index);
            const uint8_t* dp = d.ptr<uint8_t>(0);
            obs_desc.insert(obs_desc.end(), dp, dp + 32);
            obs_observer.push_back(observer_id(obs_keyframe, new_ids, new_centers));
            h = mix(mix(h, reinterpret_cast<uintptr_t>(obs_keyframe)), static_cast<uint64_t>(obs_index));
        }
        obs_offset[i + 1] = static_cast<int32_t>(obs_observer.size());
        if (hashes != nullptr) {
            hashes->push_back(h);
        }
    }
}

// MirrorMode::Scan: one walk over the map, comparing with what was last uploaded
void MapMatcherGpu::scan_for_changes() const
{
    Mirror& m = m_mirror;
    std::vector<uint8_t> seen(m.slot_point.size(), 0);
    int32_t prev_slot = -1;
    bool broken = false;
    for (auto& point : *m.map) {
        auto it = m.slot.find(&point);
        if (it == m.slot.end()) {
            m.created.push_back(&point);
            continue;
        }
        // a resident point after a new one, or out of slot order: an address was recycled -- start over
        if (!m.created.empty() || it->second <= prev_slot) {
            broken = true;
            break;
        }
        const size_t s = static_cast<size_t>(it->second);
        prev_slot = it->second;
        seen[s] = 1;
        const Eigen::Vector3f& x = point.position();
        if (float_bits(x.x()) != float_bits(m.pos[3 * s]) || float_bits(x.y()) != float_bits(m.pos[3 * s + 1]) ||
            float_bits(x.z()) != float_bits(m.pos[3 * s + 2])) {
            m.pos_dirty.insert(&point);
        }
        uint64_t h = 0x243F6A8885A308D3ULL;
        for (const auto& [obs_keyframe, obs_index] : point.observations()) {
            h = mix(mix(h, reinterpret_cast<uintptr_t>(obs_keyframe)), static_cast<uint64_t>(obs_index));
        }
        if (h != m.obs_hash[s]) {
            m.obs_dirty.insert(&point);
        }
    }
    if (broken) {
        Map* map = m.map;
        const MirrorMode mode = m.mode;
        const_cast<MapMatcherGpu*>(this)->attach(*map, mode);
        return;
    }
    for (size_t s = 0; s < m.slot_point.size(); s++) {
        if (m.slot_point[s] != nullptr && !seen[s]) {
            m.removed_slots.push_back(static_cast<int32_t>(s));
            m.slot.erase(m.slot_point[s]);
            m.slot_point[s] = nullptr;
        }
    }
    for (const KeyFrame* kf : m.observer_frame) {
        m.pose_dirty.insert(kf);                                 // flush() compares the centres
    }
}

void MapMatcherGpu::sync() const
{
    if (!attached()) {
        return;
    }
    if (m_mirror.mode == MirrorMode::Scan) {
        scan_for_changes();
    }
    flush();
}

void MapMatcherGpu::flush() const
{
    Mirror& m = m_mirror;
    if (m.created.empty() && m.removed_slots.empty() && m.obs_dirty.empty() && m.pos_dirty.empty() && m.pose_dirty.empty()) {
        return;
    }
    m_stats.syncs++;
    // 1. Map::remove_point (src/Map.cpp:63-76): tombstones
    if (!m.removed_slots.empty()) {
        check(slamgpu_points_remove(m_ctx, MIRROR_SLOT, m.removed_slots.data(), static_cast<int>(m.removed_slots.size())),
              "slamgpu_points_remove");
        m_stats.removed += m.removed_slots.size();
        m.tombstones += m.removed_slots.size();
        m.removed_slots.clear();
    }
    // 2. what changed on live, resident points (new points are flattened as they are now)
    std::vector<MapPoint*> obs_points, pos_points;
    std::vector<int32_t> obs_slots, pos_slots;
    for (MapPoint* p : m.obs_dirty) {
        auto it = m.slot.find(p);
        if (it != m.slot.end()) {
            obs_points.push_back(p);
        }
    }
    for (MapPoint* p : m.pos_dirty) {
        auto it = m.slot.find(p);
        if (it != m.slot.end()) {
            pos_points.push_back(p);
        }
    }
    // deterministic order (slot order) whatever the sets' iteration order
    auto by_slot = [&](MapPoint* a, MapPoint* b) { return m.slot[a] < m.slot[b]; };
    std::sort(obs_points.begin(), obs_points.end(), by_slot);
    std::sort(pos_points.begin(), pos_points.end(), by_slot);
    for (MapPoint* p : obs_points) {
        obs_slots.push_back(m.slot[p]);
    }
    for (MapPoint* p : pos_points) {
        pos_slots.push_back(m.slot[p]);
    }
    std::vector<int32_t> new_ids, o_off, o_kf, c_off, c_kf;
    std::vector<float> new_centers;
    std::vector<uint8_t> o_desc, c_desc;
    std::vector<uint64_t> o_hash, c_hash;
    flatten_runs(obs_points, o_off, o_desc, o_kf, new_ids, new_centers, &o_hash);
    flatten_runs(m.created, c_off, c_desc, c_kf, new_ids, new_centers, &c_hash);
    // 3. observers: new key frames, and key frames whose pose moved (bundle adjustment, pose graph)
    for (const KeyFrame* kf : m.pose_dirty) {
        auto it = m.observer.find(kf);
        if (it == m.observer.end()) {
            continue;                                            // observes nothing yet
        }
        const size_t id = static_cast<size_t>(it->second);
        if (std::find(new_ids.begin(), new_ids.end(), it->second) != new_ids.end()) {
            continue;                                            // registered a moment ago with its current centre
        }
        const Eigen::Vector3f c = kf->camera_center();
        const float cf[3] = {c.x(), c.y(), c.z()};
        if (float_bits(cf[0]) != float_bits(m.observer_center[3 * id]) || float_bits(cf[1]) != float_bits(m.observer_center[3 * id + 1]) ||
            float_bits(cf[2]) != float_bits(m.observer_center[3 * id + 2])) {
            std::copy(cf, cf + 3, m.observer_center.begin() + static_cast<long>(3 * id));
            new_ids.push_back(it->second);
            new_centers.insert(new_centers.end(), cf, cf + 3);
        }
    }
    m.pose_dirty.clear();
    if (!new_ids.empty()) {
        check(slamgpu_observer_centers_update(m_ctx, MIRROR_SLOT, new_ids.data(), static_cast<int>(new_ids.size()),
                                              new_centers.data()),
              "slamgpu_observer_centers_update");
        m_stats.observer_updates += new_ids.size();
    }
    // 4. Map::associate / disassociate / fuse (src/Map.cpp:78-124): the whole run, in the host's current order
    if (!obs_points.empty()) {
        check(slamgpu_points_set_observations(m_ctx, MIRROR_SLOT, obs_slots.data(), static_cast<int>(obs_slots.size()),
                                              o_off.data(), o_desc.data(), o_kf.data()),
              "slamgpu_points_set_observations");
        for (size_t i = 0; i < obs_slots.size(); i++) {
            m.obs_hash[static_cast<size_t>(obs_slots[i])] = o_hash[i];
        }
        m_stats.observation_updates += obs_points.size();
    }
    m.obs_dirty.clear();
    // 5. MapPoint::set_position (src/MapPoint.cpp:19-22)
    if (!pos_points.empty()) {
        std::vector<float> pos(3 * pos_points.size());
        for (size_t i = 0; i < pos_points.size(); i++) {
            const Eigen::Vector3f& x = pos_points[i]->position();
            pos[3 * i] = x.x();
            pos[3 * i + 1] = x.y();
            pos[3 * i + 2] = x.z();
            std::copy(pos.begin() + static_cast<long>(3 * i), pos.begin() + static_cast<long>(3 * i + 3),
                      m.pos.begin() + static_cast<long>(3 * static_cast<size_t>(pos_slots[i])));
        }
        check(slamgpu_points_set_position(m_ctx, MIRROR_SLOT, pos_slots.data(), static_cast<int>(pos_slots.size()), pos.data()),
              "slamgpu_points_set_position");
        m_stats.position_updates += pos_points.size();
    }
    m.pos_dirty.clear();
    // 6. Map::create_point / add_point (src/Map.cpp:9-61): new slots at the end, in map order
    if (!m.created.empty()) {
        std::vector<float> pos(3 * m.created.size());
        for (size_t i = 0; i < m.created.size(); i++) {
            const Eigen::Vector3f& x = m.created[i]->position();
            pos[3 * i] = x.x();
            pos[3 * i + 1] = x.y();
            pos[3 * i + 2] = x.z();
        }
        check(slamgpu_points_append_observed(m_ctx, MIRROR_SLOT, static_cast<int>(m.created.size()), pos.data(),
                                             c_off.data(), c_desc.data(), c_kf.data()),
              "slamgpu_points_append_observed");
        for (size_t i = 0; i < m.created.size(); i++) {
            m.slot[m.created[i]] = static_cast<int32_t>(m.slot_point.size());
            m.slot_point.push_back(m.created[i]);
            m.obs_hash.push_back(c_hash[i]);
        }
        m.pos.insert(m.pos.end(), pos.begin(), pos.end());
        m_stats.appended += m.created.size();
        m.created.clear();
    }
    // 7. squeeze the tombstones out once they outweigh a quarter of the slots (order kept: vector::erase)
    if (m.tombstones >= 256 && m.tombstones * 4 > m.slot_point.size()) {
        std::vector<int32_t> remap(m.slot_point.size());
        int n_after = 0;
        check(slamgpu_points_compact(m_ctx, MIRROR_SLOT, remap.data(), &n_after), "slamgpu_points_compact");
        std::vector<MapPoint*> sp(static_cast<size_t>(n_after));
        std::vector<float> pos(3 * static_cast<size_t>(n_after));
        std::vector<uint64_t> oh(static_cast<size_t>(n_after));
        for (size_t s = 0; s < m.slot_point.size(); s++) {
            if (remap[s] < 0) {
                continue;
            }
            const size_t j = static_cast<size_t>(remap[s]);
            sp[j] = m.slot_point[s];
            std::copy(m.pos.begin() + static_cast<long>(3 * s), m.pos.begin() + static_cast<long>(3 * s + 3),
                      pos.begin() + static_cast<long>(3 * j));
            oh[j] = m.obs_hash[s];
            m.slot[m.slot_point[s]] = remap[s];
        }
        m.slot_point.swap(sp);
        m.pos.swap(pos);
        m.obs_hash.swap(oh);
        m.tombstones = 0;
        m_stats.compactions++;
    }
}

size_t MapMatcherGpu::audit() const
{
    if (!attached()) {
        return 0;
    }
    sync();
    const Mirror& m = m_mirror;
    size_t bad = 0;
    int n_slots = 0;
    int64_t n_obs = 0;
    check(slamgpu_points_count(m_ctx, MIRROR_SLOT, &n_slots, &n_obs), "slamgpu_points_count");
    bad += static_cast<size_t>(n_slots) != m.slot_point.size();
    std::vector<float> normal(3 * static_cast<size_t>(n_slots)), range(2 * static_cast<size_t>(n_slots));
    check(slamgpu_points_get_stats(m_ctx, MIRROR_SLOT, normal.data(), range.data()), "slamgpu_points_get_stats");
    int32_t prev = -1;
    size_t live = 0, observations = 0;
    for (auto& point : *m.map) {
        auto it = m.slot.find(&point);
        if (it == m.slot.end() || it->second <= prev) {
            bad++;                                               // not resident, or map order lost
            continue;
        }
        prev = it->second;
        live++;
        const size_t s = static_cast<size_t>(it->second);
        bad += m.slot_point[s] != &point;
        const Eigen::Vector3f& x = point.position();
        bad += float_bits(x.x()) != float_bits(m.pos[3 * s]) || float_bits(x.y()) != float_bits(m.pos[3 * s + 1]) ||
               float_bits(x.z()) != float_bits(m.pos[3 * s + 2]);
        uint64_t h = 0x243F6A8885A308D3ULL;
        for (const auto& [obs_keyframe, obs_index] : point.observations()) {
            h = mix(mix(h, reinterpret_cast<uintptr_t>(obs_keyframe)), static_cast<uint64_t>(obs_index));
            observations++;
            auto o = m.observer.find(obs_keyframe);
            if (o == m.observer.end()) {
                bad++;
                continue;
            }
            const Eigen::Vector3f c = obs_keyframe->camera_center();
            const size_t id = static_cast<size_t>(o->second);
            // the library derives the CURRENT frame's centre itself (-R^T t in the order it assumes Eigen uses):
            // check that order on every observer's pose against the host's own camera_center()
            float lib_c[3];
            check(slamgpu_host_camera_center(obs_keyframe->pose().data(), lib_c), "slamgpu_host_camera_center");
            bad += float_bits(c.x()) != float_bits(lib_c[0]) || float_bits(c.y()) != float_bits(lib_c[1]) ||
                   float_bits(c.z()) != float_bits(lib_c[2]);
            bad += float_bits(c.x()) != float_bits(m.observer_center[3 * id]) || float_bits(c.y()) != float_bits(m.observer_center[3 * id + 1]) ||
                   float_bits(c.z()) != float_bits(m.observer_center[3 * id + 2]);
        }
        bad += h != m.obs_hash[s];
        // the device's re-derived statistics against the host's own (the reference's Eigen code when compiled there)
        if (!point.observations().empty()) {
            const Eigen::Vector3f nrm = point.avg_viewing_normal();              // src/MapPoint.cpp:24-33
            const std::pair<float, float> rg = point.observed_distance_range();  // :35-45
            bad += float_bits(nrm.x()) != float_bits(normal[3 * s]) || float_bits(nrm.y()) != float_bits(normal[3 * s + 1]) ||
                   float_bits(nrm.z()) != float_bits(normal[3 * s + 2]);
            bad += float_bits(rg.first) != float_bits(range[2 * s]) || float_bits(rg.second) != float_bits(range[2 * s + 1]);
        }
    }
    bad += live != m.slot.size();
    bad += static_cast<int64_t>(observations) != n_obs;
    return bad;
}

size_t MapMatcherGpu::self_check(unsigned seed, int n) const
{
    // Seeded points in front of seeded poses; the reference's own functions (Eigen inside) against the library's
    // restatement of their operation order, bit for bit.
    uint64_t state = 0x9E3779B97F4A7C15ULL * (seed + 1);
    auto uniform = [&](float lo, float hi) {
        state = state * 6364136223846793005ULL + 1442695040888963407ULL;
        return lo + (hi - lo) * static_cast<float>((state >> 40) & 0xFFFFFF) / 16777216.0F;
    };
    const slamgpu_proj_params prm = params(false);
    size_t bad = 0;
    for (int i = 0; i < n; i++) {
        // a rotation from three small angles, built in double and rounded once (any orthonormal-ish matrix will do:
        // the check is about the evaluation order, not about the pose)
        const double ax = uniform(-0.5F, 0.5F), ay = uniform(-0.5F, 0.5F), az = uniform(-0.5F, 0.5F);
        const double cx = std::cos(ax), sx = std::sin(ax), cy = std::cos(ay), sy = std::sin(ay), cz = std::cos(az), sz = std::sin(az);
        const double R[9] = {cy * cz, -cy * sz, sy, sx * sy * cz + cx * sz, -sx * sy * sz + cx * cz, -sx * cy,
                             -cx * sy * cz + sx * sz, cx * sy * sz + sx * cz, cx * cy};
        Eigen::Matrix4f pose;
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) {
                pose(r, c) = static_cast<float>(R[3 * r + c]);
            }
            pose(r, 3) = uniform(-3.0F, 3.0F);
            pose(3, r) = 0.0F;
        }
        pose(3, 3) = 1.0F;
        // points whose projection spans the whole image, including u >= 1024 px where one ulp is 1.2e-4 px
        const float z = uniform(2.0F, 60.0F);
        const float u = uniform(0.0F, static_cast<float>(prm.width)), v = uniform(0.0F, static_cast<float>(prm.height));
        const Eigen::Vector3f Xc((u - prm.cx) / prm.fx * z, (v - prm.cy) / prm.fy * z, z);
        // world point = R^T (Xc - t), any rounding is fine: it is an input
        Eigen::Vector3f X(0, 0, 0);
        {
            const float d[3] = {Xc.x() - pose(0, 3), Xc.y() - pose(1, 3), Xc.z() - pose(2, 3)};
            X = Eigen::Vector3f(pose(0, 0) * d[0] + pose(1, 0) * d[1] + pose(2, 0) * d[2],
                                pose(0, 1) * d[0] + pose(1, 1) * d[1] + pose(2, 1) * d[2],
                                pose(0, 2) * d[0] + pose(1, 2) * d[1] + pose(2, 2) * d[2]);
        }
        const float xw[3] = {X.x(), X.y(), X.z()};
        float uv_lib[2];
        check(slamgpu_host_project(pose.data(), &prm, xw, uv_lib), "slamgpu_host_project");
        const Eigen::Vector2f uv_ref = m_camera.project(pose, X);               // src/Camera.cpp:25-32
        bad += float_bits(uv_ref.x()) != float_bits(uv_lib[0]) || float_bits(uv_ref.y()) != float_bits(uv_lib[1]);
    }
    return bad;
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
