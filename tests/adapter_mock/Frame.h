#pragma once
#include <Eigen/Dense>
#include <opencv2/opencv.hpp>
#include <unordered_set>
#include <vector>
#include "MapPoint.h"
#include "features/FeatureExtractor.h"
#include "mock_math.h"
namespace slam {
class Frame {   // accessors of reference src/Frame.h:36-57; add / remove_map_match as src/Frame.cpp:80-116
  public:
    Frame(int index, const ExtractedFeatures& features) : m_index(index), m_features(features), m_map_matches(features.keypoints.size(), nullptr) {}
    const cv::Mat descriptor(size_t index) const { return m_features.descriptors.row(static_cast<int>(index)); }
    const ExtractedFeatures& features() const { return m_features; }
    const cv::KeyPoint& keypoint(size_t index) const { return m_features.keypoints[index]; }
    Eigen::Vector3f camera_center() const
    {   // -R^T t with the a0 + (a1 + a2) order (src/Frame.cpp:39-42)
        Eigen::Vector3f c;
        for (int i = 0; i < 3; i++) {
            volatile float p0 = -m_pose(0, i) * m_pose(0, 3), p1 = -m_pose(1, i) * m_pose(1, 3), p2 = -m_pose(2, i) * m_pose(2, 3);
            c.v[i] = mockmath::sum3(p0, p1, p2);
        }
        return c;
    }
    const Eigen::Matrix4f& pose() const { return m_pose; }
    void set_pose(const Eigen::Matrix4f& pose) { m_pose = pose; }
    void add_map_match(const MapPointMatch& match)
    {
        MapPoint* previous = m_map_matches[match.keypoint_index];
        if (previous == &match.point) return;
        if (previous != nullptr) m_matched_map_points.erase(previous);
        for (size_t i = 0; i < m_map_matches.size(); i++)
            if (m_map_matches[i] == &match.point && i != match.keypoint_index) m_map_matches[i] = nullptr;
        m_map_matches[match.keypoint_index] = &match.point;
        m_matched_map_points.insert(&match.point);
    }
    void remove_map_match(const MapPointMatch& match)
    {
        m_matched_map_points.erase(&match.point);
        for (auto& p : m_map_matches)
            if (p == &match.point) p = nullptr;
    }
    MapPoint& map_match(size_t index) const { return *m_map_matches[index]; }
    bool is_matched(size_t keypoint_index) const { return m_map_matches[keypoint_index] != nullptr; }
    bool is_matched(const MapPoint& point) const { return m_matched_map_points.count(&point) != 0; }
    size_t num_map_matches() const
    {
        size_t n = 0;
        for (auto* p : m_map_matches) n += p != nullptr;
        return n;
    }
    std::vector<MapPointMatch> map_matches() const   // ascending keypoint index, like MapPointIterator
    {
        std::vector<MapPointMatch> out;
        for (size_t i = 0; i < m_map_matches.size(); i++)
            if (m_map_matches[i]) out.push_back(MapPointMatch{*m_map_matches[i], i});
        return out;
    }
  private:
    size_t m_index;
    Eigen::Matrix4f m_pose;
    ExtractedFeatures m_features;
    std::vector<MapPoint*> m_map_matches;
    std::unordered_set<const MapPoint*> m_matched_map_points;
};
class KeyFrame : public Frame {
  public:
    using Frame::Frame;
};

inline Eigen::Vector3f MapPoint::avg_viewing_normal() const
{
    float acc[3] = {0.0F, 0.0F, 0.0F};
    for (const auto& [frame, index] : m_observations) {
        const Eigen::Vector3f C = frame->camera_center();
        const float d[3] = {m_position.x() - C.x(), m_position.y() - C.y(), m_position.z() - C.z()};
        float v[3], vv[3];
        mockmath::normalized3(d, v);
        mockmath::normalized3(v, vv);
        for (int k = 0; k < 3; k++) {
            volatile float t = acc[k] + vv[k];
            acc[k] = t;
        }
    }
    float n[3];
    mockmath::normalized3(acc, n);
    return Eigen::Vector3f(n[0], n[1], n[2]);
}
inline std::pair<float, float> MapPoint::observed_distance_range() const
{
    float nearest = std::numeric_limits<float>::max(), furthest = 0.0F;
    for (const auto& [frame, index] : m_observations) {
        const Eigen::Vector3f C = frame->camera_center();
        const float d[3] = {m_position.x() - C.x(), m_position.y() - C.y(), m_position.z() - C.z()};
        const float distance = mockmath::norm3(d);
        nearest = std::min(nearest, distance);
        furthest = std::max(furthest, distance);
    }
    return {nearest, furthest};
}
} // namespace slam
