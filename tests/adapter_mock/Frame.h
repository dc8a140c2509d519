#pragma once
#include <Eigen/Dense>
#include <opencv2/opencv.hpp>
#include <unordered_set>
#include <vector>
#include "features/FeatureExtractor.h"
namespace slam {
class Frame {   // read accessors of reference src/Frame.h:36-57
  public:
    Frame(int index, const ExtractedFeatures& features) : m_index(index), m_features(features), m_map_matches(features.keypoints.size(), nullptr) {}
    const cv::Mat descriptor(size_t index) const { return m_features.descriptors.row(static_cast<int>(index)); }
    const ExtractedFeatures& features() const { return m_features; }
    Eigen::Vector3f camera_center() const
    {   // -R^T t with the a0 + (a1 + a2) order of the oracle (src/Frame.cpp:39-42)
        Eigen::Vector3f c;
        for (int i = 0; i < 3; i++) {
            volatile float p0 = -m_pose(0, i) * m_pose(0, 3), p1 = -m_pose(1, i) * m_pose(1, 3), p2 = -m_pose(2, i) * m_pose(2, 3);
            volatile float s = p1 + p2;
            c.v[i] = p0 + s;
        }
        return c;
    }
    const Eigen::Matrix4f& pose() const { return m_pose; }
    void set_pose(const Eigen::Matrix4f& pose) { m_pose = pose; }
    void add_map_match(const MapPointMatch& match)
    {
        m_map_matches[match.keypoint_index] = &match.point;
        m_matched_map_points.insert(&match.point);
    }
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
} // namespace slam
