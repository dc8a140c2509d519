#pragma once
#include <Eigen/Dense>
#include <limits>
#include <unordered_map>
#include <utility>
namespace slam {
class KeyFrame;
class MapPoint {   // accessors of reference src/MapPoint.h:16-40
  public:
    explicit MapPoint(const Eigen::Vector3f& position) : m_position(position) {}
    const Eigen::Vector3f& position() const { return m_position; }
    void set_position(const Eigen::Vector3f& position) { m_position = position; }
    Eigen::Vector3f avg_viewing_normal() const;                 // src/MapPoint.cpp:24-33 (defined in Frame.h)
    std::pair<float, float> observed_distance_range() const;    // :35-45
    const std::unordered_map<KeyFrame*, size_t>& observations() const { return m_observations; }
    bool is_observed_by(KeyFrame* key_frame) const { return m_observations.find(key_frame) != m_observations.end(); }
    void add_observation(KeyFrame* key_frame, size_t index) { m_observations[key_frame] = index; }
    void remove_observation(KeyFrame* key_frame) { m_observations.erase(key_frame); }
    int creation_id = -1;                                       // test bookkeeping only
  private:
    Eigen::Vector3f m_position;
    std::unordered_map<KeyFrame*, size_t> m_observations;   // pointer-keyed, like the reference (src/MapPoint.h:38)
};
} // namespace slam
