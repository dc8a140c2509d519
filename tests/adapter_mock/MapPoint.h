#pragma once
#include <Eigen/Dense>
#include <unordered_map>
namespace slam {
class KeyFrame;
class MapPoint {   // read accessors of reference src/MapPoint.h:16-22
  public:
    explicit MapPoint(const Eigen::Vector3f& position) : m_position(position) {}
    const Eigen::Vector3f& position() const { return m_position; }
    const std::unordered_map<KeyFrame*, size_t>& observations() const { return m_observations; }
    bool is_observed_by(KeyFrame* key_frame) const { return m_observations.find(key_frame) != m_observations.end(); }
    void add_observation(KeyFrame* key_frame, size_t index) { m_observations[key_frame] = index; }
  private:
    Eigen::Vector3f m_position;
    std::unordered_map<KeyFrame*, size_t> m_observations;   // pointer-keyed, like the reference (src/MapPoint.h:38)
};
} // namespace slam
