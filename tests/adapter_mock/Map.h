#pragma once
#include <algorithm>
#include <memory>
#include <vector>
#include "Frame.h"
#include "MapChangeListener.h"   // the notifications INTEGRATION.md adds to src/Map.cpp (one line per mutator)
#include "MapPoint.h"
namespace slam {
class Map {   // iteration of reference src/Map.h:59-63 over vector<unique_ptr<MapPoint>>; mutators as src/Map.cpp:9-124
  public:
    class iterator {
      public:
        explicit iterator(std::vector<std::unique_ptr<MapPoint>>::iterator it) : m_it(it) {}
        MapPoint& operator*() const { return **m_it; }
        iterator& operator++() { ++m_it; return *this; }
        bool operator!=(const iterator& o) const { return m_it != o.m_it; }
      private:
        std::vector<std::unique_ptr<MapPoint>>::iterator m_it;
    };
    iterator begin() { return iterator(m_points.begin()); }
    iterator end() { return iterator(m_points.end()); }
    size_t size() const { return m_points.size(); }
    MapPoint& at(size_t i) { return *m_points[i]; }

    MapChangeListener* listener = nullptr;     // what a maintainer adds: `if (listener) listener->on_...(...)`

    MapPoint& add(const Eigen::Vector3f& p)    // Map::add_point
    {
        m_points.push_back(std::make_unique<MapPoint>(p));
        if (listener) listener->on_point_created(*m_points.back());
        return *m_points.back();
    }
    MapPoint& create_point(const Eigen::Vector3f& position, KeyFrame& frame, size_t keypoint_index)
    {
        MapPoint& created = add(position);
        associate(frame, created, keypoint_index);
        return created;
    }
    void remove_point(MapPoint* point)
    {
        for (auto it = m_points.begin(); it != m_points.end(); ++it) {
            if (it->get() != point) continue;
            const auto observations = (*it)->observations();
            for (const auto& [frame, index] : observations) frame->remove_map_match({**it, index});
            if (listener) listener->on_point_removed(**it);
            m_points.erase(it);
            break;
        }
    }
    void fuse(MapPoint& kept, MapPoint& discarded)
    {
        if (&kept == &discarded) return;
        const auto observations = discarded.observations();
        for (const auto& [frame, index] : observations) {
            disassociate(*frame, discarded);
            if (kept.is_observed_by(frame) || frame->is_matched(index) || frame->is_matched(kept)) continue;
            associate(*frame, kept, index);
        }
        remove_point(&discarded);
    }
    void associate(KeyFrame& frame, MapPoint& point, size_t keypoint_index)
    {
        if (frame.is_matched(keypoint_index) && &frame.map_match(keypoint_index) == &point && point.is_observed_by(&frame)) return;
        if (frame.is_matched(keypoint_index)) {
            MapPoint& existing = frame.map_match(keypoint_index);
            if (&existing != &point) disassociate(frame, existing);
        }
        if (point.is_observed_by(&frame)) disassociate(frame, point);
        point.add_observation(&frame, keypoint_index);
        frame.add_map_match(MapPointMatch{point, keypoint_index});
        if (listener) listener->on_observations_changed(point);
    }
    void disassociate(KeyFrame& frame, MapPoint& point)
    {
        auto observation = point.observations().find(&frame);
        if (observation == point.observations().end()) return;
        frame.remove_map_match(MapPointMatch{point, observation->second});
        point.remove_observation(&frame);
        if (listener) listener->on_observations_changed(point);
    }
    void set_position(MapPoint& point, const Eigen::Vector3f& position)     // MapPoint::set_position + its notification
    {
        point.set_position(position);
        if (listener) listener->on_position_changed(point);
    }
    void set_pose(KeyFrame& frame, const Eigen::Matrix4f& pose)             // Frame::set_pose of a key frame + its notification
    {
        frame.set_pose(pose);
        if (listener) listener->on_pose_changed(frame);
    }
  private:
    std::vector<std::unique_ptr<MapPoint>> m_points;
};
} // namespace slam
