#pragma once
#include <memory>
#include <vector>
#include "MapPoint.h"
namespace slam {
class Map {   // iteration of reference src/Map.h:59-63 over vector<unique_ptr<MapPoint>>
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
    MapPoint& add(const Eigen::Vector3f& p) { m_points.push_back(std::make_unique<MapPoint>(p)); return *m_points.back(); }
    MapPoint& at(size_t i) { return *m_points[i]; }
    size_t size() const { return m_points.size(); }
  private:
    std::vector<std::unique_ptr<MapPoint>> m_points;
};
} // namespace slam
