// MapChangeListener.h -- the notification interface between the reference's data model and a resident mirror of
// its map (MapMatcherGpu, MirrorMode::Hooks).  No dependencies: src/Map.h includes it and keeps one pointer.
#pragma once

namespace slam {

class KeyFrame;
class MapPoint;

/** What the data model reports to a resident mirror (MirrorMode::Hooks).  One call per reference mutator:
 *  Map::create_point / add_point (src/Map.cpp:9-61), Map::remove_point (:63-76), Map::associate / disassociate
 *  (:97-124, which fuse (:78-95) goes through), MapPoint::set_position (src/MapPoint.cpp:19-22) and
 *  Frame::set_pose of a key frame (src/Frame.cpp). */
class MapChangeListener {
  public:
    virtual ~MapChangeListener() = default;
    virtual void on_point_created(MapPoint& point) = 0;            // after push_back: the point is last in map order
    virtual void on_point_removed(MapPoint& point) = 0;            // before the erase
    virtual void on_observations_changed(MapPoint& point) = 0;
    virtual void on_position_changed(MapPoint& point) = 0;
    virtual void on_pose_changed(KeyFrame& key_frame) = 0;
};

} // namespace slam
