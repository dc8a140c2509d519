// MapMatcherGpu.h -- drop-in for slam::MapMatcher (reference src/MapMatcher.h:16-38) on libslamgpu.
//
// Same constructor and the same four entry points with identical signatures and return types, so
// the three owners (src/Tracker.cpp:42, src/Mapper.cpp:87, src/LoopDetector.cpp:319) only change
// the member type.  Compiled inside the reference tree (include path = its src/): it uses only the
// read accessors SURVEY 8b lists.  All matching arithmetic runs on the GPU through the C ABI
// (include/slamgpu.h); this class flattens the object graph to POD arrays and maps indices back.
//
// RESIDENT MAP.  By default every projection call flattens and uploads the points it visits, which is
// O(|map|) host work per call.  attach(map) switches the matcher to a persistent MIRROR of that map on the
// device (a mutable point set, slamgpu_points_create): point -> slot, key frame -> observer id.  From then on
// match_map / match_key_frame / match_for_fuse / track upload only the frame and the DELTAS since the last
// call.  The mirror learns about changes in one of two ways:
//   MirrorMode::Hooks  the data model tells it (MapChangeListener; six one-line calls in src/Map.cpp,
//                      src/MapPoint.cpp and where key-frame poses are set -- INTEGRATION.md shows them): O(changes)
//   MirrorMode::Scan   no change to the reference at all: before each call the map is walked once and every
//                      point's position / observation list / observers' centres are compared with what was
//                      last uploaded (pointer walks and a hash, no descriptor traffic): O(|map|) but light
#pragma once

#include <cstdint>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>

#include "MapChangeListener.h"
#include "features/FeatureExtractor.h"   // slam::MapPointMatch, slam::FeatureMatch, slam::ExtractedFeatures
#include "slamgpu.h"

namespace slam {

class Camera;
class Frame;
class KeyFrame;
class Map;
class MapPoint;

enum class MirrorMode { Hooks, Scan };

class MapMatcherGpu : public MapChangeListener {
  public:
    MapMatcherGpu(const Camera& camera, float max_descriptor_distance, cv::NormTypes norm_type, int device = 0);
    ~MapMatcherGpu() override;
    MapMatcherGpu(const MapMatcherGpu&) = delete;
    MapMatcherGpu& operator=(const MapMatcherGpu&) = delete;

    /** Match to every point in the map via reprojection (src/MapMatcher.cpp:107-110) */
    std::vector<MapPointMatch> match_map(const Frame& frame, Map& map) const;

    /** Match to a key frame's points via reprojection (:112-115) */
    std::vector<MapPointMatch> match_key_frame(const Frame& frame, Map& map, KeyFrame* key_frame) const;

    /** Project source points into frame, including keypoints that already have a map point (:117-127) */
    std::vector<MapPointMatch> match_for_fuse(const Frame& frame, const std::vector<MapPoint*>& points) const;

    /** Descriptor match to a key frame's points via knn (:129-163) */
    std::vector<MapPointMatch> match_descriptors(const Frame& frame, const KeyFrame& key_frame) const;

    /** verify_pnp's first step for every retrieved loop candidate at once (src/LoopDetector.cpp:181 runs
     *  match_descriptors once per candidate): result[i] == match_descriptors(frame, *candidates[i]), from ONE batched
     *  device call (slamgpu_match_descriptors_batch: the query rows go up once, every candidate's matched rows once). */
    std::vector<std::vector<MapPointMatch>> match_descriptors_batch(const Frame& frame,
                                                                    const std::vector<const KeyFrame*>& candidates) const;

    /** BaseFeatureExtractor::match_features (src/features/FeatureExtractor.cpp:15-30) on the same context */
    std::vector<FeatureMatch> match_features(const ExtractedFeatures& prev_features, const ExtractedFeatures& features,
                                             float max_distance) const;

    /** Tracker::match_with_last_key_frame followed by Tracker::match_with_map (src/Tracker.cpp:232-248) as ONE
     *  chained device call (slamgpu_track_match): both match lists are applied to `frame` with add_map_match
     *  exactly as the two loops do.  Returns {key-frame matches, map matches}; the Tracker replaces its two
     *  calls by this one. */
    std::pair<std::vector<MapPointMatch>, std::vector<MapPointMatch>> track(Frame& frame, Map& map,
                                                                            KeyFrame& last_key_frame) const;

    /** The libslamgpu context this matcher owns (for helpers that share it, e.g. LoopDetectorGpu). */
    slamgpu_ctx* context() const { return m_ctx; }

    // ---- resident map ----
    /** Keep `map` mirrored on the device from now on (full upload now, deltas afterwards). */
    void attach(Map& map, MirrorMode mode = MirrorMode::Hooks);
    void detach();
    bool attached() const { return m_mirror.map != nullptr; }
    /** Send the pending changes now (the match calls do it themselves; a Mapper may call it right after
     *  Mapper::insert so that the upload overlaps whatever the host does before the next frame). */
    void sync() const;
    /** Flatten the attached map from scratch and compare it with what the mirror believes is resident
     *  (slots, order, positions, observation lists, observer centres).  For tests and for a maintainer who
     *  wants to check that every mutator of their tree reports to the listener.  Returns the number of
     *  differences found (0 = consistent). */
    size_t audit() const;
    /** Counters of the resident path (for the bench and the tests). */
    struct MirrorStats {
        size_t syncs = 0, appended = 0, removed = 0, observation_updates = 0, position_updates = 0, observer_updates = 0,
               compactions = 0, full_uploads = 0, legacy_uploads = 0;
        double last_apply_ms = 0.0;   // host time of the last track()'s add_map_match loops (the reference's own code)
    };
    const MirrorStats& mirror_stats() const { return m_stats; }

    // MapChangeListener (MirrorMode::Hooks): record only; the work happens in sync()
    void on_point_created(MapPoint& point) override;
    void on_point_removed(MapPoint& point) override;
    void on_observations_changed(MapPoint& point) override;
    void on_position_changed(MapPoint& point) override;
    void on_pose_changed(KeyFrame& key_frame) override;

    /** Bit-level check of the float expressions the device restates against THIS build's Eigen:
     *  Camera::project (src/Camera.cpp:25-32), Frame::camera_center (src/Frame.cpp:39-42), avg_viewing_normal and
     *  observed_distance_range (src/MapPoint.cpp:24-45) on seeded inputs (including u >= 1024 px, where one float
     *  ulp exceeds the 1e-4 px tolerance).  Needs a device.  Returns the number of mismatching values (0 = the
     *  operation order the library assumes is the one this Eigen build evaluates).  See tools/eigen_order_check.cpp
     *  for the stand-alone form a maintainer runs first. */
    size_t self_check(unsigned seed = 1, int n = 4096) const;

    // Matching constants of src/MapMatcher.cpp:12-18, runtime-settable (defaults = the reference's)
    float search_radius = 20.0F;
    float min_viewing_angle_cosine = 0.5F;
    float max_nearer_ratio = 2.0F;
    float max_further_ratio = 1.25F;
    float match_ratio = 0.75F;
    /** Build the KD permutation here, with the std::nth_element of the toolchain that compiled the reference
     *  (the adapter is compiled with it), instead of inside libslamgpu.so: ties between equal coordinates are
     *  decided by libstdc++'s introselect, so a library built with another libstdc++ could in principle order
     *  them differently.  Costs the overlap of the library's asynchronous build. */
    bool kd_from_host_toolchain = false;

  private:
    struct Mirror {
        Map* map = nullptr;
        MirrorMode mode = MirrorMode::Hooks;
        std::unordered_map<const MapPoint*, int32_t> slot;      // live points
        std::vector<MapPoint*> slot_point;                      // by slot; nullptr = tombstone
        std::unordered_map<const KeyFrame*, int32_t> observer;  // key frame -> observer id
        std::vector<const KeyFrame*> observer_frame;
        std::vector<float> observer_center;                     // [id x 3] as last uploaded
        size_t tombstones = 0;
        // what was last uploaded per slot (Scan mode compares against it; audit() in both modes)
        std::vector<float> pos;                                 // [slot x 3]
        std::vector<uint64_t> obs_hash;                         // [slot]
        // pending changes (Hooks mode)
        std::vector<MapPoint*> created;
        std::unordered_set<MapPoint*> obs_dirty, pos_dirty;
        std::vector<int32_t> removed_slots;
        std::unordered_set<const KeyFrame*> pose_dirty;
    };

    std::vector<MapPointMatch> project(const Frame& frame, const std::vector<MapPoint*>& visit,
                                       KeyFrame* required_observer, bool replace) const;
    std::vector<MapPointMatch> project_resident(const Frame& frame, const std::vector<int32_t>* order,
                                                bool replace) const;
    int upload_frame(const Frame& frame) const;
    void upload_points(const std::vector<MapPoint*>& visit, std::vector<MapPoint*>& uniq,
                       std::vector<int32_t>& order) const;
    slamgpu_proj_params params(bool replace) const;

    // mirror plumbing
    void scan_for_changes() const;
    void flush() const;
    int32_t observer_id(const KeyFrame* key_frame, std::vector<int32_t>& new_ids, std::vector<float>& new_centers) const;
    void flatten_runs(const std::vector<MapPoint*>& points, std::vector<int32_t>& obs_offset, std::vector<uint8_t>& obs_desc,
                      std::vector<int32_t>& obs_observer, std::vector<int32_t>& new_ids, std::vector<float>& new_centers,
                      std::vector<uint64_t>* hashes) const;
    std::vector<int32_t> matched_slots(const Frame& frame) const;
    std::vector<int32_t> observed_slots(KeyFrame* key_frame) const;

    const Camera& m_camera;
    float m_max_descriptor_distance;
    slamgpu_ctx* m_ctx = nullptr;
    mutable Mirror m_mirror;
    mutable MirrorStats m_stats;
};

} // namespace slam
