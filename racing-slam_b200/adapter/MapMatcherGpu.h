// MapMatcherGpu.h -- drop-in for slam::MapMatcher (reference src/MapMatcher.h:16-38) on libslamgpu.
//
// Same constructor and the same four entry points with identical signatures and return types, so
// the three owners (src/Tracker.cpp:42, src/Mapper.cpp:87, src/LoopDetector.cpp:319) only change
// the member type.  Compiled inside the reference tree (include path = its src/): it uses only the
// read accessors SURVEY 8b lists.  All matching arithmetic runs on the GPU through the C ABI
// (include/slamgpu.h); this class flattens the object graph to POD arrays and maps indices back.
#pragma once

#include <cstdint>
#include <utility>
#include <vector>

#include "features/FeatureExtractor.h"   // slam::MapPointMatch, slam::FeatureMatch, slam::ExtractedFeatures
#include "slamgpu.h"

namespace slam {

class Camera;
class Frame;
class KeyFrame;
class Map;
class MapPoint;

class MapMatcherGpu {
  public:
    MapMatcherGpu(const Camera& camera, float max_descriptor_distance, cv::NormTypes norm_type, int device = 0);
    ~MapMatcherGpu();
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

    // Matching constants of src/MapMatcher.cpp:12-18, runtime-settable (defaults = the reference's)
    float search_radius = 20.0F;
    float min_viewing_angle_cosine = 0.5F;
    float max_nearer_ratio = 2.0F;
    float max_further_ratio = 1.25F;
    float match_ratio = 0.75F;

  private:
    std::vector<MapPointMatch> project(const Frame& frame, const std::vector<MapPoint*>& visit,
                                       KeyFrame* required_observer, bool replace) const;
    int upload_frame(const Frame& frame) const;
    void upload_points(const std::vector<MapPoint*>& visit, std::vector<MapPoint*>& uniq,
                       std::vector<int32_t>& order) const;
    slamgpu_proj_params params(bool replace) const;

    const Camera& m_camera;
    float m_max_descriptor_distance;
    slamgpu_ctx* m_ctx = nullptr;
};

} // namespace slam
