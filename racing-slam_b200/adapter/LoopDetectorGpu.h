// LoopDetectorGpu.h -- the data-parallel steps of slam::LoopDetector (reference src/LoopDetector.cpp) on libslamgpu,
// with the shapes of the functions they replace, so that LoopDetector::Impl only swaps calls:
//   bow_of(vocabulary, features)              src/LoopDetector.cpp:85-92    -> LoopDetectorGpu::bow_of
//   Impl::bows + vocabulary->score(a, b)      :314-330, :346-373           -> add_key_frame / score_against_all
//   verify_pnp's match_descriptors per candidate  :181                      -> MapMatcherGpu::match_descriptors_batch
//   verify_pnp's cv::solvePnPRansac scoring   :203-214                      -> score_pnp_models
// DBoW2 keeps loading the vocabulary and OpenCV keeps the minimal-set sampling and the EPnP solver; what moves is
// the word lookup of every descriptor, the L1 score against every stored key frame and the inlier count of every
// RANSAC model.  Word ids, vector values and scores are bit-identical to DBoW2's; inlier counts and masks to
// cv::solvePnPRansac's error function (tests/test_adapter.py::test_loop_detector_helper_vs_oracle).
#pragma once

#include <array>
#include <cstdint>
#include <vector>

#include "DBoW2/BowVector.h"
#include "features/FeatureExtractor.h"
#include "slamgpu.h"

namespace slam {

class LoopDetectorGpu {
  public:
    /** ctx: the context of the owner's MapMatcherGpu (MapMatcherGpu::context()) or nullptr to create one on `device`. */
    explicit LoopDetectorGpu(slamgpu_ctx* ctx = nullptr, int device = 0);
    ~LoopDetectorGpu();
    LoopDetectorGpu(const LoopDetectorGpu&) = delete;
    LoopDetectorGpu& operator=(const LoopDetectorGpu&) = delete;

    /** The vocabulary tree as arrays (TemplatedVocabulary::m_nodes flattened; INTEGRATION.md shows the accessor):
     *  node 0 = root; children of node i = child_ids[child_offset[i] .. child_offset[i + 1]) in DBoW2's order. */
    void set_vocabulary(int n_nodes, const int32_t* child_offset, const int32_t* child_ids, const uint8_t* node_descriptors,
                        const double* node_weights, const int32_t* node_word_ids);

    /** bow_of (src/LoopDetector.cpp:85-92): vocabulary.transform(descriptor rows, bow) with the vocabulary's
     *  TF-IDF weighting and L1 norm (what ORBvoc is built with).  Empty features -> empty vector. */
    DBoW2::BowVector bow_of(const ExtractedFeatures& features) const;

    /** Impl::bows[i] = bow (src/LoopDetector.cpp:346-352): stored resident; returns the key frame's index. */
    size_t add_key_frame(const DBoW2::BowVector& bow);
    size_t key_frames() const { return m_n_key_frames; }

    /** vocabulary->score(query, bows[i]) for EVERY stored key frame in one call (the loop of :354-371 without its
     *  gap / time filters, which stay on the host): scores[i] as the double DBoW2 returns. */
    std::vector<double> score_against_all(const DBoW2::BowVector& query) const;

    struct PnpScore {
        std::vector<int32_t> inlier_counts;   // per model, in iteration order
        int best = -1;                        // the model RANSACPointSetRegistrator::run keeps, -1 = none (`!solved`)
        int inliers = 0;
        std::vector<uint8_t> inlier_mask;     // per correspondence, of the kept model (the `inliers` output as a mask)
    };
    /** The scoring of cv::solvePnPRansac(object_points, image_points, K, no distortion, ..., 200 iterations,
     *  reprojection error 4.0, confidence 0.99, SOLVEPNP_EPNP) (src/LoopDetector.cpp:203-214) for models the caller
     *  has drawn (R row-major then t, 12 doubles each, iteration order): inlier count of every model, RANSAC's choice,
     *  the kept model's mask. */
    PnpScore score_pnp_models(const std::vector<cv::Point3f>& object_points, const std::vector<cv::Point2f>& image_points,
                              const std::vector<std::array<double, 12>>& models, double fx, double fy, double cx, double cy,
                              double reprojection_error = 4.0, double confidence = 0.99, int model_points = 5,
                              int max_iterations = 200) const;

  private:
    slamgpu_ctx* m_ctx = nullptr;
    bool m_owns_ctx = false;
    bool m_has_vocabulary = false;
    size_t m_n_key_frames = 0;
};

} // namespace slam
