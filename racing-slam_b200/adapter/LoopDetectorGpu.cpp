// LoopDetectorGpu.cpp -- see LoopDetectorGpu.h.  Citations are relative to the reference tree.
#include "LoopDetectorGpu.h"

#include <stdexcept>
#include <string>

namespace slam {

namespace {

void check(int rc, const char* what)
{
    if (rc != SLAMGPU_OK) {
        throw std::runtime_error(std::string("libslamgpu: ") + what + ": " + slamgpu_last_error());
    }
}

constexpr int VOCABULARY_SLOT = 0;
constexpr int BOW_DATABASE_SLOT = 0;

void flatten(const DBoW2::BowVector& bow, std::vector<int32_t>& ids, std::vector<double>& values)
{
    ids.clear();
    values.clear();
    ids.reserve(bow.size());
    values.reserve(bow.size());
    for (const auto& [word, value] : bow) {                     // std::map: ascending word id, as the ABI expects
        ids.push_back(static_cast<int32_t>(word));
        values.push_back(value);
    }
}

} // namespace

LoopDetectorGpu::LoopDetectorGpu(slamgpu_ctx* ctx, int device) : m_ctx(ctx)
{
    if (m_ctx == nullptr) {
        check(slamgpu_create(device, &m_ctx), "slamgpu_create");
        m_owns_ctx = true;
    }
}

LoopDetectorGpu::~LoopDetectorGpu()
{
    if (m_owns_ctx) {
        slamgpu_destroy(m_ctx);
    } else {
        if (m_has_vocabulary) {
            slamgpu_vocab_release(m_ctx, VOCABULARY_SLOT);
        }
        if (m_n_key_frames > 0) {
            slamgpu_bowdb_release(m_ctx, BOW_DATABASE_SLOT);
        }
    }
}

void LoopDetectorGpu::set_vocabulary(int n_nodes, const int32_t* child_offset, const int32_t* child_ids,
                                     const uint8_t* node_descriptors, const double* node_weights,
                                     const int32_t* node_word_ids)
{
    check(slamgpu_vocab_upload(m_ctx, VOCABULARY_SLOT, n_nodes, child_offset, child_ids, node_descriptors, node_weights,
                               node_word_ids),
          "slamgpu_vocab_upload");
    m_has_vocabulary = true;
}

DBoW2::BowVector LoopDetectorGpu::bow_of(const ExtractedFeatures& features) const
{
    DBoW2::BowVector bow;
    const int n = features.descriptors.rows;
    if (n == 0) {
        return bow;                                             // src/LoopDetector.cpp:88
    }
    if (!m_has_vocabulary) {
        throw std::logic_error("LoopDetectorGpu::bow_of before set_vocabulary");
    }
    std::vector<uint8_t> rows;                                  // descriptor_rows(), :75-83
    rows.reserve(32 * static_cast<size_t>(n));
    for (int r = 0; r < n; r++) {
        const uint8_t* p = features.descriptors.ptr<uint8_t>(r);
        rows.insert(rows.end(), p, p + 32);
    }
    const int32_t image_offset[2] = {0, n};
    int32_t out_offset[2] = {0, 0};
    std::vector<int32_t> ids(static_cast<size_t>(n));
    std::vector<double> values(static_cast<size_t>(n));
    check(slamgpu_vocab_bow(m_ctx, VOCABULARY_SLOT, rows.data(), image_offset, 1, SLAMGPU_BOW_TF_IDF, SLAMGPU_BOW_NORM_L1,
                            ids.data(), values.data(), out_offset),
          "slamgpu_vocab_bow");
    for (int32_t i = 0; i < out_offset[1]; i++) {               // ascending word id: the end hint makes it linear
        bow.insert(bow.end(), {static_cast<DBoW2::WordId>(ids[static_cast<size_t>(i)]), values[static_cast<size_t>(i)]});
    }
    return bow;
}

size_t LoopDetectorGpu::add_key_frame(const DBoW2::BowVector& bow)
{
    std::vector<int32_t> ids;
    std::vector<double> values;
    flatten(bow, ids, values);
    check(slamgpu_bowdb_append(m_ctx, BOW_DATABASE_SLOT, ids.data(), values.data(), static_cast<int>(ids.size())),
          "slamgpu_bowdb_append");
    return m_n_key_frames++;
}

std::vector<double> LoopDetectorGpu::score_against_all(const DBoW2::BowVector& query) const
{
    std::vector<double> scores(m_n_key_frames, 0.0);
    if (m_n_key_frames == 0) {
        return scores;
    }
    std::vector<int32_t> ids;
    std::vector<double> values;
    flatten(query, ids, values);
    check(slamgpu_bowdb_score_l1(m_ctx, BOW_DATABASE_SLOT, ids.data(), values.data(), static_cast<int>(ids.size()),
                                 scores.data()),
          "slamgpu_bowdb_score_l1");
    return scores;
}

LoopDetectorGpu::PnpScore LoopDetectorGpu::score_pnp_models(const std::vector<cv::Point3f>& object_points,
                                                            const std::vector<cv::Point2f>& image_points,
                                                            const std::vector<std::array<double, 12>>& models, double fx,
                                                            double fy, double cx, double cy, double reprojection_error,
                                                            double confidence, int model_points, int max_iterations) const
{
    PnpScore result;
    const size_t n = object_points.size();
    if (n != image_points.size()) {
        throw std::invalid_argument("LoopDetectorGpu::score_pnp_models: object / image point counts differ");
    }
    result.inlier_counts.assign(models.size(), 0);
    result.inlier_mask.assign(n, 0);
    if (n == 0 || models.empty()) {
        return result;
    }
    std::vector<float> object(3 * n), image(2 * n);
    for (size_t i = 0; i < n; i++) {                            // src/LoopDetector.cpp:192-196
        object[3 * i] = object_points[i].x;
        object[3 * i + 1] = object_points[i].y;
        object[3 * i + 2] = object_points[i].z;
        image[2 * i] = image_points[i].x;
        image[2 * i + 1] = image_points[i].y;
    }
    const int32_t corr_offset[2] = {0, static_cast<int32_t>(n)};
    const int32_t hyp_offset[2] = {0, static_cast<int32_t>(models.size())};
    const double K4[4] = {fx, fy, cx, cy};                      // intrinsics converted to CV_64F, :199-200
    int32_t best = -1, inliers = 0;
    check(slamgpu_pnp_ransac_score(m_ctx, object.data(), image.data(), corr_offset, 1, models[0].data(), hyp_offset, K4,
                                   reprojection_error, confidence, model_points, max_iterations,
                                   result.inlier_counts.data(), &best, &inliers, result.inlier_mask.data()),
          "slamgpu_pnp_ransac_score");
    result.best = best;
    result.inliers = inliers;
    return result;
}

} // namespace slam
