// extern "C" handle around the REFERENCE's vendored DBoW2 (thirdparty/DBoW2, compiled unmodified by
// oracle/Makefile against oracle/shim/opencv2).  Test infrastructure only: it builds ORB vocabularies with
// DBoW2's own create(), exports their trees as plain arrays and runs DBoW2's own transform() / score(), to
// pin oracle/slam_oracle.cpp's restatement (src/LoopDetector.cpp:85-92,369).
#include <cstdint>
#include <cstring>
#include <vector>

#include "DBoW2/FORB.h"
#include "DBoW2/TemplatedVocabulary.h"

namespace {

using Base = DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB>;   // src/LoopDetector.cpp:26

struct Voc : Base {
    using Base::Base;
    using Base::m_nodes;
    using Base::m_words;
    using Base::transform;    // the per-feature overload is protected
};

cv::Mat row_of(const uint8_t* d)
{
    cv::Mat m(1, 32, CV_8U);
    std::memcpy(m.ptr<unsigned char>(), d, 32);
    return m;
}

void fill_bow(DBoW2::BowVector& v, const int32_t* ids, const double* vals, int n)
{
    for (int i = 0; i < n; i++) v.insert(v.end(), DBoW2::BowVector::value_type(static_cast<DBoW2::WordId>(ids[i]), vals[i]));
}

} // namespace

extern "C" {

// DBoW2's own k-means++ tree over training images; weighting / scoring are DBoW2's enums
// (TF_IDF = 0, TF, IDF, BINARY; L1_NORM = 0, ...).
void* ref_voc_create(int k, int L, int weighting, int scoring, const uint8_t* desc, const int32_t* image_offset,
                     int n_images)
{
    std::vector<std::vector<cv::Mat>> training(n_images);
    for (int i = 0; i < n_images; i++)
        for (int r = image_offset[i]; r < image_offset[i + 1]; r++) training[i].push_back(row_of(desc + 32 * static_cast<size_t>(r)));
    auto* voc = new Voc(k, L, static_cast<DBoW2::WeightingType>(weighting), static_cast<DBoW2::ScoringType>(scoring));
    voc->create(training);
    return voc;
}

void ref_voc_destroy(void* h) { delete static_cast<Voc*>(h); }

int ref_voc_num_nodes(void* h) { return static_cast<int>(static_cast<Voc*>(h)->m_nodes.size()); }
int ref_voc_num_words(void* h) { return static_cast<int>(static_cast<Voc*>(h)->size()); }

// The tree as arrays: children in DBoW2's own order (the order transform() scans them in).
// child_offset [n_nodes + 1], child_ids [n_nodes - 1], desc [n_nodes x 32], weight [n_nodes], word_id [n_nodes] (-1 = inner)
void ref_voc_export(void* h, int32_t* child_offset, int32_t* child_ids, uint8_t* desc, double* weight, int32_t* word_id)
{
    auto* voc = static_cast<Voc*>(h);
    int c = 0;
    for (size_t i = 0; i < voc->m_nodes.size(); i++) {
        const auto& node = voc->m_nodes[i];
        child_offset[i] = c;
        for (auto id : node.children) child_ids[c++] = static_cast<int32_t>(id);
        if (!node.descriptor.empty()) std::memcpy(desc + 32 * i, node.descriptor.ptr<unsigned char>(), 32);
        else std::memset(desc + 32 * i, 0, 32);
        weight[i] = node.weight;
        word_id[i] = node.isLeaf() && i != 0 ? static_cast<int32_t>(node.word_id) : -1;
    }
    child_offset[voc->m_nodes.size()] = c;
}

// transform(feature, id, weight) per descriptor (TemplatedVocabulary.h:1218-1257)
void ref_voc_transform_each(void* h, const uint8_t* desc, int n, int32_t* out_word, double* out_weight)
{
    auto* voc = static_cast<Voc*>(h);
    for (int i = 0; i < n; i++) {
        DBoW2::WordId id = 0;
        DBoW2::WordValue w = 0;
        voc->transform(row_of(desc + 32 * static_cast<size_t>(i)), id, w);
        out_word[i] = static_cast<int32_t>(id);
        out_weight[i] = w;
    }
}

// transform(features, bow) (TemplatedVocabulary.h:1064-1120): returns the number of entries
int ref_voc_transform_bow(void* h, const uint8_t* desc, int n, int32_t* out_ids, double* out_vals, int cap)
{
    auto* voc = static_cast<Voc*>(h);
    std::vector<cv::Mat> rows;
    for (int i = 0; i < n; i++) rows.push_back(row_of(desc + 32 * static_cast<size_t>(i)));
    DBoW2::BowVector bow;
    voc->transform(rows, bow);
    int c = 0;
    for (const auto& kv : bow) {
        if (c < cap) {
            out_ids[c] = static_cast<int32_t>(kv.first);
            out_vals[c] = kv.second;
        }
        c++;
    }
    return c;
}

// vocabulary->score(a, b) (src/LoopDetector.cpp:369)
double ref_voc_score(void* h, const int32_t* ids_a, const double* vals_a, int na, const int32_t* ids_b,
                     const double* vals_b, int nb)
{
    DBoW2::BowVector a, b;
    fill_bow(a, ids_a, vals_a, na);
    fill_bow(b, ids_b, vals_b, nb);
    return static_cast<Voc*>(h)->score(a, b);
}
}
