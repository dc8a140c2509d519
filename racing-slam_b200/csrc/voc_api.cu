// voc_api.cu -- bag-of-words entry points of the C ABI (include/slamgpu.h): resident DBoW2
// vocabulary, per-feature word lookup, per-image bag-of-words vectors and L1 scoring against a database.
// Kernels are in voc_kernels.cu.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "slamgpu_internal.cuh"

using namespace slamgpu;

namespace {

#define SG_CTX(ctx)                                 \
    do {                                            \
        SG_CHECK((ctx) != nullptr, "null context"); \
        SG_CUDA(cudaSetDevice((ctx)->device));      \
    } while (0)

void free_bowdb(BowDbDev& d)
{
    if (d.ids) cudaFree(d.ids);
    if (d.vals) cudaFree(d.vals);
    if (d.offset) cudaFree(d.offset);
    d = BowDbDev();
}

void free_vocab(VocabDev& v)
{
    if (v.blob) cudaFree(v.blob);
    v = VocabDev();
}

} // namespace

namespace slamgpu {
void free_all_vocabs(slamgpu_ctx* ctx)
{
    for (auto& kv : ctx->vocabs) free_vocab(kv.second);
    ctx->vocabs.clear();
    for (auto& kv : ctx->bow_dbs) free_bowdb(kv.second);
    ctx->bow_dbs.clear();
}
} // namespace slamgpu

extern "C" {

int slamgpu_vocab_upload(slamgpu_ctx* ctx, int vocab_id, int n_nodes, const int32_t* child_offset,
                         const int32_t* child_ids, const uint8_t* desc, const double* weight, const int32_t* word_id)
{
    SG_CTX(ctx);
    SG_CHECK(n_nodes >= 1, "a vocabulary has at least its root");
    SG_CHECK(child_offset && desc && weight && word_id, "null vocabulary array");
    SG_CHECK(child_offset[0] == 0, "child_offset[0] must be 0");
    for (int i = 0; i < n_nodes; ++i) SG_CHECK(child_offset[i + 1] >= child_offset[i], "child_offset must be non-decreasing");
    const int nc = child_offset[n_nodes];
    SG_CHECK(nc == 0 || child_ids != nullptr, "null child_ids");
    SG_CHECK(nc < (1 << 30), "vocabulary too large");
    // every node except the root is the child of exactly one node (a tree)
    std::vector<uint8_t> seen(static_cast<size_t>(n_nodes), 0);
    int n_words = 0, max_children = 0;
    for (int c = 0; c < nc; ++c) {
        const int id = child_ids[c];
        SG_CHECK(id > 0 && id < n_nodes && !seen[id], "child_ids is not a tree over the nodes");
        seen[id] = 1;
    }
    for (int i = 0; i < n_nodes; ++i) {
        SG_CHECK(child_offset[i + 1] - child_offset[i] < (1 << 20), "too many children in one node");
        max_children = std::max(max_children, child_offset[i + 1] - child_offset[i]);
        if (child_offset[i + 1] == child_offset[i] && i != 0) {
            SG_CHECK(word_id[i] >= 0, "negative word id at a leaf");
            n_words = std::max(n_words, word_id[i] + 1);
        }
    }

    auto it = ctx->vocabs.find(vocab_id);
    if (it != ctx->vocabs.end()) {
        SG_CUDA(cudaStreamSynchronize(ctx->stream));
        free_vocab(it->second);
        ctx->vocabs.erase(it);
    }
    VocabDev v;
    v.n_nodes = n_nodes;
    v.n_children = nc;
    v.n_words = n_words;
    v.root_begin = 0;
    v.root_count = child_offset[1] - child_offset[0];
    v.max_children = max_children;
    const size_t C = static_cast<size_t>(nc), W = static_cast<size_t>(n_words);
    const size_t b_rec = Carver::padded(C * sizeof(int4)), b_desc = Carver::padded(C * 32),
                 b_w = Carver::padded(C * sizeof(double)), b_ww = Carver::padded(W * sizeof(double));
    const size_t total = b_rec + b_desc + b_w + b_ww;
    const cudaError_t e = cudaMalloc(&v.blob, total + 256);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(SLAMGPU_ERR_NOMEM, std::string("vocabulary allocation failed: ") + cudaGetErrorString(e));
    }
    Carver dv(v.blob);
    v.rec = dv.take<int4>(C);
    v.cdesc = dv.take<uint4>(2 * C);
    v.cweight = dv.take<double>(C);
    v.word_weight = dv.take<double>(W);
    if (total > 0) {
        int rc = ensure_host_stage(ctx, total + 256);
        if (rc != SLAMGPU_OK) {
            cudaFree(v.blob);
            return rc;
        }
        SG_CUDA(cudaStreamSynchronize(ctx->stream));
        Carver hv(ctx->h_stage.ptr);
        int4* h_rec = hv.take<int4>(C);
        uint8_t* h_desc = reinterpret_cast<uint8_t*>(hv.take<uint4>(2 * C));
        double* h_w = hv.take<double>(C);
        double* h_ww = hv.take<double>(W);
        std::memset(h_ww, 0, W * sizeof(double));
        for (size_t c = 0; c < C; ++c) {                 // re-layout by child position
            const int id = child_ids[c];
            const int cnt = child_offset[id + 1] - child_offset[id];
            h_rec[c] = make_int4(child_offset[id], cnt, cnt == 0 ? word_id[id] : -1, 0);
            std::memcpy(h_desc + 32 * c, desc + 32 * static_cast<size_t>(id), 32);
            h_w[c] = weight[id];
            if (cnt == 0) h_ww[word_id[id]] = weight[id];
        }
        v.word_weight_host.assign(h_ww, h_ww + W);
        SG_CUDA(cudaMemcpyAsync(v.blob, ctx->h_stage.ptr, total, cudaMemcpyHostToDevice, ctx->stream));
        SG_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    ctx->vocabs[vocab_id] = v;
    return SLAMGPU_OK;
}

int slamgpu_vocab_release(slamgpu_ctx* ctx, int vocab_id)
{
    SG_CTX(ctx);
    auto it = ctx->vocabs.find(vocab_id);
    if (it == ctx->vocabs.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown vocabulary id");
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    free_vocab(it->second);
    ctx->vocabs.erase(it);
    return SLAMGPU_OK;
}

int slamgpu_vocab_transform_dev(slamgpu_ctx* ctx, int vocab_id, const uint8_t* d_desc, int n, int32_t* d_out_word,
                                double* d_out_weight)
{
    SG_CTX(ctx);
    SG_CHECK(n >= 0, "negative descriptor count");
    auto it = ctx->vocabs.find(vocab_id);
    if (it == ctx->vocabs.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown vocabulary id");
    if (n == 0) return SLAMGPU_OK;
    SG_CHECK(d_desc != nullptr && d_out_word != nullptr, "null device pointer");
    return voc_launch_transform(ctx, it->second, d_desc, n, d_out_word, d_out_weight);
}

int slamgpu_vocab_transform(slamgpu_ctx* ctx, int vocab_id, const uint8_t* desc, int n, int32_t* out_word,
                            double* out_weight)
{
    SG_CTX(ctx);
    SG_CHECK(n >= 0, "negative descriptor count");
    auto it = ctx->vocabs.find(vocab_id);
    if (it == ctx->vocabs.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown vocabulary id");
    if (n == 0) return SLAMGPU_OK;
    SG_CHECK(desc != nullptr && out_word != nullptr, "null pointer");
    const size_t N = static_cast<size_t>(n);
    const size_t b_desc = Carver::padded(N * 32), b_word = Carver::padded(N * 4), b_wgt = Carver::padded(N * 8);
    SG_TRY(ensure_device_io(ctx, b_desc + b_word + b_wgt + 256));
    SG_TRY(ensure_host_stage(ctx, std::max(b_desc, b_word + b_wgt) + 256));
    Carver dio(ctx->d_io.ptr);
    uint8_t* d_desc = reinterpret_cast<uint8_t*>(dio.take<uint4>(2 * N));
    int32_t* d_word = dio.take<int32_t>(N);
    double* d_wgt = dio.take<double>(N);
    std::memcpy(ctx->h_stage.ptr, desc, N * 32);
    SG_CUDA(cudaMemcpyAsync(d_desc, ctx->h_stage.ptr, N * 32, cudaMemcpyHostToDevice, ctx->stream));
    SG_TRY(voc_launch_transform(ctx, it->second, d_desc, n, d_word, d_wgt));
    SG_CUDA(cudaMemcpyAsync(ctx->h_stage.ptr, d_word, b_word + (out_weight ? b_wgt : 0), cudaMemcpyDeviceToHost, ctx->stream));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    const uint8_t* h = static_cast<const uint8_t*>(ctx->h_stage.ptr);
    std::memcpy(out_word, h, N * 4);
    if (out_weight) std::memcpy(out_weight, h + b_word, N * 8);
    return SLAMGPU_OK;
}

int slamgpu_vocab_bow(slamgpu_ctx* ctx, int vocab_id, const uint8_t* desc, const int32_t* image_offset,
                      int n_images, int weighting, int norm, int32_t* out_ids, double* out_vals,
                      int32_t* out_offset)
{
    SG_CTX(ctx);
    SG_CHECK(n_images >= 0, "negative image count");
    SG_CHECK(weighting >= SLAMGPU_BOW_TF_IDF && weighting <= SLAMGPU_BOW_BINARY, "unknown weighting");
    SG_CHECK(norm >= SLAMGPU_BOW_NORM_NONE && norm <= SLAMGPU_BOW_NORM_L2, "unknown norm");
    auto it = ctx->vocabs.find(vocab_id);
    if (it == ctx->vocabs.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown vocabulary id");
    const VocabDev& v = it->second;
    SG_CHECK(out_offset != nullptr, "null out_offset");
    out_offset[0] = 0;
    if (n_images == 0) return SLAMGPU_OK;
    SG_CHECK(image_offset != nullptr && image_offset[0] == 0, "image_offset[0] must be 0");
    for (int i = 0; i < n_images; ++i) {
        const int len = image_offset[i + 1] - image_offset[i];
        SG_CHECK(len >= 0, "image_offset must be non-decreasing");
        SG_CHECK(len <= SLAMGPU_BOW_MAX_FEATURES, "more than SLAMGPU_BOW_MAX_FEATURES descriptors in one image");
    }
    const int n = image_offset[n_images];
    if (n == 0 || v.n_children == 0) {      // TemplatedVocabulary.h:1069-1072: empty vocabulary -> empty vectors
        for (int i = 0; i < n_images; ++i) out_offset[i + 1] = 0;
        return SLAMGPU_OK;
    }
    SG_CHECK(desc != nullptr && out_ids != nullptr && out_vals != nullptr, "null pointer");

    const size_t N = static_cast<size_t>(n), M = static_cast<size_t>(n_images);
    const size_t b_desc = Carver::padded(N * 32), b_off = Carver::padded((M + 1) * 4), b_word = Carver::padded(N * 4),
                 b_ids = Carver::padded(N * 4), b_cnt = Carver::padded(N * 4), b_nd = Carver::padded(M * 4);
    SG_TRY(ensure_device_io(ctx, b_desc + b_off + b_word + b_ids + b_cnt + b_nd + 256));
    SG_TRY(ensure_host_stage(ctx, std::max(b_desc + b_off, b_ids + b_cnt + b_nd) + 256));
    Carver dio(ctx->d_io.ptr), hio(ctx->h_stage.ptr);
    uint8_t* d_desc = reinterpret_cast<uint8_t*>(dio.take<uint4>(2 * N));
    int32_t* d_off = dio.take<int32_t>(M + 1);
    int32_t* d_word = dio.take<int32_t>(N);
    int32_t* d_ids = dio.take<int32_t>(N);
    int32_t* d_cnt = dio.take<int32_t>(N);
    int32_t* d_nd = dio.take<int32_t>(M);
    uint8_t* h_desc = reinterpret_cast<uint8_t*>(hio.take<uint4>(2 * N));
    int32_t* h_off = hio.take<int32_t>(M + 1);
    std::memcpy(h_desc, desc, N * 32);
    std::memcpy(h_off, image_offset, (M + 1) * 4);
    SG_CUDA(cudaMemcpyAsync(d_desc, ctx->h_stage.ptr, b_desc + b_off, cudaMemcpyHostToDevice, ctx->stream));
    SG_TRY(voc_launch_transform(ctx, v, d_desc, n, d_word, nullptr));
    SG_TRY(voc_launch_bow_count(ctx, d_word, d_off, n_images, d_ids, d_cnt, d_nd));
    SG_CUDA(cudaMemcpyAsync(ctx->h_stage.ptr, d_ids, b_ids + b_cnt + b_nd, cudaMemcpyDeviceToHost, ctx->stream));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    const uint8_t* hb = static_cast<const uint8_t*>(ctx->h_stage.ptr);
    const int32_t* h_ids = reinterpret_cast<const int32_t*>(hb);
    const int32_t* h_cnt = reinterpret_cast<const int32_t*>(hb + b_ids);
    const int32_t* h_nd = reinterpret_cast<const int32_t*>(hb + b_ids + b_cnt);

    // The double-precision half, in DBoW2's order.  TF / TF_IDF: the first feature of a word inserts its weight,
    // every further one adds it (BowVector::addWeight); IDF / BINARY keep the first (addIfNotExist).  Words whose
    // weight is not positive are "stopped" (TemplatedVocabulary.h:1092,1115).
    const bool accumulate = weighting == SLAMGPU_BOW_TF_IDF || weighting == SLAMGPU_BOW_TF;
    const bool must = norm != SLAMGPU_BOW_NORM_NONE;
    int o = 0;
    for (int i = 0; i < n_images; ++i) {
        const int b = image_offset[i], first = o;
        for (int r = 0; r < h_nd[i]; ++r) {
            const int id = h_ids[b + r];
            const double w = (id >= 0 && id < v.n_words) ? v.word_weight_host[id] : 0.0;
            if (!(w > 0)) continue;
            double val = w;
            if (accumulate)
                for (int c = 1; c < h_cnt[b + r]; ++c) val += w;
            out_ids[o] = id;
            out_vals[o] = val;
            ++o;
        }
        const int cnt = o - first;
        if (accumulate && cnt > 0 && !must) {            // :1096-1102
            const double nd = cnt;
            for (int k = first; k < o; ++k) out_vals[k] /= nd;
        }
        if (must) {                                      // BowVector::normalize, BowVector.cpp:62-84
            double s = 0.0;
            if (norm == SLAMGPU_BOW_NORM_L1) {
                for (int k = first; k < o; ++k) s += std::fabs(out_vals[k]);
            } else {
                for (int k = first; k < o; ++k) s += out_vals[k] * out_vals[k];
                s = std::sqrt(s);
            }
            if (s > 0.0)
                for (int k = first; k < o; ++k) out_vals[k] /= s;
        }
        out_offset[i + 1] = o;
    }
    return SLAMGPU_OK;
}

int slamgpu_vocab_score_l1(slamgpu_ctx* ctx, const int32_t* q_ids, const double* q_vals, int nq,
                           const int32_t* db_ids, const double* db_vals, const int32_t* db_offset, int n_db,
                           double* out_score)
{
    SG_CTX(ctx);
    SG_CHECK(nq >= 0 && n_db >= 0, "negative size");
    if (n_db == 0) return SLAMGPU_OK;
    SG_CHECK(db_offset != nullptr && out_score != nullptr && db_offset[0] == 0, "null database offsets / output");
    for (int i = 0; i < n_db; ++i) SG_CHECK(db_offset[i + 1] >= db_offset[i], "db_offset must be non-decreasing");
    const int total = db_offset[n_db];
    SG_CHECK((nq == 0 || (q_ids && q_vals)) && (total == 0 || (db_ids && db_vals)), "null vector arrays");
    for (int i = 1; i < nq; ++i) SG_CHECK(q_ids[i] > q_ids[i - 1], "query ids must be strictly ascending");
    const size_t Q = static_cast<size_t>(nq), T = static_cast<size_t>(total), D = static_cast<size_t>(n_db);
    const size_t b_qi = Carver::padded(Q * 4), b_qv = Carver::padded(Q * 8), b_di = Carver::padded(T * 4),
                 b_dv = Carver::padded(T * 8), b_do = Carver::padded((D + 1) * 4), b_s = Carver::padded(D * 8);
    const size_t up = b_qi + b_qv + b_di + b_dv + b_do;
    SG_TRY(ensure_device_io(ctx, up + b_s + 256));
    SG_TRY(ensure_host_stage(ctx, std::max(up, b_s) + 256));
    Carver dio(ctx->d_io.ptr), hio(ctx->h_stage.ptr);
    int32_t* d_qi = dio.take<int32_t>(Q);
    double* d_qv = dio.take<double>(Q);
    int32_t* d_di = dio.take<int32_t>(T);
    double* d_dv = dio.take<double>(T);
    int32_t* d_do = dio.take<int32_t>(D + 1);
    double* d_s = dio.take<double>(D);
    int32_t* h_qi = hio.take<int32_t>(Q);
    double* h_qv = hio.take<double>(Q);
    int32_t* h_di = hio.take<int32_t>(T);
    double* h_dv = hio.take<double>(T);
    int32_t* h_do = hio.take<int32_t>(D + 1);
    if (Q) {
        std::memcpy(h_qi, q_ids, Q * 4);
        std::memcpy(h_qv, q_vals, Q * 8);
    }
    if (T) {
        std::memcpy(h_di, db_ids, T * 4);
        std::memcpy(h_dv, db_vals, T * 8);
    }
    std::memcpy(h_do, db_offset, (D + 1) * 4);
    SG_CUDA(cudaMemcpyAsync(ctx->d_io.ptr, ctx->h_stage.ptr, up, cudaMemcpyHostToDevice, ctx->stream));
    SG_TRY(voc_launch_score_l1(ctx, d_qi, d_qv, nq, d_di, d_dv, d_do, n_db, d_s));
    SG_CUDA(cudaMemcpyAsync(ctx->h_stage.ptr, d_s, D * 8, cudaMemcpyDeviceToHost, ctx->stream));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    std::memcpy(out_score, ctx->h_stage.ptr, D * 8);
    return SLAMGPU_OK;
}

int slamgpu_bowdb_append(slamgpu_ctx* ctx, int db_id, const int32_t* ids, const double* vals, int n)
{
    SG_CTX(ctx);
    SG_CHECK(n >= 0, "negative vector length");
    SG_CHECK(n == 0 || (ids != nullptr && vals != nullptr), "null vector arrays");
    for (int i = 1; i < n; ++i) SG_CHECK(ids[i] > ids[i - 1], "vector ids must be strictly ascending");
    BowDbDev& d = ctx->bow_dbs[db_id];
    SG_CHECK(d.n_entries + n <= 0x7FFFFFFF, "bag-of-words database too large");
    // grow geometrically; the live part moves device-to-device
    if (d.n_vectors + 1 > d.cap_vectors) {
        const int cap = std::max(1024, 2 * d.cap_vectors);
        int32_t* p = nullptr;
        SG_CUDA(cudaMalloc(&p, (static_cast<size_t>(cap) + 1) * sizeof(int32_t)));
        if (d.offset) {
            SG_CUDA(cudaMemcpyAsync(p, d.offset, (static_cast<size_t>(d.n_vectors) + 1) * 4, cudaMemcpyDeviceToDevice, ctx->stream));
            SG_CUDA(cudaStreamSynchronize(ctx->stream));
            cudaFree(d.offset);
        } else {
            SG_CUDA(cudaMemsetAsync(p, 0, 4, ctx->stream));
        }
        d.offset = p;
        d.cap_vectors = cap;
    }
    if (d.n_entries + n > d.cap_entries) {
        const int64_t cap = std::max<int64_t>(std::max<int64_t>(1 << 20, d.n_entries + n), 2 * d.cap_entries);
        int32_t* pi = nullptr;
        double* pv = nullptr;
        SG_CUDA(cudaMalloc(&pi, static_cast<size_t>(cap) * 4));
        const cudaError_t e = cudaMalloc(&pv, static_cast<size_t>(cap) * 8);
        if (e != cudaSuccess) {
            cudaFree(pi);
            cudaGetLastError();
            return fail(SLAMGPU_ERR_NOMEM, std::string("bag-of-words database growth failed: ") + cudaGetErrorString(e));
        }
        if (d.n_entries) {
            SG_CUDA(cudaMemcpyAsync(pi, d.ids, static_cast<size_t>(d.n_entries) * 4, cudaMemcpyDeviceToDevice, ctx->stream));
            SG_CUDA(cudaMemcpyAsync(pv, d.vals, static_cast<size_t>(d.n_entries) * 8, cudaMemcpyDeviceToDevice, ctx->stream));
            SG_CUDA(cudaStreamSynchronize(ctx->stream));
        }
        if (d.ids) cudaFree(d.ids);
        if (d.vals) cudaFree(d.vals);
        d.ids = pi;
        d.vals = pv;
        d.cap_entries = cap;
    }
    const size_t N = static_cast<size_t>(n);
    const size_t b_i = Carver::padded(N * 4), b_v = Carver::padded(N * 8);
    SG_TRY(ensure_host_stage(ctx, b_i + b_v + 256));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    Carver hio(ctx->h_stage.ptr);
    int32_t* h_i = hio.take<int32_t>(N);
    double* h_v = hio.take<double>(N);
    int32_t* h_end = hio.take<int32_t>(1);
    if (n) {
        std::memcpy(h_i, ids, N * 4);
        std::memcpy(h_v, vals, N * 8);
        SG_CUDA(cudaMemcpyAsync(d.ids + d.n_entries, h_i, N * 4, cudaMemcpyHostToDevice, ctx->stream));
        SG_CUDA(cudaMemcpyAsync(d.vals + d.n_entries, h_v, N * 8, cudaMemcpyHostToDevice, ctx->stream));
    }
    *h_end = static_cast<int32_t>(d.n_entries + n);
    SG_CUDA(cudaMemcpyAsync(d.offset + d.n_vectors + 1, h_end, 4, cudaMemcpyHostToDevice, ctx->stream));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    d.n_vectors++;
    d.n_entries += n;
    return SLAMGPU_OK;
}

int slamgpu_bowdb_count(slamgpu_ctx* ctx, int db_id, int* n_vectors)
{
    SG_CHECK(ctx != nullptr && n_vectors != nullptr, "null pointer");
    auto it = ctx->bow_dbs.find(db_id);
    *n_vectors = it == ctx->bow_dbs.end() ? 0 : it->second.n_vectors;
    return SLAMGPU_OK;
}

int slamgpu_bowdb_score_l1(slamgpu_ctx* ctx, int db_id, const int32_t* q_ids, const double* q_vals, int nq,
                           double* out_score)
{
    SG_CTX(ctx);
    SG_CHECK(nq >= 0, "negative vector length");
    auto it = ctx->bow_dbs.find(db_id);
    if (it == ctx->bow_dbs.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown bag-of-words database id");
    const BowDbDev& d = it->second;
    if (d.n_vectors == 0) return SLAMGPU_OK;
    SG_CHECK(out_score != nullptr && (nq == 0 || (q_ids && q_vals)), "null pointer");
    for (int i = 1; i < nq; ++i) SG_CHECK(q_ids[i] > q_ids[i - 1], "query ids must be strictly ascending");
    const size_t Q = static_cast<size_t>(nq), D = static_cast<size_t>(d.n_vectors);
    const size_t b_qi = Carver::padded(Q * 4), b_qv = Carver::padded(Q * 8), b_s = Carver::padded(D * 8);
    SG_TRY(ensure_device_io(ctx, b_qi + b_qv + b_s + 256));
    SG_TRY(ensure_host_stage(ctx, std::max(b_qi + b_qv, b_s) + 256));
    Carver dio(ctx->d_io.ptr), hio(ctx->h_stage.ptr);
    int32_t* d_qi = dio.take<int32_t>(Q);
    double* d_qv = dio.take<double>(Q);
    double* d_s = dio.take<double>(D);
    int32_t* h_qi = hio.take<int32_t>(Q);
    double* h_qv = hio.take<double>(Q);
    if (Q) {
        std::memcpy(h_qi, q_ids, Q * 4);
        std::memcpy(h_qv, q_vals, Q * 8);
        SG_CUDA(cudaMemcpyAsync(ctx->d_io.ptr, ctx->h_stage.ptr, b_qi + b_qv, cudaMemcpyHostToDevice, ctx->stream));
    }
    SG_TRY(voc_launch_score_l1(ctx, d_qi, d_qv, nq, d.ids, d.vals, d.offset, d.n_vectors, d_s));
    SG_CUDA(cudaMemcpyAsync(ctx->h_stage.ptr, d_s, D * 8, cudaMemcpyDeviceToHost, ctx->stream));
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    std::memcpy(out_score, ctx->h_stage.ptr, D * 8);
    return SLAMGPU_OK;
}

int slamgpu_bowdb_release(slamgpu_ctx* ctx, int db_id)
{
    SG_CTX(ctx);
    auto it = ctx->bow_dbs.find(db_id);
    if (it == ctx->bow_dbs.end()) return fail(SLAMGPU_ERR_NOTFOUND, "unknown bag-of-words database id");
    SG_CUDA(cudaStreamSynchronize(ctx->stream));
    free_bowdb(it->second);
    ctx->bow_dbs.erase(it);
    return SLAMGPU_OK;
}

} // extern "C"
