// voc_kernels.cu -- bag-of-words kernels for sm_100a.
//
// Replaces, for LoopDetector's retrieval step (src/LoopDetector.cpp:85-92,346-373):
//   V1  voc_transform_kernel   DBoW2 TemplatedVocabulary::transform(feature, word, weight)
//                              (thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1257): descend the
//                              vocabulary tree, at every level the FIRST child with the strictly
//                              smallest FORB::distance (FORB.cpp:81-101 = 256-bit Hamming)
//   V2  voc_bow_count_kernel   the integer half of transform(features, bow) (:1064-1120,
//                              BowVector.cpp:33-46): the image's words sorted ascending (the std::map
//                              order) and run-length encoded -> (word id, count)
//   V3  voc_score_l1_kernel    L1Scoring::score (ScoringObject.cpp:23-68) of one query vector against
//                              every vector of a database, terms summed in ascending word order
//
// Roofline: V1 reads 32 B per child per level from an L2-resident tree (35 MB for the k = 10, L = 6
// ORB vocabulary): 2 dependent loads per level, latency-bound for one image, L2-bandwidth-bound for
// batches.  Algorithmic bytes per descriptor = 32 + sum over levels of (16 + 32 * children).
#include <climits>

#include "slamgpu_internal.cuh"

namespace slamgpu {

namespace {

constexpr int kVocThreads = 128;

__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// One group of W lanes per descriptor, 32 / W groups per warp (W = 10 for the usual k = 10: three
// descriptors per warp, 30 of 32 lanes busy; a power-of-two group of 16 would leave 6 of every 16 idle).
// State = the child-array position of the current node; one level = rec[position] -> the node's children,
// contiguous in cdesc, one child per lane -> packed key (distance << 20 | child order), minimum over the
// group = the reference's "first strictly smaller".
template <int W>
__global__ void __launch_bounds__(kVocThreads) voc_transform_kernel(const int4* __restrict__ rec,
                                                                    const uint4* __restrict__ cdesc,
                                                                    const double* __restrict__ cweight, int root_begin,
                                                                    int root_count, const uint4* __restrict__ desc, int n,
                                                                    int32_t* __restrict__ out_word,
                                                                    double* __restrict__ out_weight)
{
    constexpr int G = 32 / W;                                    // groups per warp
    const int wl = threadIdx.x & 31;
    const int g = wl / W;
    if (g >= G) return;                                          // the warp's spare lanes (32 - G * W)
    const int lane = wl - g * W;
    const unsigned gmask = (W == 32 ? 0xFFFFFFFFu : ((1u << W) - 1u)) << (g * W);
    const int warp = blockIdx.x * (kVocThreads / 32) + (threadIdx.x >> 5);
    const int i = warp * G + g;
    if (i >= n) return;
    const uint4 q0 = __ldg(&desc[2 * static_cast<size_t>(i)]), q1 = __ldg(&desc[2 * static_cast<size_t>(i) + 1]);
    int begin = root_begin, count = root_count, pos = -1, word = -1;
    while (count > 0) {
        unsigned best = 0xFFFFFFFFu;
        for (int c = lane; c < count; c += W) {                  // TemplatedVocabulary.h:1237-1247
            const uint4 b0 = __ldg(&cdesc[2 * static_cast<size_t>(begin + c)]);
            const uint4 b1 = __ldg(&cdesc[2 * static_cast<size_t>(begin + c) + 1]);
            const unsigned key = (static_cast<unsigned>(hamming256(q0, q1, b0, b1)) << 20) | static_cast<unsigned>(c);
            best = min(best, key);
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {                 // minimum into the group's first lane
            if (off < W) {
                const unsigned o = __shfl_down_sync(gmask, best, off);
                if (lane + off < W) best = min(best, o);
            }
        }
        best = __shfl_sync(gmask, best, g * W);
        pos = begin + static_cast<int>(best & 0xFFFFFu);
        const int4 r = __ldg(&rec[pos]);
        begin = r.x;
        count = r.y;
        word = r.z;
    }
    if (lane == 0) {
        out_word[i] = word;                                          // :1255-1256
        if (out_weight) out_weight[i] = pos >= 0 ? cweight[pos] : 0.0;
    }
}

// One CTA per image: bitonic sort of the image's words in shared memory (padded with INT_MAX to a power of
// two), then run-length encoding with a block-wide exclusive scan of the run heads.
constexpr int kBowThreads = 1024;

__global__ void __launch_bounds__(kBowThreads) voc_bow_count_kernel(const int32_t* __restrict__ word,
                                                                    const int32_t* __restrict__ image_offset,
                                                                    int32_t* __restrict__ out_ids,
                                                                    int32_t* __restrict__ out_counts,
                                                                    int32_t* __restrict__ n_distinct)
{
    extern __shared__ int32_t s_w[];                 // [cap] words, then [kBowThreads] scan scratch
    const int img = blockIdx.x;
    const int b = image_offset[img], n = image_offset[img + 1] - b;
    int cap = 1;
    while (cap < n) cap <<= 1;
    int32_t* s_scan = s_w + cap;
    for (int t = threadIdx.x; t < cap; t += kBowThreads) s_w[t] = t < n ? word[b + t] : INT_MAX;
    __syncthreads();
    for (int k = 2; k <= cap; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < (cap >> 1); i += kBowThreads) {    // one compare-exchange pair per iteration
                const int t = ((i & ~(j - 1)) << 1) | (i & (j - 1));
                const int p = t | j;
                const int32_t x = s_w[t], y = s_w[p];
                const bool up = (t & k) == 0;
                if ((x > y) == up) {
                    s_w[t] = y;
                    s_w[p] = x;
                }
            }
            __syncthreads();
        }
    }
    // run heads: position t starts a run when it differs from t - 1.  Each thread owns a contiguous chunk.
    const int chunk = (n + kBowThreads - 1) / kBowThreads;
    const int t0 = min(n, static_cast<int>(threadIdx.x) * chunk), t1 = min(n, t0 + chunk);
    int heads = 0;
    for (int t = t0; t < t1; ++t) heads += (t == 0 || s_w[t] != s_w[t - 1]) ? 1 : 0;
    s_scan[threadIdx.x] = heads;
    __syncthreads();
    for (int d = 1; d < kBowThreads; d <<= 1) {      // Hillis-Steele inclusive scan
        const int v = threadIdx.x >= d ? s_scan[threadIdx.x - d] : 0;
        __syncthreads();
        s_scan[threadIdx.x] += v;
        __syncthreads();
    }
    int run = s_scan[threadIdx.x] - heads;           // runs that start before this chunk
    for (int t = t0; t < t1; ++t) {
        if (t == 0 || s_w[t] != s_w[t - 1]) {
            int e = t + 1;
            while (e < n && s_w[e] == s_w[t]) ++e;   // runs are short (features per word)
            out_ids[b + run] = s_w[t];
            out_counts[b + run] = e - t;
            ++run;
        }
    }
    if (threadIdx.x == kBowThreads - 1) n_distinct[img] = s_scan[kBowThreads - 1];
}

// One warp per database vector.  Lanes look their query entries up in the vector by binary search; the
// terms |vi - wi| - |vi| - |wi| of the common words are then added in ascending word order, one after the
// other, exactly as the merge loop of ScoringObject.cpp:34-59 meets them.
__global__ void __launch_bounds__(kVocThreads) voc_score_l1_kernel(const int32_t* __restrict__ q_ids,
                                                                   const double* __restrict__ q_vals, int nq,
                                                                   const int32_t* __restrict__ db_ids,
                                                                   const double* __restrict__ db_vals,
                                                                   const int32_t* __restrict__ db_offset, int n_db,
                                                                   double* __restrict__ out_score)
{
    const int lane = threadIdx.x & 31;
    const int v = blockIdx.x * (kVocThreads / 32) + (threadIdx.x >> 5);
    if (v >= n_db) return;
    const int b = db_offset[v], m = db_offset[v + 1] - b;
    double score = 0.0;
    for (int base = 0; base < nq; base += 32) {
        const int qi = base + lane;
        double term = 0.0;
        bool hit = false;
        if (qi < nq && m > 0) {
            const int id = q_ids[qi];
            int lo = 0, hi = m;                       // lower_bound (galloping from the previous chunk's position
            while (lo < hi) {                         // was measured no faster: 175 vs 161 us per 1000 vectors)
                const int mid = (lo + hi) >> 1;
                if (db_ids[b + mid] < id) lo = mid + 1; else hi = mid;
            }
            if (lo < m && db_ids[b + lo] == id) {
                const double vi = q_vals[qi], wi = db_vals[b + lo];
                term = fabs(vi - wi) - fabs(vi) - fabs(wi);      // ScoringObject.cpp:41
                hit = true;
            }
        }
        unsigned bits = __ballot_sync(0xFFFFFFFFu, hit);
        while (bits) {                                // ascending word id = ascending lane
            const int src = __ffs(bits) - 1;
            bits &= bits - 1;
            score = __dadd_rn(score, __shfl_sync(0xFFFFFFFFu, term, src));
        }
    }
    if (lane == 0) out_score[v] = -score / 2.0;       // :65
}

} // namespace

int voc_launch_transform(slamgpu_ctx* ctx, const VocabDev& v, const uint8_t* d_desc, int n, int32_t* d_word,
                         double* d_weight)
{
    if (n <= 0) return SLAMGPU_OK;
    ProfScope prof(ctx, SLAMGPU_KERNEL_VOC_TRANSFORM);
    // group width by the widest node: 4 / 8 / 10 lanes when one pass covers every child, else 16 lanes looping
    const int k = v.max_children;
    const uint4* d4 = reinterpret_cast<const uint4*>(d_desc);
    auto grid = [&](int w) { const int per_cta = (kVocThreads / 32) * (32 / w); return (n + per_cta - 1) / per_cta; };
    if (k <= 4) {
        voc_transform_kernel<4><<<grid(4), kVocThreads, 0, ctx->stream>>>(v.rec, v.cdesc, v.cweight, v.root_begin, v.root_count, d4, n, d_word, d_weight);
    } else if (k <= 8) {
        voc_transform_kernel<8><<<grid(8), kVocThreads, 0, ctx->stream>>>(v.rec, v.cdesc, v.cweight, v.root_begin, v.root_count, d4, n, d_word, d_weight);
    } else if (k <= 10) {
        voc_transform_kernel<10><<<grid(10), kVocThreads, 0, ctx->stream>>>(v.rec, v.cdesc, v.cweight, v.root_begin, v.root_count, d4, n, d_word, d_weight);
    } else {
        voc_transform_kernel<16><<<grid(16), kVocThreads, 0, ctx->stream>>>(v.rec, v.cdesc, v.cweight, v.root_begin, v.root_count, d4, n, d_word, d_weight);
    }
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int voc_launch_bow_count(slamgpu_ctx* ctx, const int32_t* d_word, const int32_t* d_image_offset, int n_images,
                         int32_t* d_ids, int32_t* d_counts, int32_t* d_n_distinct)
{
    if (n_images <= 0) return SLAMGPU_OK;
    const size_t smem = (SLAMGPU_BOW_MAX_FEATURES + kBowThreads) * sizeof(int32_t);
    if (!ctx->bow_smem_set) {     // per context = per device
        SG_CUDA(cudaFuncSetAttribute(voc_bow_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        ctx->bow_smem_set = true;
    }
    voc_bow_count_kernel<<<n_images, kBowThreads, smem, ctx->stream>>>(d_word, d_image_offset, d_ids, d_counts, d_n_distinct);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int voc_launch_score_l1(slamgpu_ctx* ctx, const int32_t* d_q_ids, const double* d_q_vals, int nq,
                        const int32_t* d_db_ids, const double* d_db_vals, const int32_t* d_db_offset, int n_db,
                        double* d_score)
{
    if (n_db <= 0) return SLAMGPU_OK;
    voc_score_l1_kernel<<<(n_db + kVocThreads / 32 - 1) / (kVocThreads / 32), kVocThreads, 0, ctx->stream>>>(
        d_q_ids, d_q_vals, nq, d_db_ids, d_db_vals, d_db_offset, n_db, d_score);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

} // namespace slamgpu
