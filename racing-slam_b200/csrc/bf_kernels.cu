// bf_kernels.cu -- brute-force Hamming kNN-2 / cross-check kernels for sm_100a.
//
// Replaces the arithmetic of cv::BFMatcher::knnMatch as the reference calls it
// (src/MapMatcher.cpp:147-148, src/features/FeatureExtractor.cpp:20-21):
//   K1  bf_scan_kernel     XOR + POPC distances, per-thread packed-key top-2 over a train range that
//                          is streamed through shared memory with cp.async.bulk (TMA) + mbarrier
//   K3  bf_finish_*        merge of the train-split partials + the reference's filters
//   K2  bf_mutual_kernel   cross-check mutuality from row and column partials
//
// Roofline: the integer POPC pipe (8 POPC32 per 256-bit comparison), not HBM: a descriptor
// byte is reused by every query row of the tile.
#include <mutex>

#include "slamgpu_internal.cuh"

namespace slamgpu {

namespace {

constexpr int kThreads = 256;
constexpr int kChunk = 128;                 // train descriptors per shared-memory stage (4 KB)
constexpr int kStages = 4;                  // ring depth -> 16 KB of shared memory per CTA

__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) {
    }
}
// 1-D bulk async copy global -> shared (TMA engine, SASS UBLKCP); completion is counted in
// bytes on the mbarrier.  src/dst 16-byte aligned, size a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

struct ScanArgs {
    const uint8_t* q_base;
    const uint8_t* t_base;
    const int32_t* frame_offset;  // batched: row offsets of every frame; null = single pair
    const int32_t* pairs;         // batched: [n_pairs x 2] (query_frame, train_frame)
    int nq, nt;                   // single pair sizes
    int tiles, splits, t_per_split;
    int max_rows;                 // row stride of the key array per (pair, split)
    int swap_roles;               // batched only: scan pair as (train_frame, query_frame)
    uint2* keys;
    uint32_t* col_keys;           // cross-check (single pair): per train row min over query rows of (distance << 22 | query)
};

// ---- v2 inner loop -----------------------------------------------------------------------------
// ncu on v1 (8 POPC per comparison) showed the XU pipe, where POPC issues at 16 lanes/clk/SM, 96 %
// busy with the ALU pipe at 45 % (profiles/r01_ncu_full_bf_scan_v1.txt).  v2 moves work off XU:
//   * a carry-save adder tree (LOP3 xor3 / majority) compresses the 8 XOR words to 4 words of
//     weight 1,1,2,4  ->  4 POPC instead of 8 per comparison, +8 LOP3;
//   * two query rows share one register as packed 16-bit keys (distance << 7 | index within the
//     128-row chunk), so the top-2 update is 3 VIMNMX.U16x2 per TWO comparisons; the chunk's
//     top-2 is widened to the 32-bit (distance << 22 | train index) key once per chunk.
// Ordering is unchanged: lowest distance, then lowest train index.
__device__ __forceinline__ uint32_t lop_xor(uint32_t a, uint32_t b)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, 0, 0x3c;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t lop_xor3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t lop_maj(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xe8;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t mad_u32(uint32_t a, uint32_t b, uint32_t c)   // IMAD: FMA pipe, not ALU
{
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// 256-bit Hamming distance with 4 POPC: CSA tree over the 8 XOR words.
__device__ __forceinline__ uint32_t hamming256_csa(const uint32_t (&q)[8], const uint4& a, const uint4& b)
{
    const uint32_t x0 = lop_xor(q[0], a.x), x1 = lop_xor(q[1], a.y), x2 = lop_xor(q[2], a.z), x3 = lop_xor(q[3], a.w);
    const uint32_t x4 = lop_xor(q[4], b.x), x5 = lop_xor(q[5], b.y), x6 = lop_xor(q[6], b.z), x7 = lop_xor(q[7], b.w);
    const uint32_t s0 = lop_xor3(x0, x1, x2), c0 = lop_maj(x0, x1, x2);
    const uint32_t s1 = lop_xor3(x3, x4, x5), c1 = lop_maj(x3, x4, x5);
    const uint32_t s2 = lop_xor3(s0, s1, x6), c2 = lop_maj(s0, s1, x6);   // weight 1: s2, x7
    const uint32_t t0 = lop_xor3(c0, c1, c2), d0 = lop_maj(c0, c1, c2);   // weight 2: t0; weight 4: d0
    const uint32_t w1 = mad_u32(__popc(s2), 1u, __popc(x7));
    const uint32_t w2 = mad_u32(__popc(t0), 2u, w1);
    return mad_u32(__popc(d0), 4u, w2);
}

constexpr uint32_t kKey16None = 0xFFFFu;

// one train row against RP pairs of query rows; c0/c1 = packed chunk-local top-2
template <int RP, bool COLS>
__device__ __forceinline__ void scan_one(const uint4* __restrict__ sm, int j, const uint32_t (&q)[2 * RP][8],
                                         uint32_t (&c0)[RP], uint32_t (&c1)[RP], const uint32_t (&rowid)[2 * RP],
                                         uint32_t* s_col)
{
    const uint4 a = sm[2 * j];      // warp-uniform address -> shared-memory broadcast
    const uint4 b = sm[2 * j + 1];
    const uint32_t jj = static_cast<uint32_t>(j) * 0x10001u;
    uint32_t ck = kKeyNone;
#pragma unroll
    for (int p = 0; p < RP; ++p) {
        const uint32_t dlo = hamming256_csa(q[2 * p], a, b);
        const uint32_t dhi = hamming256_csa(q[2 * p + 1], a, b);
        if constexpr (COLS) {       // column minimum for the cross-check, same pass
            ck = min(ck, min(mad_u32(dlo, 1u << kIdxBits, rowid[2 * p]), mad_u32(dhi, 1u << kIdxBits, rowid[2 * p + 1])));
        }
        const uint32_t dd = mad_u32(dhi, 65536u, dlo);
        const uint32_t key = mad_u32(dd, 128u, jj);     // (distance << 7 | j) in both halves
        const uint32_t hi = __vmaxu2(key, c0[p]);
        c0[p] = __vminu2(key, c0[p]);
        c1[p] = __vminu2(c1[p], hi);
    }
    if constexpr (COLS) {
        ck = __reduce_min_sync(0xFFFFFFFFu, ck);            // REDUX over the warp's 32 x 2RP query rows
        if ((threadIdx.x & 31) == 0) atomicMin(&s_col[j], ck);
    }
}

__device__ __forceinline__ uint32_t widen_key(uint32_t k16, uint32_t t0)
{
    return k16 == kKey16None ? kKeyNone : (((k16 >> 7) << kIdxBits) | (t0 + (k16 & 127u)));
}

__device__ __forceinline__ void merge_top2(uint32_t k0, uint32_t k1, uint32_t& b0, uint32_t& b1)
{
    const uint32_t hi = max(k0, b0);
    b0 = min(k0, b0);
    b1 = min(min(b1, hi), k1);     // k1 >= k0: it can only compete for second place
}

template <int RP, bool COLS>
__global__ void __launch_bounds__(kThreads) bf_scan_kernel(const ScanArgs a)
{
    constexpr int R = 2 * RP;      // query rows per thread
    __shared__ __align__(128) uint4 s_train[kStages][kChunk * 2];
    __shared__ __align__(8) uint64_t s_full[kStages];
    __shared__ uint32_t s_col[COLS ? 2 : 1][kChunk];   // per-chunk column minima, double-buffered

    const int tid = threadIdx.x;
    int bid = blockIdx.x;
    const int tile = bid % a.tiles;
    bid /= a.tiles;
    const int split = bid % a.splits;
    const int pair = bid / a.splits;

    const uint8_t* qp = a.q_base;
    const uint8_t* tp = a.t_base;
    int nq = a.nq, nt = a.nt;
    if (a.pairs != nullptr) {
        int fq = a.pairs[2 * pair], ft = a.pairs[2 * pair + 1];
        if (a.swap_roles) {
            const int tmp = fq; fq = ft; ft = tmp;
        }
        const int q0 = a.frame_offset[fq], t0 = a.frame_offset[ft];
        nq = a.frame_offset[fq + 1] - q0;
        nt = a.frame_offset[ft + 1] - t0;
        qp += static_cast<size_t>(q0) * 32;
        tp += static_cast<size_t>(t0) * 32;
    }
    const int row0 = tile * (kThreads * R);
    if (row0 >= nq) return;   // ragged batch: this pair has fewer query tiles (finish never reads them)

    const int t_begin = min(split * a.t_per_split, nt);
    const int t_end = min(t_begin + a.t_per_split, nt);
    const int n_chunks = (t_end - t_begin + kChunk - 1) / kChunk;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) mbar_init(&s_full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if constexpr (COLS) {
        s_col[tid >> 7][tid & (kChunk - 1)] = kKeyNone;     // kThreads = 2 x kChunk
    }
    __syncthreads();

    auto issue = [&](int c) {
        const int s = c % kStages;
        const int t0 = t_begin + c * kChunk;
        const uint32_t bytes = static_cast<uint32_t>(min(kChunk, t_end - t0)) * 32u;
        mbar_expect_tx(&s_full[s], bytes);
        bulk_g2s(&s_train[s][0], tp + static_cast<size_t>(t0) * 32, bytes, &s_full[s]);
    };
    if (tid == 0) {
        for (int c = 0; c < min(kStages - 1, n_chunks); ++c) issue(c);
    }

    // query rows live in registers: R x 8 words per thread, rows tile-strided so loads coalesce
    uint32_t q[R][8];
    uint32_t b0[R], b1[R];
    uint32_t rowid[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        // rows past the end are clamped copies of the last row: with the largest index they can
        // never beat the real row in a column minimum
        rowid[r] = row0 + r * kThreads + tid < nq ? static_cast<uint32_t>(row0 + r * kThreads + tid) : kIdxMask;
        const int row = min(row0 + r * kThreads + tid, nq - 1);
        const uint4* src = reinterpret_cast<const uint4*>(qp + static_cast<size_t>(row) * 32);
        const uint4 lo = __ldg(src), hi = __ldg(src + 1);
        q[r][0] = lo.x; q[r][1] = lo.y; q[r][2] = lo.z; q[r][3] = lo.w;
        q[r][4] = hi.x; q[r][5] = hi.y; q[r][6] = hi.z; q[r][7] = hi.w;
        b0[r] = kKeyNone;
        b1[r] = kKeyNone;
    }

    for (int c = 0; c < n_chunks; ++c) {
        // the stage refilled here was consumed in iteration c-1 (the __syncthreads below orders it)
        if (tid == 0 && c + kStages - 1 < n_chunks) issue(c + kStages - 1);
        const int s = c % kStages;
        mbar_wait(&s_full[s], static_cast<uint32_t>((c / kStages) & 1));
        const int t0 = t_begin + c * kChunk;
        const int cnt = min(kChunk, t_end - t0);
        const uint4* sm = &s_train[s][0];
        uint32_t c0[RP], c1[RP];
#pragma unroll
        for (int p = 0; p < RP; ++p) c0[p] = c1[p] = 0xFFFFFFFFu;
        uint32_t* col = s_col[COLS ? (c & 1) : 0];
        if (cnt == kChunk) {
#pragma unroll 2
            for (int j = 0; j < kChunk; ++j) scan_one<RP, COLS>(sm, j, q, c0, c1, rowid, col);
        } else {
            for (int j = 0; j < cnt; ++j) scan_one<RP, COLS>(sm, j, q, c0, c1, rowid, col);
        }
        // widen the chunk's packed top-2 and fold it into the running 32-bit keys
#pragma unroll
        for (int p = 0; p < RP; ++p) {
            merge_top2(widen_key(c0[p] & 0xFFFFu, t0), widen_key(c1[p] & 0xFFFFu, t0), b0[2 * p], b1[2 * p]);
            merge_top2(widen_key(c0[p] >> 16, t0), widen_key(c1[p] >> 16, t0), b0[2 * p + 1], b1[2 * p + 1]);
        }
        __syncthreads();
        if constexpr (COLS) {
            // flush this chunk's column minima; the buffer is next written in chunk c + 2, after the
            // __syncthreads that ends chunk c + 1
            if (tid < cnt) {
                const uint32_t v = col[tid];
                col[tid] = kKeyNone;
                if (v != kKeyNone) atomicMin(&a.col_keys[t0 + tid], v);
            }
        }
    }

    uint2* out = a.keys + (static_cast<size_t>(pair) * a.splits + split) * a.max_rows;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int row = row0 + r * kThreads + tid;
        if (row < nq) out[row] = make_uint2(b0[r], b1[r]);
    }
}

// merge the split partials of one query row
__device__ __forceinline__ uint2 merge_partials(const uint2* keys, int splits, size_t stride)
{
    uint32_t m0 = kKeyNone, m1 = kKeyNone;
    for (int s = 0; s < splits; ++s) {
        const uint2 p = keys[s * stride];
        uint32_t hi = max(p.x, m0);
        m0 = min(p.x, m0);
        m1 = min(m1, hi);
        m1 = min(m1, p.y);   // p.y >= p.x, so it can only compete for second place
    }
    return make_uint2(m0, m1);
}

// The finish kernels run as CTAs of (kFinishThreads / G rows) x (G split groups): thread (x, y) merges the partials
// y, y + G, ... of its row, the G results meet in shared memory and the thread with y == 0 holds the row's merged
// top-2.  One thread per row walking all the splits is a chain of dependent L2 round trips: a single 2000 x 2000
// pair has 63 splits, and its finish kernel took longer (11.7 us) than its scan (9.2 us).  G is chosen at launch
// (finish_groups); every thread of the CTA must call this (it synchronises).
constexpr int kFinishThreads = 256;

__device__ __forceinline__ void merge_key(uint32_t& m0, uint32_t& m1, const uint2 p)
{
    const uint32_t hi = max(p.x, m0);
    m0 = min(p.x, m0);
    m1 = min(m1, hi);
    m1 = min(m1, p.y);   // p.y >= p.x, so it can only compete for second place
}

__device__ __forceinline__ uint2 merge_partials_2d(const uint2* keys, int splits, size_t stride, bool active, uint2* s_part)
{
    uint32_t m0 = kKeyNone, m1 = kKeyNone;
    if (active) {
#pragma unroll 4
        for (int s = threadIdx.y; s < splits; s += blockDim.y) merge_key(m0, m1, keys[s * stride]);
    }
    if (blockDim.y == 1) return make_uint2(m0, m1);
    s_part[threadIdx.y * blockDim.x + threadIdx.x] = make_uint2(m0, m1);
    __syncthreads();
    if (threadIdx.y == 0) {
        for (int g = 1; g < static_cast<int>(blockDim.y); ++g) merge_key(m0, m1, s_part[g * blockDim.x + threadIdx.x]);
    }
    __syncthreads();                                          // s_part may be reused by a second merge
    return make_uint2(m0, m1);
}

__global__ void __launch_bounds__(kFinishThreads) bf_finish_raw_kernel(const uint2* keys, int splits, int max_rows, int nq,
                                                                        int32_t* idx, int32_t* dist)
{
    __shared__ uint2 s_part[kFinishThreads];
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const uint2 m = merge_partials_2d(keys + q, splits, max_rows, q < nq, s_part);
    if (q >= nq || threadIdx.y != 0) return;
    int2 i, d;
    i.x = m.x == kKeyNone ? -1 : static_cast<int>(m.x & kIdxMask);
    d.x = m.x == kKeyNone ? -1 : static_cast<int>(m.x >> kIdxBits);
    i.y = m.y == kKeyNone ? -1 : static_cast<int>(m.y & kIdxMask);
    d.y = m.y == kKeyNone ? -1 : static_cast<int>(m.y >> kIdxBits);
    reinterpret_cast<int2*>(idx)[q] = i;
    reinterpret_cast<int2*>(dist)[q] = d;
}

struct FilterArgs {
    const uint2* keys;
    int splits, max_rows;
    int nq, nt;                    // single pair
    const int32_t* frame_offset;   // batched
    const int32_t* pairs;
    float max_dist, ratio;
    int32_t* out_train;            // [pair][max_rows]
    int32_t* out_count;            // [pair]
};

// src/MapMatcher.cpp:151-161: drop if d0 > max_dist (float, non-strict) or, with k = 2,
// d0 > ratio * d1 (float multiply, no contraction possible: the product feeds a compare).
__global__ void __launch_bounds__(kFinishThreads) bf_finish_filter_kernel(const FilterArgs a)
{
    __shared__ uint2 s_part[kFinishThreads];
    const int pair = blockIdx.y;
    int nq = a.nq, nt = a.nt;
    if (a.pairs != nullptr) {
        const int fq = a.pairs[2 * pair], ft = a.pairs[2 * pair + 1];
        nq = a.frame_offset[fq + 1] - a.frame_offset[fq];
        nt = a.frame_offset[ft + 1] - a.frame_offset[ft];
    }
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const uint2 m = merge_partials_2d(a.keys + static_cast<size_t>(pair) * a.splits * a.max_rows + q, a.splits, a.max_rows,
                                      q < nq, s_part);
    if (threadIdx.y != 0) return;                       // whole warps: blockDim.x is a multiple of 32
    bool keep = false;
    int train = -1;
    if (q < nq) {
        if (m.x != kKeyNone) {
            const float d0 = static_cast<float>(m.x >> kIdxBits);
            keep = !(d0 > a.max_dist);
            if (keep && nt >= 2 && m.y != kKeyNone) {
                const float d1 = static_cast<float>(m.y >> kIdxBits);
                keep = !(d0 > __fmul_rn(a.ratio, d1));
            }
            if (keep) train = static_cast<int>(m.x & kIdxMask);
        }
        a.out_train[static_cast<size_t>(pair) * a.max_rows + q] = train;
    } else if (q < a.max_rows) {
        a.out_train[static_cast<size_t>(pair) * a.max_rows + q] = -1;   // ragged batch padding
    }
    const unsigned ballot = __ballot_sync(0xFFFFFFFFu, keep);
    if (a.out_count != nullptr && (threadIdx.x & 31) == 0 && ballot != 0) atomicAdd(&a.out_count[pair], __popc(ballot));
}

// cross-check: (q,t) survives iff t = argmin row q and q = argmin column t (lowest index on ties).
// strict_max_dist < 0: raw cross-check.  >= 0: also require distance < strict_max_dist
// (src/features/FeatureExtractor.cpp:25).
__global__ void bf_mutual_kernel(const uint2* row_keys, int row_splits, int row_stride, const uint2* col_keys,
                                 int col_splits, int col_stride, int nq, float strict_max_dist, int32_t* idx,
                                 int32_t* dist)
{
    __shared__ uint2 s_part[kFinishThreads];
    __shared__ int s_t[kFinishThreads];
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const uint2 m = merge_partials_2d(row_keys + q, row_splits, row_stride, q < nq, s_part);
    // the row's nearest train row, handed to the row's other threads for the column merge
    if (threadIdx.y == 0) s_t[threadIdx.x] = (q < nq && m.x != kKeyNone) ? static_cast<int>(m.x & kIdxMask) : -1;
    __syncthreads();
    const int t = s_t[threadIdx.x];
    const uint2 c = merge_partials_2d(col_keys + max(t, 0), col_splits, col_stride, t >= 0, s_part);
    if (q >= nq || threadIdx.y != 0) return;
    int oi = -1, od = -1;
    if (t >= 0 && c.x != kKeyNone && static_cast<int>(c.x & kIdxMask) == q) {
        const int d = static_cast<int>(m.x >> kIdxBits);
        if (strict_max_dist < 0.0f || static_cast<float>(d) < strict_max_dist) {
            oi = t;
            od = d;
        }
    }
    idx[q] = oi;
    dist[q] = od;
}

// same, with the column minima produced by the fused scan (one key per train row, no splits)
__global__ void bf_mutual_direct_kernel(const uint2* row_keys, int row_splits, int row_stride,
                                        const uint32_t* col_keys, int nq, float strict_max_dist, int32_t* idx,
                                        int32_t* dist)
{
    __shared__ uint2 s_part[kFinishThreads];
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const uint2 m = merge_partials_2d(row_keys + q, row_splits, row_stride, q < nq, s_part);
    if (q >= nq || threadIdx.y != 0) return;
    int oi = -1, od = -1;
    if (m.x != kKeyNone) {
        const int t = static_cast<int>(m.x & kIdxMask);
        const uint32_t c = col_keys[t];
        if (c != kKeyNone && static_cast<int>(c & kIdxMask) == q) {
            const int d = static_cast<int>(m.x >> kIdxBits);
            if (strict_max_dist < 0.0f || static_cast<float>(d) < strict_max_dist) {
                oi = t;
                od = d;
            }
        }
    }
    idx[q] = oi;
    dist[q] = od;
}

// exclusive scan of per-pair counts (single CTA; n is at most a few hundred thousand)
__global__ void bf_scan_counts_kernel(const int32_t* count, int n, int32_t* offset, int32_t* mirror)
{
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    if (threadIdx.x == 0) s_carry = offset[0];   // running total of the slices before this one
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < n; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const int v = i < n ? count[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_warp[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int w = lane < (blockDim.x >> 5) ? s_warp[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xFFFFFFFFu, w, o);
                if (lane >= o) w += y;
            }
            s_warp[lane] = w;   // inclusive scan of warp totals
        }
        __syncthreads();
        const int warp_off = warp == 0 ? 0 : s_warp[warp - 1];
        const int carry = s_carry;
        if (i < n) {
            offset[i] = carry + warp_off + x - v;
            if (mirror) mirror[i] = carry + warp_off + x - v;
        }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) s_carry = carry + warp_off + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        offset[n] = s_carry;
        if (mirror) mirror[n] = s_carry;
    }
}

// ordered compaction: one warp per pair walks its rows 32 at a time
__global__ void bf_compact_kernel(const int32_t* out_train, const int32_t* frame_offset, const int32_t* pairs,
                                  int nq_single, int n_pairs, int max_rows, const int32_t* offset,
                                  int32_t* query_list, int32_t* train_list, long long capacity)
{
    const int pair = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (pair >= n_pairs) return;
    int nq = nq_single;
    if (pairs != nullptr) {
        const int fq = pairs[2 * pair];
        nq = frame_offset[fq + 1] - frame_offset[fq];
    }
    int pos = offset[pair];
    const int32_t* src = out_train + static_cast<size_t>(pair) * max_rows;
    for (int base = 0; base < nq; base += 32) {
        const int q = base + lane;
        const int t = q < nq ? src[q] : -1;
        const unsigned ballot = __ballot_sync(0xFFFFFFFFu, t >= 0);
        if (t >= 0) {
            const int p = pos + __popc(ballot & ((1u << lane) - 1u));
            if (p >= 0 && p < capacity) {   // an undersized list (or a wrapped int32 total) is reported by the host from the offsets
                query_list[p] = q;
                train_list[p] = t;
            }
        }
        pos += __popc(ballot);
    }
}

// ---- a short batch (a loop-closure query: at most a few hundred pairs), results to host memory ----------------
// The general form above is finish -> exclusive scan of the counts (one CTA) -> ordered compaction (one warp per
// pair, 125 dependent 32-row steps for a 4000-row frame): right for tens of thousands of pairs, slow for five.
// Here one CTA owns a pair: it filters all of the pair's rows and compacts them into the pair's own block of a
// device scratch (block-wide ballot scan, 1024 rows per step); a second launch gives every pair its offset (the
// sum of the counts before it: n_pairs <= 1024 values) and moves the block into the caller's lists.
constexpr int kFcThreads = 1024;

__global__ void __launch_bounds__(kFcThreads) bf_finish_compact_kernel(const FilterArgs a, int32_t* loc_q, int32_t* loc_t)
{
    __shared__ int s_warp[32];
    __shared__ int s_base;
    const int pair = blockIdx.x;
    const int fq = a.pairs[2 * pair], ft = a.pairs[2 * pair + 1];
    const int nq = a.frame_offset[fq + 1] - a.frame_offset[fq];
    const int nt = a.frame_offset[ft + 1] - a.frame_offset[ft];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    int32_t* lq = loc_q + static_cast<size_t>(pair) * a.max_rows;
    int32_t* lt = loc_t + static_cast<size_t>(pair) * a.max_rows;
    const uint2* keys = a.keys + static_cast<size_t>(pair) * a.splits * a.max_rows;
    for (int base = 0; base < nq; base += kFcThreads) {
        const int q = base + threadIdx.x;
        bool keep = false;
        int train = -1;
        if (q < nq) {                                         // the same tests as bf_finish_filter_kernel
            const uint2 m = merge_partials(keys + q, a.splits, a.max_rows);
            if (m.x != kKeyNone) {
                const float d0 = static_cast<float>(m.x >> kIdxBits);
                keep = !(d0 > a.max_dist);
                if (keep && nt >= 2 && m.y != kKeyNone) {
                    const float d1 = static_cast<float>(m.y >> kIdxBits);
                    keep = !(d0 > __fmul_rn(a.ratio, d1));
                }
                if (keep) train = static_cast<int>(m.x & kIdxMask);
            }
        }
        const unsigned ballot = __ballot_sync(0xFFFFFFFFu, keep);
        if (lane == 0) s_warp[warp] = __popc(ballot);
        __syncthreads();
        if (warp == 0) {
            int w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xFFFFFFFFu, w, o);
                if (lane >= o) w += y;
            }
            s_warp[lane] = w;                                 // inclusive scan of the warps' counts
        }
        __syncthreads();
        const int at = s_base + (warp ? s_warp[warp - 1] : 0) + __popc(ballot & ((1u << lane) - 1u));
        const int total = s_warp[31];
        if (keep) {
            lq[at] = q;
            lt[at] = train;
        }
        __syncthreads();
        if (threadIdx.x == 0) s_base += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) a.out_count[pair] = s_base;
}

__global__ void __launch_bounds__(256) bf_gather_lists_kernel(const int32_t* count, int n_pairs, int max_rows,
                                                              const int32_t* loc_q, const int32_t* loc_t, int32_t* offset,
                                                              int32_t* mirror, int32_t* query_list, int32_t* train_list,
                                                              long long capacity)
{
    __shared__ int s_part[8];
    __shared__ int s_off;
    const int pair = blockIdx.x;
    int v = 0;
    for (int i = threadIdx.x; i < pair; i += 256) v += count[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        int off = 0;
        for (int w = 0; w < 8; ++w) off += s_part[w];
        s_off = off;
    }
    __syncthreads();
    const int off = s_off, n = count[pair];
    if (threadIdx.x == 0) {
        offset[pair] = off;
        if (mirror) mirror[pair] = off;
        if (pair == n_pairs - 1) {
            offset[n_pairs] = off + n;
            if (mirror) mirror[n_pairs] = off + n;
        }
    }
    const int32_t* lq = loc_q + static_cast<size_t>(pair) * max_rows;
    const int32_t* lt = loc_t + static_cast<size_t>(pair) * max_rows;
    for (int i = threadIdx.x; i < n; i += 256) {
        const long long p = static_cast<long long>(off) + i;
        if (p < capacity) {       // an undersized list is reported by the host from the offsets
            query_list[p] = lq[i];
            train_list[p] = lt[i];
        }
    }
}

} // namespace

int bf_launch_finish_compact(slamgpu_ctx* ctx, const BfPlan& plan, const uint2* d_keys, const int32_t* d_frame_offset,
                             const int32_t* d_pairs, int n_pairs, int max_rows, float max_dist, float ratio,
                             int32_t* d_local /* 2 x n_pairs x max_rows */, int32_t* d_count, int32_t* d_offset,
                             int32_t* h_offset_mirror, int32_t* query_list, int32_t* train_list, int64_t capacity)
{
    if (n_pairs <= 0 || max_rows <= 0) return SLAMGPU_OK;
    SG_CHECK(n_pairs <= 65535 && d_pairs != nullptr, "finish_compact: short pair lists only");
    FilterArgs a;
    a.keys = d_keys;
    a.splits = plan.splits;
    a.max_rows = max_rows;
    a.nq = 0;
    a.nt = 0;
    a.frame_offset = d_frame_offset;
    a.pairs = d_pairs;
    a.max_dist = max_dist;
    a.ratio = ratio;
    a.out_train = nullptr;
    a.out_count = d_count;
    int32_t* loc_q = d_local;
    int32_t* loc_t = d_local + static_cast<size_t>(n_pairs) * max_rows;
    bf_finish_compact_kernel<<<n_pairs, kFcThreads, 0, ctx->stream>>>(a, loc_q, loc_t);
    bf_gather_lists_kernel<<<n_pairs, 256, 0, ctx->stream>>>(d_count, n_pairs, max_rows, loc_q, loc_t, d_offset,
                                                             h_offset_mirror, query_list, train_list,
                                                             static_cast<long long>(capacity));
    ctx->launches += 2;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

BfPlan bf_make_plan(int n_sm, int nq_max, int nt_max, int64_t n_pairs)
{
    BfPlan p;
    const int64_t target = static_cast<int64_t>(n_sm) * 4;   // ~32 resident warps per SM
    // 4 query rows per thread when that still leaves enough CTAs to fill the machine, else 2
    const int64_t work_r4 = ((nq_max + 1023) / 1024) * static_cast<int64_t>((nt_max + kChunk - 1) / kChunk) * n_pairs;
    p.r = work_r4 >= target ? 4 : 2;
    p.tile = kThreads * p.r;
    p.tiles = max(1, (nq_max + p.tile - 1) / p.tile);
    const int64_t ctas_no_split = static_cast<int64_t>(p.tiles) * n_pairs;
    int64_t splits = (target + ctas_no_split - 1) / ctas_no_split;
    const int64_t max_splits = max(1, (nt_max + 31) / 32);
    splits = min(max(splits, int64_t(1)), max_splits);
    int tps = static_cast<int>((nt_max + splits - 1) / splits);
    tps = max(8, (tps + 7) & ~7);
    p.t_per_split = tps;
    p.splits = max(1, (nt_max + tps - 1) / tps);
    return p;
}

size_t bf_partial_elems(const BfPlan& p, int64_t n_pairs, int max_rows)
{
    return static_cast<size_t>(n_pairs) * p.splits * max_rows;
}

// CTA shape of the finish kernels: (rows, split groups) with about four partials per thread, at most 8 groups
static dim3 finish_block(int splits)
{
    const int g = splits <= 4 ? 1 : splits <= 8 ? 2 : splits <= 16 ? 4 : 8;
    return dim3(kFinishThreads / g, g);
}

int bf_launch_scan(slamgpu_ctx* ctx, const BfPlan& plan, const uint8_t* d_query, int nq, const uint8_t* d_train,
                   int nt, const int32_t* d_frame_offset, const int32_t* d_pairs, int n_pairs, int max_rows,
                   bool swap_roles, uint2* d_keys, uint32_t* d_col_keys)
{
    ScanArgs a;
    a.q_base = d_query;
    a.t_base = d_train;
    a.frame_offset = d_frame_offset;
    a.pairs = d_pairs;
    a.nq = nq;
    a.nt = nt;
    a.tiles = plan.tiles;
    a.splits = plan.splits;
    a.t_per_split = plan.t_per_split;
    a.max_rows = max_rows;
    a.swap_roles = swap_roles ? 1 : 0;
    a.keys = d_keys;
    a.col_keys = d_col_keys;
    SG_CHECK(d_col_keys == nullptr || d_pairs == nullptr, "fused column minima are for single-pair scans");
    const int64_t n_ctas = static_cast<int64_t>(plan.tiles) * plan.splits * n_pairs;
    if (n_ctas <= 0) return SLAMGPU_OK;
    SG_CHECK(n_ctas < (int64_t(1) << 31), "brute-force grid too large; split the pair list");
    const dim3 grid(static_cast<unsigned>(n_ctas));
    ProfScope prof(ctx, SLAMGPU_KERNEL_BF_SCAN);
    if (d_col_keys != nullptr) {
        if (plan.r == 4) {
            bf_scan_kernel<2, true><<<grid, kThreads, 0, ctx->stream>>>(a);
        } else {
            bf_scan_kernel<1, true><<<grid, kThreads, 0, ctx->stream>>>(a);
        }
    } else if (plan.r == 4) {
        bf_scan_kernel<2, false><<<grid, kThreads, 0, ctx->stream>>>(a);
    } else {
        bf_scan_kernel<1, false><<<grid, kThreads, 0, ctx->stream>>>(a);
    }
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int bf_launch_finish_raw(slamgpu_ctx* ctx, const BfPlan& plan, const uint2* d_keys, int nq, int nt, int max_rows,
                         int32_t* d_idx, int32_t* d_dist)
{
    (void)nt;
    if (nq <= 0) return SLAMGPU_OK;
    const dim3 blk = finish_block(plan.splits);
    bf_finish_raw_kernel<<<(nq + blk.x - 1) / blk.x, blk, 0, ctx->stream>>>(d_keys, plan.splits, max_rows, nq, d_idx, d_dist);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int bf_launch_finish_filter(slamgpu_ctx* ctx, const BfPlan& plan, const uint2* d_keys, int nq, int nt,
                            const int32_t* d_frame_offset, const int32_t* d_pairs, int n_pairs, int max_rows,
                            float max_dist, float ratio, int32_t* d_out_train, int32_t* d_out_count, cudaStream_t stream)
{
    if (n_pairs <= 0 || max_rows <= 0) return SLAMGPU_OK;
    if (stream == nullptr) stream = ctx->stream;
    // d_out_count null: the caller counts the surviving rows itself (the single-pair host call does)
    if (d_out_count) SG_CUDA(cudaMemsetAsync(d_out_count, 0, sizeof(int32_t) * n_pairs, stream));
    FilterArgs a;
    a.keys = d_keys;
    a.splits = plan.splits;
    a.max_rows = max_rows;
    a.nq = nq;
    a.nt = nt;
    a.frame_offset = d_frame_offset;
    a.pairs = d_pairs;
    a.max_dist = max_dist;
    a.ratio = ratio;
    a.out_train = d_out_train;
    a.out_count = d_out_count;
    const int rows = d_pairs ? max_rows : nq;
    for (int p0 = 0; p0 < n_pairs; p0 += 65535) {   // gridDim.y limit
        const int np = min(65535, n_pairs - p0);
        FilterArgs b = a;
        b.keys = a.keys + static_cast<size_t>(p0) * plan.splits * max_rows;
        b.pairs = d_pairs ? d_pairs + 2 * static_cast<size_t>(p0) : nullptr;
        b.out_train = d_out_train + static_cast<size_t>(p0) * max_rows;
        b.out_count = d_out_count ? d_out_count + p0 : nullptr;
        const dim3 blk = finish_block(plan.splits);
        bf_finish_filter_kernel<<<dim3((rows + blk.x - 1) / blk.x, np), blk, 0, stream>>>(b);
        ctx->launches++;
    }
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int bf_launch_mutual(slamgpu_ctx* ctx, const BfPlan& row_plan, const uint2* d_row_keys, int row_stride,
                     const BfPlan& col_plan, const uint2* d_col_keys, int col_stride, int nq, int nt,
                     float strict_max_dist, int32_t* d_idx, int32_t* d_dist)
{
    (void)nt;
    if (nq <= 0) return SLAMGPU_OK;
    const dim3 blk = finish_block(std::max(row_plan.splits, col_plan.splits));
    bf_mutual_kernel<<<(nq + blk.x - 1) / blk.x, blk, 0, ctx->stream>>>(d_row_keys, row_plan.splits, row_stride, d_col_keys,
                                                                       col_plan.splits, col_stride, nq, strict_max_dist,
                                                                       d_idx, d_dist);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int bf_launch_mutual_direct(slamgpu_ctx* ctx, const BfPlan& row_plan, const uint2* d_row_keys, int row_stride,
                            const uint32_t* d_col_keys, int nq, float strict_max_dist, int32_t* d_idx, int32_t* d_dist)
{
    if (nq <= 0) return SLAMGPU_OK;
    const dim3 blk = finish_block(row_plan.splits);
    bf_mutual_direct_kernel<<<(nq + blk.x - 1) / blk.x, blk, 0, ctx->stream>>>(d_row_keys, row_plan.splits, row_stride,
                                                                              d_col_keys, nq, strict_max_dist, d_idx, d_dist);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int bf_launch_scan_counts(slamgpu_ctx* ctx, const int32_t* d_count, int n, int32_t* d_offset, int32_t* h_offset_mirror,
                          cudaStream_t stream)
{
    bf_scan_counts_kernel<<<1, 1024, 0, stream ? stream : ctx->stream>>>(d_count, n, d_offset, h_offset_mirror);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int bf_launch_compact(slamgpu_ctx* ctx, const int32_t* d_out_train, const int32_t* d_frame_offset,
                      const int32_t* d_pairs, int nq_single, int n_pairs, int max_rows, const int32_t* d_offset,
                      int32_t* d_query_list, int32_t* d_train_list, int64_t capacity, cudaStream_t stream)
{
    if (n_pairs <= 0) return SLAMGPU_OK;
    const int warps_per_cta = 8;
    bf_compact_kernel<<<(n_pairs + warps_per_cta - 1) / warps_per_cta, warps_per_cta * 32, 0, stream ? stream : ctx->stream>>>(
        d_out_train, d_frame_offset, d_pairs, nq_single, n_pairs, max_rows, d_offset, d_query_list, d_train_list,
        static_cast<long long>(capacity));
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

namespace {
__global__ void __launch_bounds__(256) bf_push_lists_kernel(const int32_t* __restrict__ offset, int n_pairs,
                                                            const int32_t* __restrict__ ql, const int32_t* __restrict__ tl,
                                                            int32_t* h_ql, int32_t* h_tl, long long capacity)
{
    const long long begin = offset[0];
    long long end = offset[n_pairs];
    if (end > capacity) end = capacity;                  // an undersized list is reported by the host from the offsets
    if (begin < 0 || end <= begin) return;               // (a wrapped int32 total as well)
    const long long n = end - begin, stride = static_cast<long long>(gridDim.x) * blockDim.x;
    // the two lists as one index space: warps write whole 128-byte lines towards the host
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < 2 * n; i += stride) {
        if (i < n) h_ql[begin + i] = ql[begin + i];
        else h_tl[begin + i - n] = tl[begin + i - n];
    }
}
} // namespace

int bf_launch_push_lists(slamgpu_ctx* ctx, const int32_t* d_offset, int n_pairs, const int32_t* d_query_list,
                         const int32_t* d_train_list, int32_t* h_query_list, int32_t* h_train_list, int64_t capacity,
                         cudaStream_t stream)
{
    if (n_pairs <= 0) return SLAMGPU_OK;
    // few CTAs: the copy shares the SMs with the next slice's scan and is bound by the PCIe link, not by threads
    bf_push_lists_kernel<<<32, 256, 0, stream>>>(d_offset, n_pairs, d_query_list, d_train_list, h_query_list, h_train_list,
                                                 static_cast<long long>(capacity));
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

// ---------------------------------------------------------------------------------------------
// POPC-pipe microbenchmark: long chains of independent POPCs, nothing else on the critical pipe.
// ---------------------------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) popc_bench_kernel(uint32_t seed, int iters, uint32_t* sink, long long* cycles)
{
    uint32_t x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u + blockIdx.x;
    uint32_t acc = 0;
    long long c0, c1;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(c0)::"memory");
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0)::"memory");
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            x[i] = __popc(x[i]) + seed;   // dependent per chain, 8 independent chains per thread
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc ^= x[i];
    if (acc == 0xDEADBEEFu) sink[0] = acc;   // keeps the chains alive; orders the clock reads after them
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(c1)::"memory");
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1)::"memory");
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        cycles[0] = c1 - c0;
        cycles[1] = static_cast<long long>(t1 - t0);   // ns
    }
}
} // namespace

int popc_microbench(slamgpu_ctx* ctx, double* popc_per_s, int* n_sm, double* sm_mhz_est)
{
    SG_TRY(ensure_device_ws(ctx, 4096));
    uint32_t* sink = static_cast<uint32_t*>(ctx->d_ws.ptr);
    long long* cyc = reinterpret_cast<long long*>(sink + 16);
    const int iters = 20000, ctas = ctx->n_sm * 8;
    cudaEvent_t e0, e1;
    SG_CUDA(cudaEventCreate(&e0));
    SG_CUDA(cudaEventCreate(&e1));
    popc_bench_kernel<<<ctas, 256, 0, ctx->stream>>>(12345u, 2000, sink, cyc);   // warm-up
    SG_CUDA(cudaEventRecord(e0, ctx->stream));
    popc_bench_kernel<<<ctas, 256, 0, ctx->stream>>>(12345u, iters, sink, cyc);
    SG_CUDA(cudaEventRecord(e1, ctx->stream));
    ctx->launches += 2;
    SG_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    SG_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    long long h_cyc[2] = {0, 0};
    SG_CUDA(cudaMemcpy(h_cyc, cyc, sizeof(h_cyc), cudaMemcpyDeviceToHost));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    const double ops = static_cast<double>(ctas) * 256.0 * 8.0 * iters;
    if (popc_per_s) *popc_per_s = ops / (ms * 1e-3);
    if (n_sm) *n_sm = ctx->n_sm;
    // SM cycles over wall nanoseconds seen by one thread of CTA 0: the SM clock under this load, in MHz
    if (sm_mhz_est) *sm_mhz_est = h_cyc[1] > 0 ? static_cast<double>(h_cyc[0]) / static_cast<double>(h_cyc[1]) * 1e3 : 0.0;
    return SLAMGPU_OK;
}

} // namespace slamgpu

// ---------------------------------------------------------------------------------------------
// slamgpu_knn2_dev_noalloc: the integer-pipe kNN-2 scan for callers that must not allocate (code under CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The scan's split partials -- the scratch slamgpu_knn2_dev takes from the context's
// growable workspace (ensure_device_ws: cudaMalloc on growth) -- live in a file-scope device array of fixed size.
// Every GPU that loads the library holds these 8 MiB until exit.  ALL calls on a device share the array: two calls
// that overlap on the GPU (two contexts = two streams) race on it; calls on one context are stream-ordered and safe.
// The array keeps the last call's contents; nothing in it has to start at zero: a call writes every (split, row)
// entry its finish kernel reads (bf_scan_kernel stores all rows < nq of every split) before reading it.
// ---------------------------------------------------------------------------------------------
namespace slamgpu {
namespace {

constexpr int kNoAllocMaxRows = 8192;                         // nq and nt limit of the entry point
constexpr size_t kNoAllocKeys = size_t(1) << 20;              // uint2 entries: 8 MiB
static __device__ __align__(256) uint2 g_noalloc_keys[kNoAllocKeys];

// In host code `g_noalloc_keys` is a host-side stand-in for the symbol, not a device address: the address a kernel
// can use comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr int kNoAllocMaxDevices = 64;
uint2* g_noalloc_ptr[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mutex;

int noalloc_scratch(int device, uint2** out)
{
    SG_CHECK(device >= 0 && device < kNoAllocMaxDevices, "device index beyond the no-alloc scratch table");
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[device] == nullptr) {
        void* p = nullptr;
        SG_CUDA(cudaGetSymbolAddress(&p, g_noalloc_keys));
        SG_CHECK(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
        g_noalloc_ptr[device] = static_cast<uint2*>(p);
    }
    *out = g_noalloc_ptr[device];
    return SLAMGPU_OK;
}

} // namespace

// set-up hook of slamgpu_create (device current): resolves the scratch address so that no later call has to
int bf_noalloc_setup(int device)
{
    uint2* p = nullptr;
    return noalloc_scratch(device, &p);
}

} // namespace slamgpu

extern "C" int slamgpu_knn2_dev_noalloc(slamgpu_ctx* ctx, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                                        int32_t* d_out_idx, int32_t* d_out_dist)
{
    using namespace slamgpu;
    SG_CHECK(ctx != nullptr, "null context");
    SG_CHECK(nq >= 0 && nt >= 0, "negative descriptor count");
    SG_CHECK(nq <= kNoAllocMaxRows && nt <= kNoAllocMaxRows,
             "slamgpu_knn2_dev_noalloc takes at most 8192 query and 8192 train descriptors (fixed scratch); "
             "use slamgpu_knn2_dev for larger sets");
    if (nq == 0) return SLAMGPU_OK;
    SG_CHECK(d_query != nullptr && d_out_idx != nullptr && d_out_dist != nullptr && (nt == 0 || d_train != nullptr), "null pointer");
    SG_CHECK(((reinterpret_cast<uintptr_t>(d_query) | reinterpret_cast<uintptr_t>(d_train)) & 15u) == 0,
             "device descriptor pointers must be 16-byte aligned");
    SG_CUDA(cudaSetDevice(ctx->device));
    if (nt == 0) {   // every neighbour missing (as slamgpu_knn2_dev)
        SG_CUDA(cudaMemsetAsync(d_out_idx, 0xFF, sizeof(int32_t) * 2 * nq, ctx->stream));
        SG_CUDA(cudaMemsetAsync(d_out_dist, 0xFF, sizeof(int32_t) * 2 * nq, ctx->stream));
        return SLAMGPU_OK;
    }
    uint2* keys = nullptr;
    SG_TRY(noalloc_scratch(ctx->device, &keys));
    const BfPlan plan = bf_make_plan(ctx->n_sm, nq, nt, 1);
    SG_CHECK(bf_partial_elems(plan, 1, nq) <= kNoAllocKeys, "scan partials exceed the fixed scratch of slamgpu_knn2_dev_noalloc");
    // the launches of bf_launch_scan / bf_launch_finish_raw without their profiling scope (which creates events)
    ScanArgs a;
    a.q_base = d_query;
    a.t_base = d_train;
    a.frame_offset = nullptr;
    a.pairs = nullptr;
    a.nq = nq;
    a.nt = nt;
    a.tiles = plan.tiles;
    a.splits = plan.splits;
    a.t_per_split = plan.t_per_split;
    a.max_rows = nq;
    a.swap_roles = 0;
    a.keys = keys;
    a.col_keys = nullptr;
    const dim3 grid(static_cast<unsigned>(plan.tiles) * static_cast<unsigned>(plan.splits));
    if (plan.r == 4) bf_scan_kernel<2, false><<<grid, kThreads, 0, ctx->stream>>>(a);
    else bf_scan_kernel<1, false><<<grid, kThreads, 0, ctx->stream>>>(a);
    const dim3 blk = finish_block(plan.splits);
    bf_finish_raw_kernel<<<(nq + blk.x - 1) / blk.x, blk, 0, ctx->stream>>>(keys, plan.splits, nq, nq, d_out_idx, d_out_dist);
    ctx->launches += 2;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}
