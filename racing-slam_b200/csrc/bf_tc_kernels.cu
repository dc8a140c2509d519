// bf_tc_kernels.cu -- tensor-core brute-force Hamming kNN-2 scan for sm_100a (tcgen05 + TMEM).
//
// Same contract as bf_scan_kernel (bf_kernels.cu): per query row the best two packed keys
// (distance << 22 | train index) of cv::BFMatcher::knnMatch as the reference calls it
// (src/MapMatcher.cpp:147-148), written as split partials that the finish
// kernels merge.  north_star's optional K7 ("tensor-core path ... only if ncu shows the popc
// pipe is the limiter"): ncu showed exactly that for the integer kernels (XU 96 % in v1, ALU 87 %
// in v2, profiles/r01_ncu_full_*), and `mma.sync ... b1.and.popc` is emulated on sm_100a, so the
// path uses the native 5th-generation tensor cores instead:
//
//   * every descriptor bit becomes one int8 (+8 / -8): a 256-bit descriptor is a K = 256 row and
//     the Hamming distance is an int8 dot product, dot = 64 * (256 - 2 * distance), exact in int32;
//   * one extra K = 32 slice (constant in shared memory) adds  16385 + 127 - (column & 127)  so
//     that the accumulator IS the packed 16-bit sort key  32896 - (distance << 7 | column & 127):
//     larger = closer, ties to the lower column -- the reference's order -- with no epilogue math;
//   * tcgen05.mma.cta_group::1.kind::i8, M = 128 query rows x N = 256 train rows per tile,
//     9 MMAs (K = 8 x 32 + the key slice) into a double-buffered 2 x 256-column TMEM accumulator;
//   * operands are staged by cp.async.bulk (TMA, SASS UBLKCP) from an expanded copy of the
//     descriptors laid out in the UMMA no-swizzle K-major canonical form (8-row x 16-byte core
//     matrices), so a tile is one contiguous block of global memory;
//   * 8 epilogue warps read TMEM with tcgen05.ld (thread = query row), pack columns c and c + 128
//     into one register and keep the top-2 with 3 VIMNMX.U16x2 per two comparisons.
//
// Warp roles (320 threads, one persistent CTA per SM): warp 0 = bulk-copy producer, warp 1 = TMEM
// allocation + MMA issue (one thread), warps 2..9 = epilogue.
#include "slamgpu_internal.cuh"
#include "bf_tc_common.cuh"

namespace slamgpu {

namespace {

constexpr int TM = 128;                    // query rows per tile (UMMA M)
constexpr int TN = 256;                    // train rows per tile (UMMA N)
constexpr int kRowBytes = 256;             // expanded descriptor: 256 int8
constexpr int kGroupBytes = 8 * kRowBytes; // 8 rows = 16 core matrices of 128 bytes
constexpr int kABytes = TM * kRowBytes;    // 32 KB
constexpr int kBBytes = TN * kRowBytes;    // 64 KB
constexpr int kAXBytes = TM * 32;          // key slice of A: 4 KB
constexpr int kBXBytes = TN * 32;          // key slice of B: 8 KB
constexpr int kTcThreads = 320;
constexpr int kTc2Threads = 64 + 8 * 32;    // CTA-pair kernel: producer, MMA/relay, 8 epilogue warps
constexpr uint32_t kKeyBias = 32896u;      // accumulator = kKeyBias - (distance << 7 | column & 127), in [1, 32896]

// shared-memory map (dynamic, 1024-byte aligned base)
constexpr int kOffA = 0;                         // 2 x 32 KB
constexpr int kOffB = 2 * kABytes;               // 2 x 64 KB
constexpr int kOffAX = kOffB + 2 * kBBytes;      // 4 KB
constexpr int kOffBX = kOffAX + kAXBytes;        // 8 KB
constexpr int kOffBar = kOffBX + kBXBytes;       // 12 mbarriers
constexpr int kOffTmemSlot = kOffBar + 16 * 8;
constexpr int kTcSmemBytes = kOffTmemSlot + 16;

// cute::UMMA::InstrDescriptor for kind::i8: c_format S32 (2) at [4,6), a/b format INT8 (1) at [7,10) / [10,13),
// K-major A and B (bits 15, 16 = 0), N >> 3 at [17,23), M >> 4 at [24,29)
constexpr uint32_t kIdesc = (2u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(TN >> 3) << 17) |
                            (static_cast<uint32_t>(TM >> 4) << 24);

struct TcArgs {
    const uint8_t* exp_q;          // expanded descriptors of the query side (frame 0 of the set)
    const uint8_t* exp_t;          // expanded descriptors of the train side
    long long frame_pitch_bytes;   // batched: expanded bytes per frame
    const int32_t* frame_offset;   // batched: row offsets (for the row counts only); null = single pair
    const int32_t* pairs;          // batched: [n_pairs x 2]
    int nq, nt;                    // single pair
    int tiles, splits, t_per_split;
    int max_rows;
    long long n_units;
    int variant;                   // debug: bit 0 swaps LBO / SBO
    int thr_raw;                   // filtered mode: an element matters iff 16384 - 128 * distance >= thr_raw
    uint2* keys;
};

struct Unit {
    bool valid;
    int pair, split, row0, nq, t_begin, n_btiles, t_end;
    const uint8_t* q;
    const uint8_t* t;
};

__device__ __forceinline__ Unit decode_unit(const TcArgs& a, long long u)
{
    Unit un;
    const int tile = static_cast<int>(u % a.tiles);
    u /= a.tiles;
    un.split = static_cast<int>(u % a.splits);
    un.pair = static_cast<int>(u / a.splits);
    int nq = a.nq, nt = a.nt;
    un.q = a.exp_q;
    un.t = a.exp_t;
    if (a.pairs != nullptr) {
        const int fq = a.pairs[2 * un.pair], ft = a.pairs[2 * un.pair + 1];
        nq = a.frame_offset[fq + 1] - a.frame_offset[fq];
        nt = a.frame_offset[ft + 1] - a.frame_offset[ft];
        un.q = a.exp_q + static_cast<long long>(fq) * a.frame_pitch_bytes;
        un.t = a.exp_t + static_cast<long long>(ft) * a.frame_pitch_bytes;
    }
    un.nq = nq;
    un.row0 = tile * TM;
    un.t_begin = min(un.split * a.t_per_split, nt);
    un.t_end = min(un.t_begin + a.t_per_split, nt);
    un.n_btiles = (un.t_end - un.t_begin + TN - 1) / TN;
    un.valid = un.row0 < nq;     // a valid unit with no train rows still writes "none" keys
    return un;
}

// accumulator key (16 bit, larger = better) -> library key (distance << 22 | train index)
__device__ __forceinline__ uint32_t widen_acc_key(uint32_t k, uint32_t col_base)
{
    const uint32_t key16 = kKeyBias - k;
    return k == 0u ? kKeyNone : (((key16 >> 7) << kIdxBits) | (col_base + (key16 & 127u)));
}
__device__ __forceinline__ void merge_min2(uint32_t k0, uint32_t k1, uint32_t& b0, uint32_t& b1)
{
    const uint32_t hi = max(k0, b0);
    b0 = min(k0, b0);
    b1 = min(min(b1, hi), k1);
}

// 32 packed (column c | column c + 128 << 16) keys into two independent max-top-2 chains
template <bool MASKED>
__device__ __forceinline__ void fold32(const uint32_t (&x)[32], const uint32_t (&y)[32], int col0, int valid,
                                       uint32_t (&c0)[2], uint32_t (&c1)[2])
{
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        uint32_t lo = x[j], hi = y[j];
        if constexpr (MASKED) {      // ragged last tile: columns past the train set hold garbage
            lo = col0 + j < valid ? lo : 0u;
            hi = col0 + j + 128 < valid ? hi : 0u;
        }
        const uint32_t p = mad_u32(hi, 65536u, lo);      // IMAD: FMA pipe, keeps the ALU pipe for the min/max
        const uint32_t m = __vminu2(p, c0[j & 1]);
        c0[j & 1] = __vmaxu2(p, c0[j & 1]);
        c1[j & 1] = __vmaxu2(c1[j & 1], m);
    }
}

__global__ void __launch_bounds__(kTcThreads, 1) bf_tc_scan_kernel(const TcArgs a)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // barriers: a_full[2], a_empty[2], b_full[2], b_empty[2], acc_full[2], acc_empty[2]
    const uint32_t bar = sbase + kOffBar;
    auto A_FULL = [&](uint32_t s) { return bar + 8u * s; };
    auto A_EMPTY = [&](uint32_t s) { return bar + 8u * (2 + s); };
    auto B_FULL = [&](uint32_t s) { return bar + 8u * (4 + s); };
    auto B_EMPTY = [&](uint32_t s) { return bar + 8u * (6 + s); };
    auto ACC_FULL = [&](uint32_t s) { return bar + 8u * (8 + s); };
    auto ACC_EMPTY = [&](uint32_t s) { return bar + 8u * (10 + s); };
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kOffTmemSlot);

    // ---- one-time setup ----
    if (threadIdx.x == 0) {
        for (uint32_t s = 0; s < 2; ++s) {
            mbar_init(A_FULL(s), 1);
            mbar_init(A_EMPTY(s), 1);
            mbar_init(B_FULL(s), 1);
            mbar_init(B_EMPTY(s), 1);
            mbar_init(ACC_FULL(s), 1);
            mbar_init(ACC_EMPTY(s), 8);     // one arrive per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // the constant key slices (canonical layout: 8-row group = [k-chunk 0..1][row 0..7][16 B])
    for (int i = threadIdx.x; i < (kAXBytes + kBXBytes) / 16; i += kTcThreads) {
        const bool is_a = i < kAXBytes / 16;
        const int e = is_a ? i : i - kAXBytes / 16;       // 16-byte element within the slice
        const int group = e >> 4, chunk = (e >> 3) & 1, r = group * 8 + (e & 7);
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (chunk == 0) {
            // A: [1, 127, 127, 1]   B: [127 - (col & 127), 127, 2, 2]   ->  sum = 16385 + 127 - (col & 127)
            v.x = is_a ? (1u | (127u << 8) | (127u << 16) | (1u << 24))
                       : (static_cast<uint32_t>(127 - (r & 127)) | (127u << 8) | (2u << 16) | (2u << 24));
        }
        *reinterpret_cast<uint4*>(smem + (is_a ? kOffAX : kOffBX) + e * 16) = v;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the tensor core
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + kOffTmemSlot),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const long long stride = gridDim.x;
    if (warp == 0) {
        // ===== producer: bulk copies of the expanded tiles =====
        if (lane == 0) {
            uint32_t a_it = 0, b_it = 0;
            for (long long u = blockIdx.x; u < a.n_units; u += stride) {
                const Unit un = decode_unit(a, u);
                if (!un.valid) continue;
                const uint32_t sa = a_it & 1u;
                mbar_wait(A_EMPTY(sa), ((a_it >> 1) & 1u) ^ 1u);
                mbar_expect_tx(A_FULL(sa), kABytes);
                const uint8_t* qs = un.q + static_cast<long long>(un.row0) * kRowBytes;
#pragma unroll
                for (int c = 0; c < 2; ++c)
                    bulk_g2s(sbase + kOffA + sa * kABytes + c * 16384, qs + c * 16384, 16384u, A_FULL(sa));
                ++a_it;
                for (int bt = 0; bt < un.n_btiles; ++bt) {
                    const uint32_t sb = b_it & 1u;
                    mbar_wait(B_EMPTY(sb), ((b_it >> 1) & 1u) ^ 1u);
                    mbar_expect_tx(B_FULL(sb), kBBytes);
                    const uint8_t* ts = un.t + static_cast<long long>(un.t_begin + bt * TN) * kRowBytes;
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        bulk_g2s(sbase + kOffB + sb * kBBytes + c * 16384, ts + c * 16384, 16384u, B_FULL(sb));
                    ++b_it;
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: one thread =====
        if (lane == 0) {
            const bool swap = SG_TC_VARIANT(a.variant, 1);
            const uint32_t lbo = swap ? kGroupBytes : 128u, sbo = swap ? 128u : kGroupBytes;
            const uint32_t lbo_x = swap ? 256u : 128u, sbo_x = swap ? 128u : 256u;
            const uint64_t desc_ax = umma_desc(sbase + kOffAX, lbo_x, sbo_x);
            const uint64_t desc_bx = umma_desc(sbase + kOffBX, lbo_x, sbo_x);
            uint32_t a_it = 0, b_it = 0, c_it = 0;
            for (long long u = blockIdx.x; u < a.n_units; u += stride) {
                const Unit un = decode_unit(a, u);
                if (!un.valid) continue;
                const uint32_t sa = a_it & 1u;
                mbar_wait(A_FULL(sa), (a_it >> 1) & 1u);
                for (int bt = 0; bt < un.n_btiles; ++bt) {
                    const uint32_t sb = b_it & 1u, sc = c_it & 1u;
                    mbar_wait(ACC_EMPTY(sc), ((c_it >> 1) & 1u) ^ 1u);
                    mbar_wait(B_FULL(sb), (b_it >> 1) & 1u);
                    tc_fence_after();
                    const uint32_t d = tmem_base + sc * TN;
                    const uint32_t a0 = sbase + kOffA + sa * kABytes, b0 = sbase + kOffB + sb * kBBytes;
#pragma unroll
                    for (int k = 0; k < 8; ++k)     // K = 32 bytes per MMA = two 16-byte k-chunks = 256 bytes of the tile
                        tc_mma_i8(d, umma_desc(a0 + k * 256, lbo, sbo), umma_desc(b0 + k * 256, lbo, sbo), kIdesc,
                                  k > 0 ? 1u : 0u);
                    tc_mma_i8(d, desc_ax, desc_bx, kIdesc, 1u);
                    tc_commit(B_EMPTY(sb));          // smem stage free when these MMAs have read it
                    tc_commit(ACC_FULL(sc));         // accumulator ready for the epilogue
                    ++b_it;
                    ++c_it;
                }
                tc_commit(A_EMPTY(sa));
                ++a_it;
            }
        }
    } else {
        // ===== epilogue: 8 warps, thread = query row, warp pair splits the columns =====
        const int quad = warp & 3;                 // TMEM lanes [32 * quad, +32) are the ones this warp may read
        const int g = (warp - 2) >> 2;             // 0: columns [0,64) u [128,192)   1: [64,128) u [192,256)
        const int row_in_tile = quad * 32 + lane;
        uint32_t c_it = 0;
        for (long long u = blockIdx.x; u < a.n_units; u += stride) {
            const Unit un = decode_unit(a, u);
            if (!un.valid) continue;
            uint32_t b0 = kKeyNone, b1 = kKeyNone;
            for (int bt = 0; bt < un.n_btiles; ++bt) {
                const uint32_t sc = c_it & 1u;
                mbar_wait(ACC_FULL(sc), (c_it >> 1) & 1u);
                tc_fence_after();
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + sc * TN + g * 64;
                const int cb = un.t_begin + bt * TN;
                const int valid = un.t_end - cb;
                uint32_t x[32], y[32], x2[32], y2[32];
                uint32_t c0[2] = {0u, 0u}, c1[2] = {0u, 0u};
                tmem_ld32(taddr, x);
                tmem_ld32(taddr + 128, y);
                tmem_wait_ld();
                tmem_pin_regs(x);
                tmem_pin_regs(y);
                tmem_ld32(taddr + 32, x2);         // in flight while the first 32 columns are folded
                tmem_ld32(taddr + 160, y2);
                if (valid >= TN) fold32<false>(x, y, g * 64, valid, c0, c1);
                else fold32<true>(x, y, g * 64, valid, c0, c1);
                tmem_wait_ld();
                tmem_pin_regs(x2);
                tmem_pin_regs(y2);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(ACC_EMPTY(sc));   // this warp is done reading the accumulator
                if (valid >= TN) fold32<false>(x2, y2, g * 64 + 32, valid, c0, c1);
                else fold32<true>(x2, y2, g * 64 + 32, valid, c0, c1);
                // widen the tile's packed top-2 (two chains x two column halves) into the running keys
#pragma unroll
                for (int ch = 0; ch < 2; ++ch) {
                    merge_min2(widen_acc_key(c0[ch] & 0xFFFFu, cb), widen_acc_key(c1[ch] & 0xFFFFu, cb), b0, b1);
                    merge_min2(widen_acc_key(c0[ch] >> 16, cb + 128), widen_acc_key(c1[ch] >> 16, cb + 128), b0, b1);
                }
                ++c_it;
            }
            const int row = un.row0 + row_in_tile;
            if (row < un.nq) {
                a.keys[(static_cast<size_t>(un.pair) * a.splits + un.split) * 2 * a.max_rows +
                       static_cast<size_t>(g) * a.max_rows + row] = make_uint2(b0, b1);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// =================================================================================================
// v2: CTA pair (cta_group::2).  ncu on the single-CTA kernel (profiles/r01_ncu_full_bf_tc_scan_v1.txt)
// showed the tensor pipe 72.6 % busy and l1tex__m_xbar2l1tex_read_bytes at 42.6 B/clk/SM -- the
// L2 -> SM delivery cap -- because every SM streams the whole 64 KB train tile.  Here two CTAs of a
// cluster share one M = 256 x N = 256 tile: each holds 128 query rows and HALF of the train tile
// (128 rows, 32 KB), the leader CTA issues tcgen05.mma.cta_group::2 which reads both halves from the
// two SMs' shared memory, and each SM's TMEM receives its own 128 x 256 block of keys.
//   full barriers : local per CTA (bulk-copy bytes); the peer's warp 1 relays its completions to the
//                   leader's PEER_*_FULL barriers with a remote mbarrier.arrive.  (Default .cta semantics as
//                   in CUTLASS' ClusterBarrier: a first version with release.cluster / acquire.cluster on
//                   every arrive and wait ran at 2160 cycles per tile instead of 1152.)
//   empty / ready : tcgen05.commit.cta_group::2 ... multicast::cluster to both CTAs
//   accumulator   : all 16 epilogue warps of the pair arrive on the leader's ACC_EMPTY
// =================================================================================================
constexpr int TM2 = 256;                         // query rows per cluster tile
constexpr int kB2Bytes = 128 * kRowBytes;        // this CTA's half of a train tile: 32 KB
constexpr int kB2Stages = 4;
constexpr int kOff2A = 0;                        // 2 x 32 KB
constexpr int kOff2B = 2 * kABytes;              // 4 x 32 KB
constexpr int kOff2AX = kOff2B + kB2Stages * kB2Bytes;   // 4 KB
constexpr int kOff2BX = kOff2AX + kAXBytes;      // 4 KB (128 rows)
constexpr int kOff2Bar = kOff2BX + kAXBytes;     // 24 mbarriers
constexpr int kOff2TmemSlot = kOff2Bar + 32 * 8;
constexpr int kTc2SmemBytes = kOff2TmemSlot + 16;
constexpr uint32_t kIdesc2 = (2u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(TN >> 3) << 17) |
                             (static_cast<uint32_t>(TM2 >> 4) << 24);

struct Unit2 {
    bool valid;
    int pair, split, row0, nq, t_begin, n_btiles, t_end;
    const uint8_t* q;
    const uint8_t* t;
};

// a.tiles counts 256-row cluster tiles here
__device__ __forceinline__ Unit2 decode_unit2(const TcArgs& a, long long u)
{
    Unit2 un;
    const int tile = static_cast<int>(u % a.tiles);
    u /= a.tiles;
    un.split = static_cast<int>(u % a.splits);
    un.pair = static_cast<int>(u / a.splits);
    int nq = a.nq, nt = a.nt;
    un.q = a.exp_q;
    un.t = a.exp_t;
    if (a.pairs != nullptr) {
        const int fq = a.pairs[2 * un.pair], ft = a.pairs[2 * un.pair + 1];
        nq = a.frame_offset[fq + 1] - a.frame_offset[fq];
        nt = a.frame_offset[ft + 1] - a.frame_offset[ft];
        un.q = a.exp_q + static_cast<long long>(fq) * a.frame_pitch_bytes;
        un.t = a.exp_t + static_cast<long long>(ft) * a.frame_pitch_bytes;
    }
    un.nq = nq;
    un.row0 = tile * TM2;
    un.t_begin = min(un.split * a.t_per_split, nt);
    un.t_end = min(un.t_begin + a.t_per_split, nt);
    un.n_btiles = (un.t_end - un.t_begin + TN - 1) / TN;
    un.valid = un.row0 < nq;
    return un;
}

// ---- filtered mode ---------------------------------------------------------------------------
// match_descriptors (src/MapMatcher.cpp:151-161) and match_features (src/features/FeatureExtractor.cpp:25)
// drop every row whose best distance exceeds max_dist, and the ratio test only looks at a second
// distance below max_dist / ratio: train rows farther than that cut can be present or absent in the
// top-2 without changing the result.  The accumulator then holds the raw 64 * dot = 16384 - 128 * d
// (no key slice: 8 MMAs per tile), the epilogue takes the maximum of 16 accumulators with 8
// three-input VIMNMX3 and only when one reaches the cut builds the keys and folds that group.
__device__ __forceinline__ int max3s(int a, int b, int c) { return max(max(a, b), c); }

// ---- epilogue of the CTA-pair kernel -----------------------------------------------------------
// 8 epilogue warps = 4 TMEM lane quadrants x 2 column halves.  Warp (quad, g) reads columns [64 g, +64)
// and [128 + 64 g, +64) of its 32 rows with four tcgen05.ld.x32 issued back to back, frees the accumulator
// stage as soon as they have landed, then folds.  Measured on the way here (2048 pairs of 4000 x 4000,
// filtered mode, scan kernel only):
//   this layout ........................................................ 5.13 ms
//   2 teams x 8 warps alternating tiles, .x16 double-buffered .......... 5.58 ms  (stage held too long)
//   16 warps x 64 columns on the same tile ............................. 5.67 ms
//   same with tcgen05.ld.pack::16b ..................................... 6.24 ms  (packed read-out is slower)
//   N = 128 MMAs into four 128-column stages, 16 warps x 32 columns .... 7.91 ms  (an N = 128 MMA costs 110
//                                                                                 cycles against 134 for N = 256)
// and with the MMAs or the read-out switched off: MMAs alone 4.1 ms (134 cycles per 256 x 256 x 32 MMA = 95 %
// of the int8 peak), read-out alone 3.9 ms: the TMEM -> register path moves ~128 B/clk/SM, i.e. reading a
// tile takes as long as computing it, and the two overlap only partly.
template <bool FILTERED, class ArriveFn>
__device__ __forceinline__ void epilogue_tile2(uint32_t taddr, int g, int cb, int t_end, int thr_raw, uint32_t& b0,
                                               uint32_t& b1, ArriveFn arrive_empty)
{
    uint32_t x[32], y[32], x2[32], y2[32];     // columns 64 g + [0,32), 128 + 64 g + [0,32), 64 g + [32,64), 128 + 64 g + [32,64)
    tmem_ld32(taddr, x);
    tmem_ld32(taddr + 128, y);
    tmem_ld32(taddr + 32, x2);
    tmem_ld32(taddr + 160, y2);
    tmem_wait_ld();
    tmem_pin_regs(x);
    tmem_pin_regs(y);
    tmem_pin_regs(x2);
    tmem_pin_regs(y2);
    tc_fence_before();
    __syncwarp();
    arrive_empty();                    // the accumulator stage is free again before any arithmetic
    const int valid = t_end - cb;
    if constexpr (FILTERED) {
        // raw accumulators 16384 - 128 * d: skip a group of 8 + 8 columns unless one reaches the cut
        uint32_t c0[2] = {0u, 0u}, c1[2] = {0u, 0u};
        auto fold = [&](uint32_t (&lo)[32], uint32_t (&hi)[32], int col0) {
            if (valid < TN) {          // ragged last tile: columns past the train set hold garbage
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    lo[j] = col0 + j < valid ? lo[j] : 0x80000000u;
                    hi[j] = col0 + j + 128 < valid ? hi[j] : 0x80000000u;
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int* ls = reinterpret_cast<const int*>(&lo[8 * q]);
                const int* hs = reinterpret_cast<const int*>(&hi[8 * q]);
                int m = max3s(ls[0], ls[1], ls[2]);
                m = max3s(m, ls[3], ls[4]);
                m = max3s(m, ls[5], ls[6]);
                m = max3s(m, ls[7], hs[0]);
                int n = max3s(hs[1], hs[2], hs[3]);
                n = max3s(n, hs[4], hs[5]);
                n = max3s(n, hs[6], hs[7]);
                if (max(m, n) >= thr_raw) {
                    // rare: a train row of this group is within the cut -- 16-bit keys 32896 - (d << 7 | col & 127),
                    // columns c and c + 128 packed in one register, top-2 with 3 VIMNMX.U16x2 per pair
                    const uint32_t kbase = 16512u - static_cast<uint32_t>(col0 + 8 * q);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        uint32_t kl = lo[8 * q + j] + (kbase - j), kh = hi[8 * q + j] + (kbase - j);
                        if (valid < TN) {
                            kl = col0 + 8 * q + j < valid ? kl : 0u;
                            kh = col0 + 8 * q + j + 128 < valid ? kh : 0u;
                        }
                        const uint32_t p = mad_u32(kh, 65536u, kl);
                        const uint32_t mn = __vminu2(p, c0[j & 1]);
                        c0[j & 1] = __vmaxu2(p, c0[j & 1]);
                        c1[j & 1] = __vmaxu2(c1[j & 1], mn);
                    }
                }
            }
        };
        fold(x, y, g * 64);
        fold(x2, y2, g * 64 + 32);
        if ((c0[0] | c0[1]) != 0u) {       // nothing within the cut in this tile: the running keys stand
#pragma unroll
            for (int ch = 0; ch < 2; ++ch) {
                merge_min2(widen_acc_key(c0[ch] & 0xFFFFu, cb), widen_acc_key(c1[ch] & 0xFFFFu, cb), b0, b1);
                merge_min2(widen_acc_key(c0[ch] >> 16, cb + 128), widen_acc_key(c1[ch] >> 16, cb + 128), b0, b1);
            }
        }
    } else {
        uint32_t c0[2] = {0u, 0u}, c1[2] = {0u, 0u};
        if (valid >= TN) {
            fold32<false>(x, y, g * 64, valid, c0, c1);
            fold32<false>(x2, y2, g * 64 + 32, valid, c0, c1);
        } else {
            fold32<true>(x, y, g * 64, valid, c0, c1);
            fold32<true>(x2, y2, g * 64 + 32, valid, c0, c1);
        }
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            merge_min2(widen_acc_key(c0[ch] & 0xFFFFu, cb), widen_acc_key(c1[ch] & 0xFFFFu, cb), b0, b1);
            merge_min2(widen_acc_key(c0[ch] >> 16, cb + 128), widen_acc_key(c1[ch] >> 16, cb + 128), b0, b1);
        }
    }
}

template <bool FILTERED>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTc2Threads, 1) bf_tc2_scan_kernel(const TcArgs a)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const uint32_t bar = sbase + kOff2Bar;
    auto A_FULL = [&](uint32_t s) { return bar + 8u * s; };                  // [2] local bulk-copy bytes
    auto B_FULL = [&](uint32_t s) { return bar + 8u * (2 + s); };            // [4]
    auto PEER_A_FULL = [&](uint32_t s) { return bar + 8u * (6 + s); };       // [2] leader only: the peer's relay
    auto PEER_B_FULL = [&](uint32_t s) { return bar + 8u * (8 + s); };       // [4]
    auto A_EMPTY = [&](uint32_t s) { return bar + 8u * (12 + s); };          // [2] multicast commit
    auto B_EMPTY = [&](uint32_t s) { return bar + 8u * (14 + s); };          // [4]
    auto ACC_FULL = [&](uint32_t s) { return bar + 8u * (18 + s); };         // [2] multicast commit
    auto ACC_EMPTY = [&](uint32_t s) { return bar + 8u * (20 + s); };        // [2] leader only: 16 epilogue warps
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kOff2TmemSlot);

    if (threadIdx.x == 0) {
        for (uint32_t s = 0; s < 2; ++s) {
            mbar_init(A_FULL(s), 1);
            mbar_init(PEER_A_FULL(s), 1);
            mbar_init(A_EMPTY(s), 1);
            mbar_init(ACC_FULL(s), 1);
            mbar_init(ACC_EMPTY(s), 16);      // 8 epilogue warps in each CTA of the pair
        }
        for (uint32_t s = 0; s < kB2Stages; ++s) {
            mbar_init(B_FULL(s), 1);
            mbar_init(PEER_B_FULL(s), 1);
            mbar_init(B_EMPTY(s), 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // constant key slices: A [1,127,127,1], B [127 - row, 127, 2, 2] (this CTA's 128 train rows: column & 127 = row)
    for (int i = threadIdx.x; i < 2 * kAXBytes / 16; i += kTc2Threads) {
        const bool is_a = i < kAXBytes / 16;
        const int e = is_a ? i : i - kAXBytes / 16;
        const int group = e >> 4, chunk = (e >> 3) & 1, r = group * 8 + (e & 7);
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (chunk == 0) {
            v.x = is_a ? (1u | (127u << 8) | (127u << 16) | (1u << 24))
                       : (static_cast<uint32_t>(127 - r) | (127u << 8) | (2u << 16) | (2u << 24));
        }
        *reinterpret_cast<uint4*>(smem + (is_a ? kOff2AX : kOff2BX) + e * 16) = v;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + kOff2TmemSlot),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();            // both CTAs' barriers initialised before any remote arrive / multicast commit
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const long long n_clusters = gridDim.x >> 1, cid = blockIdx.x >> 1;
    if (warp == 0) {
        // ===== producer: this CTA's 128 query rows and its half of every train tile =====
        if (lane == 0) {
            uint32_t a_it = 0, b_it = 0;
            for (long long u = cid; u < a.n_units; u += n_clusters) {
                const Unit2 un = decode_unit2(a, u);
                if (!un.valid) continue;
                const uint32_t sa = a_it & 1u;
                mbar_wait(A_EMPTY(sa), ((a_it >> 1) & 1u) ^ 1u);
                mbar_expect_tx(A_FULL(sa), kABytes);
                const uint8_t* qs = un.q + static_cast<long long>(un.row0 + rank * 128) * kRowBytes;
#pragma unroll
                for (int c = 0; c < 2; ++c)
                    bulk_g2s(sbase + kOff2A + sa * kABytes + c * 16384, qs + c * 16384, 16384u, A_FULL(sa));
                ++a_it;
                for (int bt = 0; bt < un.n_btiles; ++bt) {
                    const uint32_t sb = b_it & (kB2Stages - 1);
                    mbar_wait(B_EMPTY(sb), ((b_it / kB2Stages) & 1u) ^ 1u);
                    mbar_expect_tx(B_FULL(sb), kB2Bytes);
                    const uint8_t* ts = un.t + static_cast<long long>(un.t_begin + bt * TN + rank * 128) * kRowBytes;
#pragma unroll
                    for (int c = 0; c < 2; ++c)
                        bulk_g2s(sbase + kOff2B + sb * kB2Bytes + c * 16384, ts + c * 16384, 16384u, B_FULL(sb));
                    ++b_it;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            // ===== MMA issuer: one thread of the leader CTA, for the pair =====
            const uint64_t desc_ax = umma_desc(sbase + kOff2AX, 128u, 256u);
            const uint64_t desc_bx = umma_desc(sbase + kOff2BX, 128u, 256u);
            uint32_t a_it = 0, b_it = 0, c_it = 0;
            for (long long u = cid; u < a.n_units; u += n_clusters) {
                const Unit2 un = decode_unit2(a, u);
                if (!un.valid) continue;
                const uint32_t sa = a_it & 1u;
                mbar_wait(A_FULL(sa), (a_it >> 1) & 1u);
                mbar_wait_cluster(PEER_A_FULL(sa), (a_it >> 1) & 1u);
                for (int bt = 0; bt < un.n_btiles; ++bt) {
                    const uint32_t sb = b_it & (kB2Stages - 1), sc = c_it & 1u;
                    mbar_wait_cluster(ACC_EMPTY(sc), ((c_it >> 1) & 1u) ^ 1u);
                    mbar_wait(B_FULL(sb), (b_it / kB2Stages) & 1u);
                    mbar_wait_cluster(PEER_B_FULL(sb), (b_it / kB2Stages) & 1u);
                    tc_fence_after();
                    const uint32_t d = tmem_base + sc * TN;
                    const uint32_t a0 = sbase + kOff2A + sa * kABytes, b0 = sbase + kOff2B + sb * kB2Bytes;
                    if (!SG_TC_VARIANT(a.variant, 4)) {       // timing experiments only: bit 2 skips the MMAs
#pragma unroll
                        for (int k = 0; k < 8; ++k)
                            tc2_mma_i8(d, umma_desc(a0 + k * 256, 128u, kGroupBytes),
                                       umma_desc(b0 + k * 256, 128u, kGroupBytes), kIdesc2, k > 0 ? 1u : 0u);
                        if constexpr (!FILTERED) tc2_mma_i8(d, desc_ax, desc_bx, kIdesc2, 1u);   // the key slice
                    }
                    tc2_commit_mc(B_EMPTY(sb));
                    tc2_commit_mc(ACC_FULL(sc));
                    ++b_it;
                    ++c_it;
                }
                tc2_commit_mc(A_EMPTY(sa));
                ++a_it;
            }
        } else if (lane == 0) {
            // ===== relay (peer CTA): forward this CTA's bulk-copy completions to the leader =====
            uint32_t a_it = 0, b_it = 0;
            for (long long u = cid; u < a.n_units; u += n_clusters) {
                const Unit2 un = decode_unit2(a, u);
                if (!un.valid) continue;
                const uint32_t sa = a_it & 1u;
                mbar_wait(A_FULL(sa), (a_it >> 1) & 1u);
                mbar_arrive_remote(PEER_A_FULL(sa), 0u);
                ++a_it;
                for (int bt = 0; bt < un.n_btiles; ++bt) {
                    const uint32_t sb = b_it & (kB2Stages - 1);
                    mbar_wait(B_FULL(sb), (b_it / kB2Stages) & 1u);
                    mbar_arrive_remote(PEER_B_FULL(sb), 0u);
                    ++b_it;
                }
            }
        }
    } else {
        // ===== epilogue: 8 warps = 4 TMEM lane quadrants x 2 column halves =====
        // each query row ends with 2 partials (one per column half) that the finish kernels merge
        const int quad = warp & 3;
        const int g = (warp - 2) >> 2;
        const int row_in_tile = static_cast<int>(rank) * 128 + quad * 32 + lane;
        uint32_t c_it = 0;
        for (long long u = cid; u < a.n_units; u += n_clusters) {
            const Unit2 un = decode_unit2(a, u);
            if (!un.valid) continue;
            uint32_t b0 = kKeyNone, b1 = kKeyNone;
            for (int bt = 0; bt < un.n_btiles; ++bt, ++c_it) {
                const uint32_t sc = c_it & 1u;
                mbar_wait(ACC_FULL(sc), (c_it >> 1) & 1u);
                tc_fence_after();
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + sc * TN + g * 64;
                auto arrive = [&]() {
                    if (lane == 0) {
                        if (rank == 0) mbar_arrive(ACC_EMPTY(sc));
                        else mbar_arrive_remote(ACC_EMPTY(sc), 0u);
                    }
                };
                if (SG_TC_VARIANT(a.variant, 8)) {          // timing experiments only: bit 3 skips the TMEM reads and the fold
                    tc_fence_before();
                    __syncwarp();
                    arrive();
                } else {
                    epilogue_tile2<FILTERED>(taddr, g, un.t_begin + bt * TN, un.t_end, a.thr_raw, b0, b1, arrive);
                }
            }
            const int row = un.row0 + row_in_tile;
            if (row < un.nq) {
                a.keys[((static_cast<size_t>(un.pair) * a.splits + un.split) * 2 + g) * a.max_rows + row] =
                    make_uint2(b0, b1);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    __syncwarp();
    cluster_sync_all();            // the peer's shared memory and barriers stay alive until both CTAs are done
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// Expansion: N x 32-byte descriptors -> int8 rows (+8 for a set bit, -8 for a clear one) in the UMMA
// canonical layout.  Frame f starts at f * frame_pitch_rows; inside a frame, rows are grouped by 8:
//   byte address = (row / 8) * 2048 + kchunk * 128 + (row % 8) * 16 + byte_in_chunk,  kchunk = 0..15
// One thread = one (row, kchunk) = 16 output bytes from 16 descriptor bits; a warp writes 512
// contiguous bytes.  Rows past the frame's end inside its last 8-row group are zero-filled.
struct ExpandArgs {
    const uint8_t* src;
    const uint8_t* src2 = nullptr; // single pair: the train set, expanded as frame 1 by the same launch
    int n_rows2 = 0;
    const int32_t* frame_list = nullptr;   // batched: expand frame frame_list[blockIdx.y] instead of frame blockIdx.y
    const int32_t* frame_offset;   // null: one set of n_rows rows
    int n_rows;                    // single set
    int rows_per_frame_max;        // batched: upper bound of rows per frame (grid sizing)
    long long frame_pitch_bytes;
    uint8_t* dst;
};

__device__ __forceinline__ uint32_t expand4(uint32_t nibble)
{
    // bit i of the nibble -> byte i = 0x08 if set else 0xF8
    const uint32_t spread = (nibble * 0x00204081u) & 0x01010101u;
    return 0xF8F8F8F8u ^ (spread * 0xF0u);
}

__global__ void __launch_bounds__(256) bf_tc_expand_kernel(const ExpandArgs a)
{
    const int frame = a.frame_list ? a.frame_list[blockIdx.y] : static_cast<int>(blockIdx.y);
    int n = a.n_rows;
    const uint8_t* src = a.src;
    if (a.frame_offset != nullptr) {
        const int o = a.frame_offset[frame];
        n = a.frame_offset[frame + 1] - o;
        src += static_cast<size_t>(o) * 32;
    } else if (frame == 1) {
        n = a.n_rows2;
        src = a.src2;
    }
    const int group = blockIdx.x * 2 + (threadIdx.x >> 7);      // two 8-row groups per CTA
    if (group * 8 >= n) return;
    const int t = threadIdx.x & 127;
    const int row = group * 8 + (t & 7), kchunk = t >> 3;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (row < n) {
        const uint32_t bits = *reinterpret_cast<const uint16_t*>(src + static_cast<size_t>(row) * 32 + kchunk * 2);
        v.x = expand4(bits & 15u);
        v.y = expand4((bits >> 4) & 15u);
        v.z = expand4((bits >> 8) & 15u);
        v.w = expand4(bits >> 12);
    }
    uint8_t* dst = a.dst + static_cast<long long>(frame) * a.frame_pitch_bytes + static_cast<size_t>(group) * kGroupBytes;
    *reinterpret_cast<uint4*>(dst + kchunk * 128 + (t & 7) * 16) = v;
}

} // namespace

int bf_tc_pitch_rows(int max_rows) { return ((max_rows + TN - 1) / TN) * TN; }

size_t bf_tc_expanded_bytes(int n_frames, int max_rows)
{
    return static_cast<size_t>(n_frames) * bf_tc_pitch_rows(max_rows) * kRowBytes;
}

// SLAMGPU_TC_VARIANT bit 1 selects the single-CTA kernel (v1); default is the CTA-pair kernel (v2).
static bool tc_single_cta(const slamgpu_ctx* ctx) { return SG_TC_VARIANT(ctx->tc_variant, 2); }

BfPlan bf_tc_make_plan(const slamgpu_ctx* ctx, int nq_max, int nt_max, int64_t n_pairs)
{
    BfPlan p;
    const int tm = tc_single_cta(ctx) ? TM : TM2;
    const int slots = tc_single_cta(ctx) ? ctx->n_sm : ctx->n_sm / 2;     // CTAs or CTA pairs resident at once
    p.r = 0;
    p.tile = tm;
    p.tiles = max(1, (nq_max + tm - 1) / tm);
    // split the train range only when the query tiles alone cannot occupy every SM
    const int n_btiles = max(1, (nt_max + TN - 1) / TN);
    // Split the train range when the query tiles alone cannot occupy every CTA pair.  A unit costs its train tiles
    // plus about one tile's worth of fixed work (query-tile load, pipeline fill and drain), units run in waves over
    // the resident CTA pairs: take the split count with the cheapest last wave (8000 x 8000: 2 splits = 64 units of
    // 16 tiles in one wave, where "as many units as slots" gave 96 units of 11 tiles in two).
    const int64_t base_units = static_cast<int64_t>(p.tiles) * n_pairs;
    int best_tps = n_btiles;
    int64_t best_cost = -1;
    for (int tps = n_btiles; tps >= 1; --tps) {
        const int s = (n_btiles + tps - 1) / tps;
        const int64_t waves = (base_units * s + slots - 1) / slots;
        const int64_t cost = waves * (tps + 1);
        if (best_cost < 0 || cost < best_cost) {
            best_cost = cost;
            best_tps = tps;
        }
        if (base_units * s >= 8 * static_cast<int64_t>(slots)) break;          // many waves already: finer splits only add overhead
    }
    const int tiles_per_split = best_tps;
    p.t_per_split = tiles_per_split * TN;
    p.splits = max(1, (n_btiles + tiles_per_split - 1) / tiles_per_split);
    return p;
}

// Expand `n_frames` frames (batched: frame_offset on the device) or one set of n_rows rows.
int bf_tc_launch_expand(slamgpu_ctx* ctx, const uint8_t* d_src, const int32_t* d_frame_offset, int n_frames, int n_rows,
                        int max_rows, int pitch_rows, uint8_t* d_dst)
{
    if (n_frames <= 0 || max_rows <= 0) return SLAMGPU_OK;
    SG_CHECK(pitch_rows % TN == 0 && pitch_rows >= max_rows, "bad expanded pitch");
    ExpandArgs a;
    a.src = d_src;
    a.frame_offset = d_frame_offset;
    a.n_rows = n_rows;
    a.rows_per_frame_max = max_rows;
    a.frame_pitch_bytes = static_cast<long long>(pitch_rows) * kRowBytes;
    a.dst = d_dst;
    const int groups = (max_rows + 7) / 8;
    for (int f0 = 0; f0 < n_frames; f0 += 65535) {     // gridDim.y limit
        const int nf = min(65535, n_frames - f0);
        ExpandArgs b = a;
        if (d_frame_offset) b.frame_offset = d_frame_offset + f0;
        b.dst = d_dst + static_cast<long long>(f0) * a.frame_pitch_bytes;
        bf_tc_expand_kernel<<<dim3((groups + 1) / 2, nf), 256, 0, ctx->stream>>>(b);
        ctx->launches++;
    }
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

// the frames a (short) pair list names, one launch
int bf_tc_launch_expand_list(slamgpu_ctx* ctx, const uint8_t* d_src, const int32_t* d_frame_offset, const int32_t* d_frame_list,
                             int n_list, int max_rows, int pitch_rows, uint8_t* d_dst)
{
    if (n_list <= 0 || max_rows <= 0) return SLAMGPU_OK;
    SG_CHECK(pitch_rows % TN == 0 && pitch_rows >= max_rows && n_list <= 65535, "bad expanded pitch / list too long");
    ExpandArgs a;
    a.src = d_src;
    a.frame_offset = d_frame_offset;
    a.frame_list = d_frame_list;
    a.n_rows = 0;
    a.rows_per_frame_max = max_rows;
    a.frame_pitch_bytes = static_cast<long long>(pitch_rows) * kRowBytes;
    a.dst = d_dst;
    const int groups = (max_rows + 7) / 8;
    bf_tc_expand_kernel<<<dim3((groups + 1) / 2, n_list), 256, 0, ctx->stream>>>(a);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

// query set -> frame 0, train set -> frame 1 of d_dst, one launch
int bf_tc_launch_expand_pair(slamgpu_ctx* ctx, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int pitch_rows,
                             uint8_t* d_dst)
{
    SG_CHECK(pitch_rows % TN == 0 && pitch_rows >= max(nq, nt), "bad expanded pitch");
    ExpandArgs a;
    a.src = d_q;
    a.src2 = d_t;
    a.n_rows2 = nt;
    a.frame_offset = nullptr;
    a.n_rows = nq;
    a.rows_per_frame_max = max(nq, nt);
    a.frame_pitch_bytes = static_cast<long long>(pitch_rows) * kRowBytes;
    a.dst = d_dst;
    const int groups = (max(nq, nt) + 7) / 8;
    bf_tc_expand_kernel<<<dim3((groups + 1) / 2, 2), 256, 0, ctx->stream>>>(a);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

int bf_tc_partials(const slamgpu_ctx* ctx) { (void)ctx; return 2; }

// Scan over expanded descriptors.  Writes keys[(pair * splits + split) * P + p][row], P = bf_tc_partials():
// the caller's finish kernels must merge P * plan.splits partials per row.  d_cut: train rows at Hamming distance
// >= d_cut cannot change the caller's result and may be left out of the top-2 (257 = exact kNN-2).
int bf_tc_launch_scan(slamgpu_ctx* ctx, const BfPlan& plan, const uint8_t* d_exp_q, int nq, const uint8_t* d_exp_t,
                      int nt, int pitch_rows, const int32_t* d_frame_offset, const int32_t* d_pairs, int n_pairs,
                      int max_rows, int d_cut, uint2* d_keys)
{
    // the opt-in to > 48 KB of dynamic shared memory is a per-DEVICE function attribute: one flag per context
    // (a context is bound to one device; a process may hold contexts on several)
    if (!ctx->tc_attr_set) {
        SG_CUDA(cudaFuncSetAttribute(bf_tc_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
        SG_CUDA(cudaFuncSetAttribute(bf_tc2_scan_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTc2SmemBytes));
        SG_CUDA(cudaFuncSetAttribute(bf_tc2_scan_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTc2SmemBytes));
        ctx->tc_attr_set = true;
    }
    TcArgs a;
    a.exp_q = d_exp_q;
    a.exp_t = d_exp_t;
    a.frame_pitch_bytes = static_cast<long long>(pitch_rows) * kRowBytes;
    a.frame_offset = d_frame_offset;
    a.pairs = d_pairs;
    a.nq = nq;
    a.nt = nt;
    a.tiles = plan.tiles;
    a.splits = plan.splits;
    a.t_per_split = plan.t_per_split;
    a.max_rows = max_rows;
    a.n_units = static_cast<long long>(plan.tiles) * plan.splits * n_pairs;
    a.variant = ctx->tc_variant;
    a.thr_raw = 16384 - 128 * (d_cut - 1);     // rows at distance >= d_cut may be dropped (257: none)
    a.keys = d_keys;
    if (a.n_units <= 0) return SLAMGPU_OK;
    ProfScope prof(ctx, SLAMGPU_KERNEL_BF_SCAN);
    if (tc_single_cta(ctx)) {
        const unsigned grid = static_cast<unsigned>(min(a.n_units, static_cast<long long>(ctx->n_sm)));
        bf_tc_scan_kernel<<<grid, kTcThreads, kTcSmemBytes, ctx->stream>>>(a);
    } else {
        const unsigned grid = 2u * static_cast<unsigned>(min(a.n_units, static_cast<long long>(ctx->n_sm / 2)));
        if (d_cut <= 256 && !SG_TC_VARIANT(ctx->tc_variant, 16)) {      // __cluster_dims__(2, 1, 1)
            bf_tc2_scan_kernel<true><<<grid, kTc2Threads, kTc2SmemBytes, ctx->stream>>>(a);
        } else {
            bf_tc2_scan_kernel<false><<<grid, kTc2Threads, kTc2SmemBytes, ctx->stream>>>(a);
        }
    }
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

} // namespace slamgpu
