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
constexpr uint32_t kKeyBias = 32896u;      // accumulator = kKeyBias - (distance << 7 | column & 127), in [1, 32896]

// shared-memory map (dynamic, 1024-byte aligned base)
constexpr int kOffA = 0;                         // 2 x 32 KB
constexpr int kOffB = 2 * kABytes;               // 2 x 64 KB
constexpr int kOffAX = kOffB + 2 * kBBytes;      // 4 KB
constexpr int kOffBX = kOffAX + kAXBytes;        // 8 KB
constexpr int kOffBar = kOffBX + kBXBytes;       // 12 mbarriers
constexpr int kOffTmemSlot = kOffBar + 16 * 8;
constexpr int kTcSmemBytes = kOffTmemSlot + 16;

__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) {
    }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], int8 x int8 -> int32, issued by ONE thread for the CTA
__device__ __forceinline__ void tc_mma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 32 consecutive TMEM columns of this thread's lane (lane = 32 * (warp % 4) + laneid)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// The registers of a tcgen05.ld are only defined after tcgen05.wait::ld, a dependence the compiler
// cannot see: pass them through an empty volatile asm placed after the wait so no use moves above it.
__device__ __forceinline__ void tmem_pin_regs(uint32_t (&v)[32])
{
#pragma unroll
    for (int i = 0; i < 32; i += 8)
        asm volatile("" : "+r"(v[i]), "+r"(v[i + 1]), "+r"(v[i + 2]), "+r"(v[i + 3]), "+r"(v[i + 4]), "+r"(v[i + 5]),
                          "+r"(v[i + 6]), "+r"(v[i + 7])::"memory");
}

// UMMA shared-memory matrix descriptor, no swizzle, K-major (cute::UMMA::SmemDescriptor):
//   bits [0,14)  start address >> 4      bits [16,30) leading-dimension byte offset >> 4
//   bits [32,46) stride byte offset >> 4 bits [46,48) version = 1 (sm_100)   bits [61,64) layout = 0
// Canonical K-major INTERLEAVE layout, in 16-byte units: ((8,n),2):((1,SBO),LBO) -- a core matrix is
// 8 rows x 16 bytes stored contiguously; LBO steps to the next 16 bytes of K, SBO to the next 8 rows.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo)
{
    return static_cast<uint64_t>((saddr & 0x3FFFFu) >> 4) | (static_cast<uint64_t>(lbo >> 4) << 16) |
           (static_cast<uint64_t>(sbo >> 4) << 32) | (1ull << 46);
}
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

__device__ __forceinline__ uint32_t mad_u32(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
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
            const bool swap = (a.variant & 1) != 0;
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

// Expansion: N x 32-byte descriptors -> int8 rows (+8 for a set bit, -8 for a clear one) in the UMMA
// canonical layout.  Frame f starts at f * frame_pitch_rows; inside a frame, rows are grouped by 8:
//   byte address = (row / 8) * 2048 + kchunk * 128 + (row % 8) * 16 + byte_in_chunk,  kchunk = 0..15
// One thread = one (row, kchunk) = 16 output bytes from 16 descriptor bits; a warp writes 512
// contiguous bytes.  Rows past the frame's end inside its last 8-row group are zero-filled.
struct ExpandArgs {
    const uint8_t* src;
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
    const int frame = blockIdx.y;
    int n = a.n_rows;
    const uint8_t* src = a.src;
    if (a.frame_offset != nullptr) {
        const int o = a.frame_offset[frame];
        n = a.frame_offset[frame + 1] - o;
        src += static_cast<size_t>(o) * 32;
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

BfPlan bf_tc_make_plan(int n_sm, int nq_max, int nt_max, int64_t n_pairs)
{
    BfPlan p;
    p.r = 0;
    p.tile = TM;
    p.tiles = max(1, (nq_max + TM - 1) / TM);
    // split the train range only when the query tiles alone cannot occupy every SM
    const int64_t units = static_cast<int64_t>(p.tiles) * n_pairs;
    const int n_btiles = max(1, (nt_max + TN - 1) / TN);
    int64_t splits = units >= n_sm ? 1 : (n_sm + units - 1) / units;
    splits = min(splits, static_cast<int64_t>(n_btiles));
    const int tiles_per_split = static_cast<int>((n_btiles + splits - 1) / splits);
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

// Scan over expanded descriptors.  Writes keys[(pair * splits + split) * 2 + g][row]: the caller's
// finish kernels must merge 2 * plan.splits partials per row.
int bf_tc_launch_scan(slamgpu_ctx* ctx, const BfPlan& plan, const uint8_t* d_exp_q, int nq, const uint8_t* d_exp_t,
                      int nt, int pitch_rows, const int32_t* d_frame_offset, const int32_t* d_pairs, int n_pairs,
                      int max_rows, uint2* d_keys)
{
    static bool attr_set = false;
    if (!attr_set) {
        SG_CUDA(cudaFuncSetAttribute(bf_tc_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
        attr_set = true;
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
    a.keys = d_keys;
    if (a.n_units <= 0) return SLAMGPU_OK;
    const unsigned grid = static_cast<unsigned>(min(a.n_units, static_cast<long long>(ctx->n_sm)));
    ProfScope prof(ctx, SLAMGPU_KERNEL_BF_SCAN);
    bf_tc_scan_kernel<<<grid, kTcThreads, kTcSmemBytes, ctx->stream>>>(a);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

} // namespace slamgpu
