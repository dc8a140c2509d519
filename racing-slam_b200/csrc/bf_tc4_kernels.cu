// bf_tc4_kernels.cu -- filtered brute-force Hamming scan on the block-scaled FP4 tensor cores (sm_100a).
//
// For match_descriptors (src/MapMatcher.cpp:129-163) and match_features
// (src/features/FeatureExtractor.cpp:15-30) only train rows within a distance cut can change the result
// (bf_api.cu: cut_for_filter / cut_for_crosscheck), so the scan only has to FIND those rows.  Two facts
// measured on the int8 CTA-pair kernel (bf_tc_kernels.cu, profiles/r01_ncu_full_bf_tc_scan_v2.txt) shape
// this kernel: the int8 MMAs already run at 95 % of the int8 peak, and reading the 32-bit accumulators
// out of TMEM (~128 B/clk/SM) takes as long as computing them.  Hence
//   * tcgen05.mma.cta_group::2.kind::mxf4.block_scale: every descriptor bit is one E2M1 value (+1 / -1),
//     K = 64 per MMA at the int8 cadence -> twice the comparisons per tensor-pipe cycle;
//   * TWO train rows per accumulator column: the K = 256 of train row cb + c runs with block scale 2^0 and
//     is followed by the K = 256 of train row cb + 240 + c with UE8M0 block scale 2^10 into the same
//     accumulator, acc = dot1 + 1024 * dot2 (exact in fp32, |dot| <= 256) -> half the TMEM read-out per
//     comparison.  The scale factors are constants, so their TMEM columns are filled once with tcgen05.st;
//   * one more K = 64 MMA per tile adds the constant 1.5 * 2^23 + 512 - 1025 * T (T = the dot threshold of the
//     cut), so the bits of the accumulator are the integer u = 0x4B400000 + (dot1 - T + 512) + 1024 * (dot2 - T)
//     and "dot2 >= T", "dot1 >= T" are bit 22 and bit 9 of u: the epilogue ORs 16 accumulators with 8 LOP3 and
//     tests two bits; only on a hit are the two dots decoded and folded as (distance << 22 | train index) keys.
// Layout, warp roles and the barrier protocol are those of the int8 CTA-pair kernel: 2 CTAs share a
// 256-query x 240-column tile (each loads 128 query rows and 120 + 120 train rows), warp 0 = bulk-copy
// producer, warp 1 = MMA issue (leader) / completion relay (peer), warps 2..9 = epilogue.
#include <type_traits>

#include "slamgpu_internal.cuh"
#include "bf_tc_common.cuh"

namespace slamgpu {

namespace {

constexpr int kQ4 = 256;                     // query rows per cluster tile (UMMA M of the pair)
#ifdef SLAMGPU_TC4_N_EXPERIMENT                 // timing experiment (MMA cost against N): results are wrong, debug build only
constexpr int kN4 = SLAMGPU_TC4_N_EXPERIMENT;
#else
constexpr int kN4 = 240;                     // accumulator columns per tile (UMMA N of the pair)
#endif
constexpr int kT4 = 2 * kN4;                 // train rows per tile
constexpr int kRow4 = 128;                   // expanded descriptor: 256 x E2M1 = 128 bytes
constexpr int kGroup4 = 8 * kRow4;           // 8 rows = 8 core matrices of 128 bytes
constexpr int kA4Bytes = 128 * kRow4;        // this CTA's 128 query rows: 16 KB
constexpr int kB4Half = (kN4 / 2) * kRow4;   // this CTA's 120 train rows of one K-half: 15 KB
constexpr int kB4Bytes = 2 * kB4Half;        // both K-halves: 30 KB
constexpr int kB4Stages = 4;
constexpr int kTc4Threads = 64 + 8 * 32;
constexpr int kOff4A = 0;                                  // 2 x 16 KB
constexpr int kOff4B = 2 * kA4Bytes;                       // 4 x 30 KB
constexpr int kOff4AX = kOff4B + kB4Stages * kB4Bytes;     // constant K = 64 slice of A: 128 rows x 32 B
constexpr int kOff4BX = kOff4AX + 128 * 32;                // ... of B: this CTA's 120 rows x 32 B (4 KB reserved)
constexpr int kOff4Bar = kOff4BX + 128 * 32;               // 24 mbarriers
constexpr int kOff4TmemSlot = kOff4Bar + 32 * 8;
constexpr int kOff4Xch = kOff4TmemSlot + 16;               // fused finish: 2 buffers x 128 rows x uint2 (2 KB)
constexpr int kTc4SmemBytes = kOff4Xch + 2 * 128 * 8;
// TMEM columns: accumulator stage 0 at 0, stage 1 at 256, scale factors 1.0 at 496, 2^10 at 504, and the
// constant slice's (2^0 for block 0, 2^7 for block 1) at 240
constexpr uint32_t kAccStride = 256, kSfOne = 496, kSfBig = 504, kSfMix = 240;
constexpr float kMagic = 12582912.0f;        // 1.5 * 2^23: bits(x + kMagic) = 0x4B400000 + x for integer |x| < 2^22
constexpr int kMagicBits = 0x4B400000;

// cute::UMMA::InstrDescriptorBlockScaled: a/b format E2M1 (MXF4Format 1) at [7,10) / [10,13), K-major,
// N >> 3 at [17,23), scale format UE8M0 (1) at bit 23, M >> 4 at [24,29), scale-factor ids 0, K = 64 (bit 31 = 0)
__host__ __device__ constexpr uint32_t idesc4(int n_cols)
{
    return (1u << 7) | (1u << 10) | (static_cast<uint32_t>(n_cols >> 3) << 17) | (1u << 23) |
           (static_cast<uint32_t>(kQ4 >> 4) << 24);
}

__device__ __forceinline__ void tc4_mma(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                        uint32_t accumulate, uint32_t tmem_sfa, uint32_t tmem_sfb)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb)
        : "memory");
}

template <int N>
__device__ __forceinline__ void tmem_pin(uint32_t (&v)[N])
{
#pragma unroll
    for (int i = 0; i < N; i += 8)
        asm volatile("" : "+r"(v[i]), "+r"(v[i + 1]), "+r"(v[i + 2]), "+r"(v[i + 3]), "+r"(v[i + 4]), "+r"(v[i + 5]),
                          "+r"(v[i + 6]), "+r"(v[i + 7])::"memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
// 120 consecutive columns in one statement: five tcgen05.ld off ONE address register (as five statements ptxas
// moved the address into five uniform registers per tile)
__device__ __forceinline__ void tmem_ld120(uint32_t taddr, uint32_t (&a)[32], uint32_t (&b)[32], uint32_t (&c)[32],
                                           uint32_t (&d)[16], uint32_t (&e)[8])
{
    asm volatile(
        "{\n\t.reg .b32 t1, t2, t3, t4;\n\t"
        "add.u32 t1, %120, 32;\n\tadd.u32 t2, %120, 64;\n\tadd.u32 t3, %120, 96;\n\tadd.u32 t4, %120, 112;\n\t"
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%120];\n\t"
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
        "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [t1];\n\t"
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%64, %65, %66, %67, %68, %69, %70, %71, %72, %73, %74, %75, %76, %77, %78, %79, "
        "%80, %81, %82, %83, %84, %85, %86, %87, %88, %89, %90, %91, %92, %93, %94, %95}, [t2];\n\t"
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%96, %97, %98, %99, %100, %101, %102, %103, %104, %105, %106, %107, %108, %109, %110, %111}, [t3];\n\t"
        "tcgen05.ld.sync.aligned.32x32b.x8.b32 "
        "{%112, %113, %114, %115, %116, %117, %118, %119}, [t4];\n\t}"
        : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]), "=r"(a[4]), "=r"(a[5]), "=r"(a[6]), "=r"(a[7]), "=r"(a[8]),
          "=r"(a[9]), "=r"(a[10]), "=r"(a[11]), "=r"(a[12]), "=r"(a[13]), "=r"(a[14]), "=r"(a[15]), "=r"(a[16]),
          "=r"(a[17]), "=r"(a[18]), "=r"(a[19]), "=r"(a[20]), "=r"(a[21]), "=r"(a[22]), "=r"(a[23]), "=r"(a[24]),
          "=r"(a[25]), "=r"(a[26]), "=r"(a[27]), "=r"(a[28]), "=r"(a[29]), "=r"(a[30]), "=r"(a[31]),
          "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3]), "=r"(b[4]), "=r"(b[5]), "=r"(b[6]), "=r"(b[7]), "=r"(b[8]),
          "=r"(b[9]), "=r"(b[10]), "=r"(b[11]), "=r"(b[12]), "=r"(b[13]), "=r"(b[14]), "=r"(b[15]), "=r"(b[16]),
          "=r"(b[17]), "=r"(b[18]), "=r"(b[19]), "=r"(b[20]), "=r"(b[21]), "=r"(b[22]), "=r"(b[23]), "=r"(b[24]),
          "=r"(b[25]), "=r"(b[26]), "=r"(b[27]), "=r"(b[28]), "=r"(b[29]), "=r"(b[30]), "=r"(b[31]),
          "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]), "=r"(c[4]), "=r"(c[5]), "=r"(c[6]), "=r"(c[7]), "=r"(c[8]),
          "=r"(c[9]), "=r"(c[10]), "=r"(c[11]), "=r"(c[12]), "=r"(c[13]), "=r"(c[14]), "=r"(c[15]), "=r"(c[16]),
          "=r"(c[17]), "=r"(c[18]), "=r"(c[19]), "=r"(c[20]), "=r"(c[21]), "=r"(c[22]), "=r"(c[23]), "=r"(c[24]),
          "=r"(c[25]), "=r"(c[26]), "=r"(c[27]), "=r"(c[28]), "=r"(c[29]), "=r"(c[30]), "=r"(c[31]),
          "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]), "=r"(d[8]),
          "=r"(d[9]), "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]), "=r"(d[15]),
          "=r"(e[0]), "=r"(e[1]), "=r"(e[2]), "=r"(e[3]), "=r"(e[4]), "=r"(e[5]), "=r"(e[6]), "=r"(e[7])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, uint32_t x)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(taddr), "r"(x)
                 : "memory");
}

struct Tc4Args {
    const uint8_t* exp_q;          // expanded (E2M1) descriptors of the query side
    const uint8_t* exp_t;          // ... of the train side
    long long frame_pitch_bytes;   // batched: expanded bytes per frame
    const int32_t* frame_offset;   // batched: row offsets (row counts only); null = single pair
    const int32_t* pairs;          // batched: [n_pairs x 2]
    int nq, nt;                    // single pair
    int tiles, splits, t_per_split;
    int max_rows;
    long long n_units;
    int variant;
    int narrow_last;               // size the last train tile of a unit to the rows that are left (SLAMGPU_TC4_NARROW=0: A/B)
    int thr_dot;                   // a train row matters iff dot = 256 - 2 * distance >= thr_dot
    int thr4;                      // thr_dot rounded down to a multiple of 4: what the bit tests of the hot path use
    uint2* keys;
    // FUSED FINISH (splits == 1: a unit sees every train row of its pair).  The two warps that hold the halves of a
    // row's columns merge their partial top-2 through shared memory at the end of the unit and apply the caller's
    // filter (src/MapMatcher.cpp:151-161, as bf_finish_filter_kernel) themselves: out_train[pair][row] and the
    // pair's count are written by the scan, no key records and no finish launch (2 % of a batch step).
    int fuse;
    float max_dist, ratio;
    int32_t* out_train;            // [pair][max_rows]
    int32_t* out_count;            // [pair], zeroed by the launcher; may be null
    uint32_t* col_keys;            // single pair: column minima of the cross-check (see ColOut), set to all-ones by the caller
};

struct Unit4 {
    bool valid, narrow_last;
    int pair, split, row0, nq, nt, t_begin, n_btiles, t_end;
    const uint8_t* q;
    const uint8_t* t;
};

__device__ __forceinline__ Unit4 decode_unit4(const Tc4Args& a, long long u)
{
    Unit4 un;
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
    un.nt = nt;
    un.row0 = tile * kQ4;
    un.t_begin = min(un.split * a.t_per_split, nt);
    un.t_end = min(un.t_begin + a.t_per_split, nt);
    un.n_btiles = (un.t_end - un.t_begin + kT4 - 1) / kT4;
    un.valid = un.row0 < nq;
    un.narrow_last = a.narrow_last != 0;
    return un;
}

// Accumulator columns of train tile bt of a unit.  Every tile but the last has kN4 columns = 2 * kN4 train rows.
// The LAST tile is sized to what is left: rem rows -> ceil(rem / 2) columns rounded up to 16 (the N granularity of
// cta_group::2), its second rows starting n_cols after the first ones.  The kernel is bound by the TMEM read-out
// (about 64 B per clock per SM: 2 bytes per comparison), so reading 80 columns instead of 240 for the last 160
// of 4000 train rows is 7 % of the whole unit's time.
__device__ __forceinline__ int tile_cols(const Unit4& un, int bt)
{
    const int rem = un.t_end - (un.t_begin + bt * kT4);
    return (rem >= kT4 || !un.narrow_last) ? kN4 : min(kN4, (((rem + 1) >> 1) + 15) & ~15);
}

// ---- epilogue arithmetic --------------------------------------------------------------------------
// The constant that the extra MMA adds is  C = 1.5 * 2^23 + 512 - 1025 * T4  (T4 = the dot threshold rounded
// down to a multiple of 4), so the accumulator bits are
//     u = 0x4B400000 + (dot1 - T4 + 512) + 1024 * (dot2 - T4),   0 <= dot1 - T4 + 512 < 1024,
// and the two threshold tests are single bits:  dot2 >= T4 <=> bit 22 of u,  dot1 >= T4 <=> bit 9 of u.
// Hot path per group of 16 accumulators: 8 three-input ORs (LOP3) and one test.  (The first versions took
// per-accumulator FADD + shift + two VIMNMX3 trees: 3 ops instead of 0.5, and the epilogue, not the MMAs,
// set the pace.)  A group with a hit is only PARKED in local memory (4 STL.128); the decoding runs once per
// chunk as a rolled loop over the parked groups -- with the decode unrolled inline (92 KB of SASS) ncu
// attributed 48 % of all stall samples to instruction fetch.
constexpr uint32_t kHitBits = (1u << 22) | (1u << 9);

__device__ __forceinline__ uint32_t or3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xfe;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// N accumulators (columns col0 .. col0 + N - 1 of the tile): scan; a group of 16 that hits is parked in
// `park` (16 ints per group) and its per-element hit bits are added to `hits` (bit 16 q + j = element j of
// group q; N = 8 is the low half of one group)
template <int N>
__device__ __forceinline__ uint64_t scan4(const uint32_t (&v)[N], int4* park)
{
    uint64_t hits = 0ull;
    if constexpr (N % 16 == 0) {
#pragma unroll
        for (int q = 0; q < N / 16; ++q) {
            const uint32_t* x = &v[16 * q];
            uint32_t r = or3(x[0], x[1], x[2]);
            r = or3(r, x[3], x[4]);
            r = or3(r, x[5], x[6]);
            r = or3(r, x[7], x[8]);
            r = or3(r, x[9], x[10]);
            r = or3(r, x[11], x[12]);
            r = or3(r, x[13], x[14]);
            r |= x[15];
            if (__builtin_expect((r & kHitBits) != 0u, 0)) {
                uint32_t hb = 0u;
#pragma unroll
                for (int j = 0; j < 16; ++j) hb |= (x[j] & kHitBits) != 0u ? 1u << j : 0u;
                hits |= static_cast<uint64_t>(hb) << (16 * q);
#pragma unroll
                for (int k = 0; k < 4; ++k) park[4 * q + k] = make_int4(x[4 * k], x[4 * k + 1], x[4 * k + 2], x[4 * k + 3]);
            }
        }
    } else {
        uint32_t r = or3(v[0], v[1], v[2]);
        r = or3(r, v[3], v[4]);
        r = or3(r, v[5], v[6]);
        r |= v[7];
        if (__builtin_expect((r & kHitBits) != 0u, 0)) {
            uint32_t hb = 0u;
#pragma unroll
            for (int j = 0; j < 8; ++j) hb |= (v[j] & kHitBits) != 0u ? 1u << j : 0u;
            hits = hb;
            park[0] = make_int4(v[0], v[1], v[2], v[3]);
            park[1] = make_int4(v[4], v[5], v[6], v[7]);
        }
    }
    return hits;
}

// OR of all N accumulators (N = 8, 16, 32): N / 2 three-input LOP3, chains of 16 independent of each other
template <int N>
__device__ __forceinline__ uint32_t or_all(const uint32_t (&v)[N])
{
    if constexpr (N == 8) {
        uint32_t r = or3(v[0], v[1], v[2]);
        r = or3(r, v[3], v[4]);
        r = or3(r, v[5], v[6]);
        return r | v[7];
    } else {
        uint32_t acc = 0u;
#pragma unroll
        for (int q = 0; q < N / 16; ++q) {
            const uint32_t* x = &v[16 * q];
            uint32_t r = or3(x[0], x[1], x[2]);
            r = or3(r, x[3], x[4]);
            r = or3(r, x[5], x[6]);
            r = or3(r, x[7], x[8]);
            r = or3(r, x[9], x[10]);
            r = or3(r, x[11], x[12]);
            r = or3(r, x[13], x[14]);
            acc = q == 0 ? (r | x[15]) : or3(acc, r, x[15]);
        }
        return acc;
    }
}

// one accumulator with a hit bit: both dots decoded, those within the exact cut folded as
// (distance << 22 | train index); t1 = the train row of dot1, dot2 belongs to t1 + n_cols (240 but for a last tile)
// Cross-check in ONE scan (single pair): every hit also competes for its train row's column minimum,
// atomicMin(col_keys[t], distance << 22 | query row) -- lowest query row on ties, as the column scan would give.  Only
// hits can be mutual nearest neighbours under the distance cut (bf_api.cu: cut_for_crosscheck), and hits are rare.
struct ColOut {
    uint32_t* keys;     // [nt] or null
    int row;            // this thread's query row, -1 when it is a padding row
};

__device__ __forceinline__ void decode4_one(uint32_t u, int t1, int t_end, int thr_dot, int thr4, int n_cols, uint32_t& b0,
                                            uint32_t& b1, const ColOut& colo)
{
    const int x = static_cast<int>(u) - kMagicBits;
    const int dot2 = (x >> 10) + thr4;
    const int dot1 = (x & 1023) - 512 + thr4;
    const int t2 = t1 + n_cols;
    const uint32_t k1 = dot1 >= thr_dot && t1 < t_end
                            ? (static_cast<uint32_t>(256 - dot1) >> 1 << kIdxBits) | static_cast<uint32_t>(t1)
                            : kKeyNone;
    const uint32_t k2 = dot2 >= thr_dot && t2 < t_end
                            ? (static_cast<uint32_t>(256 - dot2) >> 1 << kIdxBits) | static_cast<uint32_t>(t2)
                            : kKeyNone;
    if (colo.keys != nullptr && colo.row >= 0) {
        if (k1 != kKeyNone) atomicMin(&colo.keys[t1], (k1 & ~kIdxMask) | static_cast<uint32_t>(colo.row));
        if (k2 != kKeyNone) atomicMin(&colo.keys[t2], (k2 & ~kIdxMask) | static_cast<uint32_t>(colo.row));
    }
    const uint32_t lo = min(k1, k2), hi = max(k1, k2);
    const uint32_t top = max(lo, b0);
    b0 = min(lo, b0);
    b1 = min(min(b1, top), hi);
}

// one iteration per accumulator that hit (usually one): parked value -> keys
__device__ __forceinline__ void decode4(const int4* park, uint64_t hits, int t1_first, int t_end, int thr_dot, int thr4,
                                        int n_cols, uint32_t& b0, uint32_t& b1, const ColOut& colo)
{
    const uint32_t* p = reinterpret_cast<const uint32_t*>(park);
#pragma unroll 1
    while (hits != 0ull) {
        const int j = __ffsll(static_cast<long long>(hits)) - 1;
        hits &= hits - 1ull;
        decode4_one(p[j], t1_first + j, t_end, thr_dot, thr4, n_cols, b0, b1, colo);
    }
}

// The constant slices.  Element i of a 32-element block holds A = a(i), B = b(i) from the E2M1 values
// {0, 0.5, 1, 1.5, 2, 3, 4, 6} such that the block's sum of a * b is target4 / 4:
//   block at scale 2^14: (3072 - T4 / 4) / 4 = 768 - T4 / 16   ->  1.5 * 2^23 - 1024 * T4
//   block at scale 2^0 : 512 - T4                               (256 .. 756)
// Greedy: n elements of 6 * 6, then 6 * q (q < 6, as 6*q or 6*2 + 6*3), 1 * s (s < 6, as s or 2 + 3) and
// 0.5 * {0.5, 1, 1.5} for the quarters.  Values are returned doubled (0.5 -> 1 ... 6 -> 12).
__device__ __forceinline__ void magic_pair(int target4, int i, int& a2, int& b2)
{
    const int n36 = target4 / 144, rem4 = target4 % 144, ip = rem4 / 4, f = rem4 % 4, q = ip / 6, sv = ip % 6;
    a2 = 0;
    b2 = 0;
    if (i < n36) {
        a2 = 12;
        b2 = 12;
    } else if (i == n36) {
        a2 = 12;
        b2 = 2 * (q == 5 ? 2 : q);
    } else if (i == n36 + 1) {
        a2 = 12;
        b2 = q == 5 ? 6 : 0;
    } else if (i == n36 + 2) {
        a2 = 2;
        b2 = 2 * (sv == 5 ? 2 : sv);
    } else if (i == n36 + 3) {
        a2 = 2;
        b2 = sv == 5 ? 6 : 0;
    } else if (i == n36 + 4) {
        a2 = 1;
        b2 = f;        // 0.5 * {0, 0.5, 1, 1.5}
    }
}
__device__ __forceinline__ uint32_t e2m1_code2(int v2)      // doubled value 0, 1, 2, 3, 4, 6, 8, 12 -> E2M1 bits
{
    return v2 <= 4 ? static_cast<uint32_t>(v2) : v2 == 6 ? 5u : v2 == 8 ? 6u : 7u;
}

// TEAMS = false: every epilogue warp reads half of the columns (120) of EVERY tile.  TEAMS = true: the eight epilogue
// warps form two teams of four (one warp per TMEM lane quadrant); team t owns accumulator stage t, i.e. every other
// tile, and each of its warps reads all 240 columns of its 32 lanes.
template <bool TEAMS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTc4Threads, 1) bf_tc4_scan_kernel(const Tc4Args a)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const uint32_t bar = sbase + kOff4Bar;
    auto A_FULL = [&](uint32_t s) { return bar + 8u * s; };                  // [2] local bulk-copy bytes
    auto B_FULL = [&](uint32_t s) { return bar + 8u * (2 + s); };            // [4]
    auto PEER_A_FULL = [&](uint32_t s) { return bar + 8u * (6 + s); };       // [2] leader only: the peer's relay
    auto PEER_B_FULL = [&](uint32_t s) { return bar + 8u * (8 + s); };       // [4]
    auto A_EMPTY = [&](uint32_t s) { return bar + 8u * (12 + s); };          // [2] multicast commit
    auto B_EMPTY = [&](uint32_t s) { return bar + 8u * (14 + s); };          // [4]
    auto ACC_FULL = [&](uint32_t s) { return bar + 8u * (18 + s); };         // [2] multicast commit
    auto ACC_EMPTY = [&](uint32_t s) { return bar + 8u * (20 + s); };        // [2] leader only: 16 epilogue warps
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kOff4TmemSlot);

    if (threadIdx.x == 0) {
        for (uint32_t s = 0; s < 2; ++s) {
            mbar_init(A_FULL(s), 1);
            mbar_init(PEER_A_FULL(s), 1);
            mbar_init(A_EMPTY(s), 1);
            mbar_init(ACC_FULL(s), 1);
            mbar_init(ACC_EMPTY(s), TEAMS ? 8 : 16);
        }
        for (uint32_t s = 0; s < kB4Stages; ++s) {
            mbar_init(B_FULL(s), 1);
            mbar_init(PEER_B_FULL(s), 1);
            mbar_init(B_EMPTY(s), 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // The constant slice (see magic_pair): one extra K = 64 MMA per tile adds 1.5 * 2^23 + 512 - 1025 * T4, so the
    // epilogue reads integers whose bits 22 and 9 are the two threshold tests -- no FADD, no shift, no min/max.
    // Canonical layout: 8-row group = [k-chunk 0..1][row][16 B]; k-chunk c holds elements 32 c .. 32 c + 31 = block c.
    // block 0 (k-chunk 0, scale 2^0): 512 - T4      block 1 (k-chunk 1, scale 2^7 * 2^7): (768 - T4 / 16) * 2^14
    // (mixing the two magnitudes in one MMA is exact: checked against a two-MMA version on the B200)
    for (int i = threadIdx.x; i < 2 * 128 * 32 / 16; i += kTc4Threads) {
        const bool is_a = i < 128 * 32 / 16;
        const int e = is_a ? i : i - 128 * 32 / 16;
        const int chunk = (e >> 3) & 1;
        const int target4 = chunk == 0 ? 4 * (512 - a.thr4) : 3072 - a.thr4 / 4;
        uint32_t wds[4] = {0u, 0u, 0u, 0u};
#pragma unroll
        for (int n = 0; n < 32; ++n) {          // element n of the block -> nibble n of the 16 bytes
            int av, bv;
            magic_pair(target4, n, av, bv);
            wds[n >> 3] |= e2m1_code2(is_a ? av : bv) << (4 * (n & 7));
        }
        *reinterpret_cast<uint4*>(smem + (is_a ? kOff4AX : kOff4BX) + e * 16) = make_uint4(wds[0], wds[1], wds[2], wds[3]);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + kOff4TmemSlot),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // constant scale factors (UE8M0): 2^0 = 0x7F, 2^10 = 0x89, in every byte of 8 columns each, all 128 lanes
    if (warp >= 2 && warp < 6) {
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        tmem_st8(tmem_base + lane_base + kSfOne, 0x7F7F7F7Fu);
        tmem_st8(tmem_base + lane_base + kSfBig, 0x89898989u);
        tmem_st8(tmem_base + lane_base + kSfMix, 0x867F867Fu);     // byte b = scale of K-block b: 2^0, 2^7 (A and B: 2^14)
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();            // both CTAs' barriers and scale factors are in place
    tc_fence_after();

    const long long n_clusters = gridDim.x >> 1, cid = blockIdx.x >> 1;
    if (warp == 0) {
        // ===== producer: this CTA's 128 query rows; per tile its 120 rows of each K-half =====
        if (lane == 0) {
            uint32_t a_it = 0, b_it = 0;
            for (long long u = cid; u < a.n_units; u += n_clusters) {
                const Unit4 un = decode_unit4(a, u);
                if (!un.valid) continue;
                const uint32_t sa = a_it & 1u;
                mbar_wait(A_EMPTY(sa), ((a_it >> 1) & 1u) ^ 1u);
                mbar_expect_tx(A_FULL(sa), kA4Bytes);
                bulk_g2s(sbase + kOff4A + sa * kA4Bytes, un.q + static_cast<long long>(un.row0 + rank * 128) * kRow4,
                         kA4Bytes, A_FULL(sa));
                ++a_it;
                for (int bt = 0; bt < un.n_btiles; ++bt) {
                    const uint32_t sb = b_it & (kB4Stages - 1);
                    mbar_wait(B_EMPTY(sb), ((b_it / kB4Stages) & 1u) ^ 1u);
                    const int n_cols = tile_cols(un, bt);
                    const uint32_t half_bytes = static_cast<uint32_t>(n_cols >> 1) * kRow4;       // this CTA's rows of one K-half
                    mbar_expect_tx(B_FULL(sb), 2u * half_bytes);
                    const uint8_t* ts = un.t + static_cast<long long>(un.t_begin + bt * kT4 + rank * (n_cols >> 1)) * kRow4;
                    const uint32_t dst = sbase + kOff4B + sb * kB4Bytes;
                    bulk_g2s(dst, ts, half_bytes, B_FULL(sb));                                              // rows cb + c
                    bulk_g2s(dst + kB4Half, ts + static_cast<long long>(n_cols) * kRow4, half_bytes, B_FULL(sb));   // rows cb + n_cols + c
                    ++b_it;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            // ===== MMA issuer: one thread of the leader CTA, for the pair =====
            const uint32_t sf_one = tmem_base + kSfOne, sf_big = tmem_base + kSfBig, sf_mix = tmem_base + kSfMix;
            const uint64_t desc_ax = umma_desc(sbase + kOff4AX, 128u, 256u), desc_bx = umma_desc(sbase + kOff4BX, 128u, 256u);
            uint32_t a_it = 0, b_it = 0, c_it = 0;
            for (long long u = cid; u < a.n_units; u += n_clusters) {
                const Unit4 un = decode_unit4(a, u);
                if (!un.valid) continue;
                const uint32_t sa = a_it & 1u;
                mbar_wait(A_FULL(sa), (a_it >> 1) & 1u);
                mbar_wait(PEER_A_FULL(sa), (a_it >> 1) & 1u);
                for (int bt = 0; bt < un.n_btiles; ++bt) {
                    const uint32_t sb = b_it & (kB4Stages - 1), sc = c_it & 1u;
                    mbar_wait(ACC_EMPTY(sc), ((c_it >> 1) & 1u) ^ 1u);
                    mbar_wait(B_FULL(sb), (b_it / kB4Stages) & 1u);
                    mbar_wait(PEER_B_FULL(sb), (b_it / kB4Stages) & 1u);
                    tc_fence_after();
                    const uint32_t d = tmem_base + sc * kAccStride;
                    const uint32_t a0 = sbase + kOff4A + sa * kA4Bytes, b0 = sbase + kOff4B + sb * kB4Bytes;
                    const int n_cols = tile_cols(un, bt);
                    const uint32_t idesc = idesc4(n_cols);
                    if (!SG_TC_VARIANT(a.variant, 4)) {       // timing experiments only: bit 2 skips the MMAs
                        tc4_mma(d, desc_ax, desc_bx, idesc, 0u, sf_mix, sf_mix);        // D = 1.5 * 2^23 + 512 - 1025 * T4
#pragma unroll
                        for (int k = 0; k < 4; ++k)     // K = 64 values = 32 bytes = two 16-byte k-chunks per MMA
                            tc4_mma(d, umma_desc(a0 + k * 256, 128u, kGroup4), umma_desc(b0 + k * 256, 128u, kGroup4),
                                    idesc, 1u, sf_one, sf_one);
                        // second train row of every column, scaled by 2^10.  A last tile whose first rows already
                        // cover what is left has none: its columns keep dot2 = 0, what zero padding rows would give
                        if (un.t_begin + bt * kT4 + n_cols < un.t_end || SG_TC_VARIANT(a.variant, 256)) {
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                tc4_mma(d, umma_desc(a0 + k * 256, 128u, kGroup4),
                                        umma_desc(b0 + kB4Half + k * 256, 128u, kGroup4), idesc, 1u, sf_one, sf_big);
                        }
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
                const Unit4 un = decode_unit4(a, u);
                if (!un.valid) continue;
                const uint32_t sa = a_it & 1u;
                mbar_wait(A_FULL(sa), (a_it >> 1) & 1u);
                mbar_arrive_remote(PEER_A_FULL(sa), 0u);
                ++a_it;
                for (int bt = 0; bt < un.n_btiles; ++bt) {
                    const uint32_t sb = b_it & (kB4Stages - 1);
                    mbar_wait(B_FULL(sb), (b_it / kB4Stages) & 1u);
                    mbar_arrive_remote(PEER_B_FULL(sb), 0u);
                    ++b_it;
                }
            }
        }
    } else {

        if constexpr (TEAMS) {
        // ===== epilogue, two teams: team t = 4 warps (one per TMEM lane quadrant) owns accumulator stage t =====
        // Measured with the timing build (8192 pairs): MMAs alone 12.2 ms, TMEM reads alone 8.9 ms, but reads +
        // arithmetic 14.5 ms -- a warp cannot keep a tcgen05.ld in flight under its own arithmetic (ptxas schedules
        // the OR tree of the chunk that just landed BEFORE the loads of the next one), and with every warp working
        // on the same tile all of them compute at the same moment while the TMEM read port idles.  With two teams
        // on alternate tiles one team computes while the other one reads.  Per tile a warp reads its 32 lanes x 240
        // columns as chunks of 64 / 64 / 64 / 48 into two register buffers.
        const int quad = warp & 3;
        const int g = (warp - 2) >> 2;                      // team = accumulator stage = partial-key slot
        const int row_in_tile = static_cast<int>(rank) * 128 + quad * 32 + lane;
        const bool skip = SG_TC_VARIANT(a.variant, 8);     // timing experiments only: bit 3 skips the TMEM reads and the fold
        const uint32_t tbase = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(g) * kAccStride;
        uint32_t xa[32], xb[32], ya[32], yb[32], zc[16];
        int4 park[16];
        uint32_t b0 = kKeyNone, b1 = kKeyNone;
        uint32_t c_it = 0;
        ColOut colo{a.col_keys, -1};
        for (long long u = cid; u < a.n_units; u += n_clusters) {
            const Unit4 un = decode_unit4(a, u);
            if (!un.valid) continue;
            const int row = un.row0 + row_in_tile;
            colo.row = row < un.nq ? row : -1;
            for (int bt = 0; bt < un.n_btiles; ++bt, ++c_it) {
                if ((c_it & 1u) != static_cast<uint32_t>(g)) continue;       // the other team's tile
                const int cb = un.t_begin + bt * kT4;
                const int t_end = SG_TC_VARIANT(a.variant, 32) ? -1 : un.t_end;
                const int n_cols = tile_cols(un, bt);
                mbar_wait(ACC_FULL(g), (c_it >> 1) & 1u);
                tc_fence_after();
                if (skip) {
                } else if (n_cols == kN4) {
                    tmem_ld32(tbase, xa);
                    tmem_ld32(tbase + 32, xb);
                    tmem_wait_ld();
                    tmem_ld32(tbase + 64, ya);
                    tmem_ld32(tbase + 96, yb);
                    tmem_pin(xa);
                    tmem_pin(xb);
                    uint64_t m = SG_TC_VARIANT(a.variant, 128) ? 0ull : scan4(xa, park) | (scan4(xb, park + 8) << 32);
                    if (__builtin_expect(m != 0ull && t_end >= 0, 0)) decode4(park, m, cb, t_end, a.thr_dot, a.thr4, kN4, b0, b1, colo);
                    tmem_wait_ld();
                    tmem_ld32(tbase + 128, xa);
                    tmem_ld32(tbase + 160, xb);
                    tmem_pin(ya);
                    tmem_pin(yb);
                    m = SG_TC_VARIANT(a.variant, 128) ? 0ull : scan4(ya, park) | (scan4(yb, park + 8) << 32);
                    if (__builtin_expect(m != 0ull && t_end >= 0, 0)) decode4(park, m, cb + 64, t_end, a.thr_dot, a.thr4, kN4, b0, b1, colo);
                    tmem_wait_ld();
                    tmem_ld32(tbase + 192, ya);
                    tmem_ld16(tbase + 224, zc);
                    tmem_pin(xa);
                    tmem_pin(xb);
                    m = SG_TC_VARIANT(a.variant, 128) ? 0ull : scan4(xa, park) | (scan4(xb, park + 8) << 32);
                    if (__builtin_expect(m != 0ull && t_end >= 0, 0)) decode4(park, m, cb + 128, t_end, a.thr_dot, a.thr4, kN4, b0, b1, colo);
                    tmem_wait_ld();                 // every TMEM read of this stage is done: hand it back before the last fold
                    tmem_pin(ya);
                    tmem_pin(zc);
                } else {
                    // a last tile narrower than kN4: all of its columns, in pieces of 32 / 16 / 8
                    int c = 0;
                    for (; c + 32 <= n_cols; c += 32) {
                        tmem_ld32(tbase + c, xa);
                        tmem_wait_ld();
                        tmem_pin(xa);
                        const uint64_t m = scan4(xa, park);
                        if (__builtin_expect(m != 0ull && t_end >= 0, 0)) decode4(park, m, cb + c, t_end, a.thr_dot, a.thr4, n_cols, b0, b1, colo);
                    }
                    if (c + 16 <= n_cols) {
                        tmem_ld16(tbase + c, zc);
                        tmem_wait_ld();
                        tmem_pin(zc);
                        const uint64_t m = scan4(zc, park);
                        if (__builtin_expect(m != 0ull && t_end >= 0, 0)) decode4(park, m, cb + c, t_end, a.thr_dot, a.thr4, n_cols, b0, b1, colo);
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (rank == 0) mbar_arrive(ACC_EMPTY(g));
                    else mbar_arrive_remote(ACC_EMPTY(g), 0u);
                }
                if (!skip && n_cols == kN4) {
                    const uint64_t m = SG_TC_VARIANT(a.variant, 128) ? 0ull : scan4(ya, park) | (scan4(zc, park + 8) << 32);
                    if (__builtin_expect(m != 0ull && t_end >= 0, 0)) decode4(park, m, cb + 192, t_end, a.thr_dot, a.thr4, kN4, b0, b1, colo);
                }
            }
            // this team's partial top-2 of the unit (all of it when the unit had no tile for this team)
            if (row < un.nq)
                a.keys[((static_cast<size_t>(un.pair) * a.splits + un.split) * 2 + g) * a.max_rows + row] = make_uint2(b0, b1);
            b0 = kKeyNone;
            b1 = kKeyNone;
        }
        } else {
        // ===== epilogue: 8 warps = 4 TMEM lane quadrants x 2 column halves of 120 =====
        // Read FIRST, compute AFTER.  What bounds this kernel is the turn-around of an accumulator stage: with two
        // stages of 240 columns (all of TMEM) the MMAs of tile i + 2 cannot start before tile i has been read out, so
        // the time from "accumulator full" to "accumulator empty" must stay under the MMA time of one tile (9 MMAs of
        // ~158 cycles; measured with the timing build: MMAs alone 11.5 ms per 8192 pairs, this kernel 15.3 ms before).
        // A warp therefore issues the tcgen05.ld of all its 120 columns at once, waits once, hands the stage back and
        // only then folds the 120 accumulators -- round 1 interleaved two chunks with their arithmetic, which kept
        // the OR tree of the first chunk (and a second wait) inside the turn-around.
        const int quad = warp & 3;
        const int g = (warp - 2) >> 2;
        const int row_in_tile = static_cast<int>(rank) * 128 + quad * 32 + lane;
        const uint32_t lane_col = (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(g * 120);
        const int c0 = g * 120;
        const bool skip = SG_TC_VARIANT(a.variant, 8);     // timing experiments only: bit 3 skips the TMEM reads and the fold
        uint32_t va0[32], va1[32], vb0[32], vb1[16], vb2[8];
        int4 park[16];                  // up to 4 parked groups of 16 accumulators (local memory)
        uint32_t b0 = kKeyNone, b1 = kKeyNone;
        uint32_t c_it = 0;
        // this warp's TMEM address in either stage, kept in registers (opaque to the compiler: recomputed per tile it
        // sat between "accumulator full" and the first load)
        uint32_t taddr_s0 = tmem_base + lane_col, taddr_s1 = tmem_base + lane_col + kAccStride;
        asm volatile("" : "+r"(taddr_s0), "+r"(taddr_s1));
        uint2* const xch = reinterpret_cast<uint2*>(smem + kOff4Xch);
        uint32_t u_it = 0;              // valid units seen: parity selects the exchange buffer of the fused finish
        ColOut colo{a.col_keys, -1};
        for (long long u = cid; u < a.n_units; u += n_clusters) {
            const Unit4 un = decode_unit4(a, u);
            if (!un.valid) {
                // a ragged batch: this pair has no query rows in this tile; its rows of the padded output read -1
                if (a.fuse && g == 1 && un.row0 + row_in_tile < a.max_rows)
                    a.out_train[static_cast<size_t>(un.pair) * a.max_rows + un.row0 + row_in_tile] = -1;
                continue;
            }
            colo.row = un.row0 + row_in_tile < un.nq ? un.row0 + row_in_tile : -1;
            for (int bt = 0; bt < un.n_btiles; ++bt, ++c_it) {
                const uint32_t sc = c_it & 1u;
                const uint32_t taddr = sc ? taddr_s1 : taddr_s0;
                const int cb = un.t_begin + bt * kT4;
                const int t_end = SG_TC_VARIANT(a.variant, 32) ? -1 : un.t_end;     // timing experiments only: bit 5 disables the rare path
                const int n_cols = tile_cols(un, bt);
                const bool full = n_cols == kN4;
                mbar_wait(ACC_FULL(sc), (c_it >> 1) & 1u);
                tc_fence_after();
                if (!skip && full) {
                    tmem_ld120(taddr, va0, va1, vb0, vb1, vb2);
                    tmem_wait_ld();
                    tmem_pin(va0);
                    tmem_pin(va1);
                    tmem_pin(vb0);
                    tmem_pin(vb1);
                    tmem_pin(vb2);
                } else if (!skip) {
                    // a last tile narrower than kN4 (16 .. 224 columns): this warp's half of its columns, in pieces
                    // of 32 / 16 / 8, folded as they land
                    const int w = n_cols >> 1;
                    const uint32_t tcol = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + sc * kAccStride +
                                          static_cast<uint32_t>(g * w);
                    const int t_first = cb + g * w;
                    int c = 0;
                    for (; c + 32 <= w; c += 32) {
                        tmem_ld32(tcol + c, vb0);
                        tmem_wait_ld();
                        tmem_pin(vb0);
                        const uint64_t m = scan4(vb0, park);
                        if (__builtin_expect(m != 0ull && t_end >= 0, 0)) decode4(park, m, t_first + c, t_end, a.thr_dot, a.thr4, n_cols, b0, b1, colo);
                    }
                    if (c + 16 <= w) {
                        tmem_ld16(tcol + c, vb1);
                        tmem_wait_ld();
                        tmem_pin(vb1);
                        const uint64_t m = scan4(vb1, park);
                        if (__builtin_expect(m != 0ull && t_end >= 0, 0)) decode4(park, m, t_first + c, t_end, a.thr_dot, a.thr4, n_cols, b0, b1, colo);
                        c += 16;
                    }
                    if (c + 8 <= w) {
                        tmem_ld8(tcol + c, vb2);
                        tmem_wait_ld();
                        tmem_pin(vb2);
                        const uint64_t m = scan4(vb2, park);
                        if (__builtin_expect(m != 0ull && t_end >= 0, 0)) decode4(park, m, t_first + c, t_end, a.thr_dot, a.thr4, n_cols, b0, b1, colo);
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (rank == 0) mbar_arrive(ACC_EMPTY(sc));
                    else mbar_arrive_remote(ACC_EMPTY(sc), 0u);
                }
                if (!skip && full && !SG_TC_VARIANT(a.variant, 128)) {      // timing experiments only: bit 7 skips the scan arithmetic
                    // hot path: ONE test for the warp's 120 accumulators (60 LOP3 + 2 to combine); the per-group scan
                    // with its eight branches runs only when some accumulator has a hit bit
                    const uint32_t any = or3(or3(or_all(va0), or_all(va1), or_all(vb0)), or_all(vb1), or_all(vb2));
                    if (__builtin_expect((any & kHitBits) != 0u && t_end >= 0, 0)) {
                        const uint64_t ma = scan4(va0, park) | (scan4(va1, park + 8) << 32);
                        if (ma != 0ull) decode4(park, ma, cb + c0, t_end, a.thr_dot, a.thr4, kN4, b0, b1, colo);
                        const uint64_t mb = scan4(vb0, park) | (scan4(vb1, park + 8) << 32) | (scan4(vb2, park + 12) << 48);
                        if (mb != 0ull) decode4(park, mb, cb + c0 + 64, t_end, a.thr_dot, a.thr4, kN4, b0, b1, colo);
                    }
                }
            }
            // this warp's partial top-2 of the unit (no keys at all when the unit had no train rows)
            const int row = un.row0 + row_in_tile;
            if (a.fuse) {
                // the warp with the first 120 columns of these 32 rows hands its partial to the warp with the other 120
                // (same quad): named barrier of the two warps, exchange buffer by unit parity (the first warp can be at
                // most one barrier ahead)
                uint2* const slot = xch + (u_it & 1u) * 128 + quad * 32 + lane;
                if (g == 0) *slot = make_uint2(b0, b1);
                asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory");
                if (g == 1) {
                    const uint2 p = *slot;
                    uint32_t m0 = min(b0, p.x);
                    uint32_t m1 = min(min(max(b0, p.x), b1), p.y);         // second smallest of the four keys
                    bool keep = false;
                    int train = -1;
                    if (row < un.nq && m0 != kKeyNone) {
                        const float d0 = static_cast<float>(m0 >> kIdxBits);
                        keep = !(d0 > a.max_dist);
                        if (keep && un.nt >= 2 && m1 != kKeyNone) keep = !(d0 > __fmul_rn(a.ratio, static_cast<float>(m1 >> kIdxBits)));
                        if (keep) train = static_cast<int>(m0 & kIdxMask);
                    }
                    if (row < a.max_rows) a.out_train[static_cast<size_t>(un.pair) * a.max_rows + row] = train;
                    const unsigned ballot = __ballot_sync(0xFFFFFFFFu, keep);
                    if (a.out_count != nullptr && lane == 0 && ballot != 0u) atomicAdd(&a.out_count[un.pair], __popc(ballot));
                }
                ++u_it;
            } else if (row < un.nq) {
                a.keys[((static_cast<size_t>(un.pair) * a.splits + un.split) * 2 + g) * a.max_rows + row] = make_uint2(b0, b1);
            }
            b0 = kKeyNone;
            b1 = kKeyNone;
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

// Expansion: N x 32-byte descriptors -> 256 x E2M1 (0x2 = +1.0 for a set bit, 0xA = -1.0 for a clear one),
// 128 bytes per row, in the UMMA no-swizzle K-major canonical layout: inside a frame, rows are grouped by 8,
//   byte address = (row / 8) * 1024 + word * 128 + (row % 8) * 16 + byte,   word = 0..7 (32 descriptor bits)
// One thread = one (row, word).  Every row up to the frame pitch is written (zeros = +0.0 past the frame's
// end): tiles may read past the end and a zero row has dot 0 with everything, never NaN.
struct Expand4Args {
    const uint8_t* src;
    const int32_t* frame_offset;   // null: one set of n_rows rows (and, with src2, a second one of n_rows2 as frame 1)
    const uint8_t* src2 = nullptr;
    int n_rows2 = 0;
    const int32_t* frame_list = nullptr;   // batched: expand frame frame_list[blockIdx.y] instead of frame blockIdx.y
    int n_rows;
    int pitch_rows;
    uint8_t* dst;
};

__device__ __forceinline__ uint32_t expand8_e2m1(uint32_t x)     // 8 bits -> 8 nibbles
{
    uint32_t t = x & 0xFFu;
    t = (t | (t << 12)) & 0x000F000Fu;
    t = (t | (t << 6)) & 0x03030303u;
    t = (t | (t << 3)) & 0x11111111u;
    return 0xAAAAAAAAu ^ (t << 3);
}

__global__ void __launch_bounds__(256) bf_tc4_expand_kernel(const Expand4Args a)
{
    const int frame = a.frame_list ? a.frame_list[blockIdx.y] : static_cast<int>(blockIdx.y);
    int n = a.n_rows;
    const uint8_t* src = a.src;
    if (a.frame_offset != nullptr) {
        const int o = a.frame_offset[frame];
        n = a.frame_offset[frame + 1] - o;
        src += static_cast<size_t>(o) * 32;
    } else if (frame == 1) {       // the train set of a single pair: both sets expanded by one launch
        n = a.n_rows2;
        src = a.src2;
    }
    const int group = blockIdx.x * 4 + (threadIdx.x >> 6);       // four 8-row groups per CTA
    if (group * 8 >= a.pitch_rows) return;
    const int t = threadIdx.x & 63;
    const int row = group * 8 + (t & 7), word = t >> 3;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (row < n) {
        const uint32_t bits = *reinterpret_cast<const uint32_t*>(src + static_cast<size_t>(row) * 32 + word * 4);
        v.x = expand8_e2m1(bits);
        v.y = expand8_e2m1(bits >> 8);
        v.z = expand8_e2m1(bits >> 16);
        v.w = expand8_e2m1(bits >> 24);
    }
    uint8_t* dst = a.dst + static_cast<long long>(frame) * a.pitch_rows * kRow4 + static_cast<size_t>(group) * kGroup4;
    *reinterpret_cast<uint4*>(dst + word * 128 + (t & 7) * 16) = v;
}

} // namespace

// rows per frame in the expanded buffer: whole query tiles (256) and whole train tiles (480)
int bf_tc4_pitch_rows(int max_rows)
{
    const int a = ((max_rows + kQ4 - 1) / kQ4) * kQ4, b = ((max_rows + kT4 - 1) / kT4) * kT4;
    return max(max(a, b), 8);
}
size_t bf_tc4_expanded_bytes(int n_frames, int max_rows)
{
    return static_cast<size_t>(n_frames) * bf_tc4_pitch_rows(max_rows) * kRow4;
}

BfPlan bf_tc4_make_plan(const slamgpu_ctx* ctx, int nq_max, int nt_max, int64_t n_pairs)
{
    BfPlan p;
    const int slots = ctx->n_sm / 2;
    p.r = 0;
    p.tile = kQ4;
    p.tiles = max(1, (nq_max + kQ4 - 1) / kQ4);
    const int n_btiles = max(1, (nt_max + kT4 - 1) / kT4);
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
    p.t_per_split = tiles_per_split * kT4;
    p.splits = max(1, (n_btiles + tiles_per_split - 1) / tiles_per_split);
    return p;
}

int bf_tc4_launch_expand(slamgpu_ctx* ctx, const uint8_t* d_src, const int32_t* d_frame_offset, int n_frames, int n_rows,
                         int pitch_rows, uint8_t* d_dst)
{
    if (n_frames <= 0 || pitch_rows <= 0) return SLAMGPU_OK;
    SG_CHECK(pitch_rows % 8 == 0, "bad expanded pitch");
    Expand4Args a;
    a.src = d_src;
    a.frame_offset = d_frame_offset;
    a.n_rows = n_rows;
    a.pitch_rows = pitch_rows;
    a.dst = d_dst;
    const int groups = pitch_rows / 8;
    for (int f0 = 0; f0 < n_frames; f0 += 65535) {     // gridDim.y limit
        const int nf = min(65535, n_frames - f0);
        Expand4Args b = a;
        if (d_frame_offset) b.frame_offset = d_frame_offset + f0;
        b.dst = d_dst + static_cast<long long>(f0) * pitch_rows * kRow4;
        bf_tc4_expand_kernel<<<dim3((groups + 3) / 4, nf), 256, 0, ctx->stream>>>(b);
        ctx->launches++;
    }
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

// the frames a (short) pair list names, one launch: CTA row y expands frame d_frame_list[y] (a frame named twice is
// written twice with the same bytes)
int bf_tc4_launch_expand_list(slamgpu_ctx* ctx, const uint8_t* d_src, const int32_t* d_frame_offset, const int32_t* d_frame_list,
                              int n_list, int pitch_rows, uint8_t* d_dst)
{
    if (n_list <= 0 || pitch_rows <= 0) return SLAMGPU_OK;
    SG_CHECK(pitch_rows % 8 == 0 && n_list <= 65535, "bad expanded pitch / list too long");
    Expand4Args a;
    a.src = d_src;
    a.frame_offset = d_frame_offset;
    a.frame_list = d_frame_list;
    a.n_rows = 0;
    a.pitch_rows = pitch_rows;
    a.dst = d_dst;
    bf_tc4_expand_kernel<<<dim3((pitch_rows / 8 + 3) / 4, n_list), 256, 0, ctx->stream>>>(a);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

// query set -> frame 0, train set -> frame 1 of d_dst, one launch
int bf_tc4_launch_expand_pair(slamgpu_ctx* ctx, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int pitch_rows,
                              uint8_t* d_dst)
{
    SG_CHECK(pitch_rows % 8 == 0 && pitch_rows > 0, "bad expanded pitch");
    Expand4Args a;
    a.src = d_q;
    a.frame_offset = nullptr;
    a.src2 = d_t;
    a.n_rows2 = nt;
    a.n_rows = nq;
    a.pitch_rows = pitch_rows;
    a.dst = d_dst;
    bf_tc4_expand_kernel<<<dim3((pitch_rows / 8 + 3) / 4, 2), 256, 0, ctx->stream>>>(a);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

// Filtered scan over E2M1-expanded descriptors; writes 2 partials per (pair, split, row) like the int8 scans.
// d_cut in [1, 256]: train rows at Hamming distance >= d_cut may be left out of the top-2.
int bf_tc4_launch_scan(slamgpu_ctx* ctx, const BfPlan& plan, const uint8_t* d_exp_q, int nq, const uint8_t* d_exp_t,
                       int nt, int pitch_rows, const int32_t* d_frame_offset, const int32_t* d_pairs, int n_pairs,
                       int max_rows, int d_cut, uint2* d_keys, const Tc4Fuse* fuse, uint32_t* d_col_keys)
{
    // per-DEVICE function attribute: one flag per context (see bf_tc_launch_scan)
    if (!ctx->tc4_attr_set) {
        SG_CUDA(cudaFuncSetAttribute(bf_tc4_scan_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTc4SmemBytes));
        SG_CUDA(cudaFuncSetAttribute(bf_tc4_scan_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTc4SmemBytes));
        ctx->tc4_attr_set = true;
    }
    SG_CHECK(d_cut >= 1 && d_cut <= 250, "the fp4 scan needs a distance cut of at most 250");
    Tc4Args a;
    a.exp_q = d_exp_q;
    a.exp_t = d_exp_t;
    a.frame_pitch_bytes = static_cast<long long>(pitch_rows) * kRow4;
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
    a.narrow_last = ctx->tc4_narrow_last ? 1 : 0;
    a.thr_dot = 256 - 2 * (d_cut - 1);                      // -242 .. 256
    a.thr4 = (a.thr_dot + 256) / 4 * 4 - 256;               // rounded DOWN to a multiple of 4: -244 .. 256
    a.keys = d_keys;
    SG_CHECK(d_col_keys == nullptr || d_pairs == nullptr, "column minima are for single-pair scans");
    a.col_keys = d_col_keys;
    a.fuse = 0;
    a.max_dist = a.ratio = 0.0f;
    a.out_train = a.out_count = nullptr;
    if (fuse != nullptr) {
        SG_CHECK(plan.splits == 1 && !ctx->tc4_teams, "the fused finish needs one train split and the default epilogue");
        a.fuse = 1;
        a.max_dist = fuse->max_dist;
        a.ratio = fuse->ratio;
        a.out_train = fuse->d_out_train;
        a.out_count = fuse->d_out_count;
        if (fuse->d_out_count && n_pairs > 0) SG_CUDA(cudaMemsetAsync(fuse->d_out_count, 0, sizeof(int32_t) * n_pairs, ctx->stream));
    }
    if (a.n_units <= 0) return SLAMGPU_OK;
    const unsigned grid = 2u * static_cast<unsigned>(min(a.n_units, static_cast<long long>(ctx->n_sm / 2)));
    ProfScope prof(ctx, SLAMGPU_KERNEL_BF_SCAN);
    if (ctx->tc4_teams) bf_tc4_scan_kernel<true><<<grid, kTc4Threads, kTc4SmemBytes, ctx->stream>>>(a);     // __cluster_dims__(2, 1, 1)
    else bf_tc4_scan_kernel<false><<<grid, kTc4Threads, kTc4SmemBytes, ctx->stream>>>(a);
    ctx->launches++;
    SG_CUDA(cudaGetLastError());
    return SLAMGPU_OK;
}

} // namespace slamgpu
