// call_latency.cpp -- the latency of single C-ABI calls as a C++ host (the reference is C++) sees them: no
// interpreter, no array marshalling, steady_clock around the call, copies inside.  bench.py writes the inputs
// (the same synthetic scenes its Python legs use) to a blob file, runs this program and merges the JSON it prints
// into `also.c_abi_call_latency`.
//
//   g++ -O2 -std=c++17 -I include tools/call_latency.cpp -L racing-slam_b200 -lslamgpu
//       -Wl,-rpath,'$ORIGIN/../racing-slam_b200' -o tools/call_latency        (python __graft_entry__.py builds it)
//   tools/call_latency <inputs.blob> [reps = 300]
//
// Blob file: records of  u32 name length | name | u64 byte count | bytes.
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <map>
#include <string>
#include <vector>

#include "slamgpu.h"

using Blob = std::vector<uint8_t>;
static std::map<std::string, Blob> g_in;

static bool load(const char* path)
{
    std::ifstream f(path, std::ios::binary);
    if (!f) return false;
    for (;;) {
        uint32_t nl = 0;
        f.read(reinterpret_cast<char*>(&nl), 4);
        if (!f) break;
        std::string name(nl, '\0');
        f.read(&name[0], nl);
        uint64_t nb = 0;
        f.read(reinterpret_cast<char*>(&nb), 8);
        Blob b(nb);
        f.read(reinterpret_cast<char*>(b.data()), static_cast<std::streamsize>(nb));
        if (!f) return false;
        g_in[name] = std::move(b);
    }
    return true;
}

template <class T> static const T* in(const char* name)
{
    auto it = g_in.find(name);
    if (it == g_in.end()) {
        std::fprintf(stderr, "call_latency: input '%s' missing\n", name);
        std::exit(2);
    }
    return reinterpret_cast<const T*>(it->second.data());
}
static size_t in_bytes(const char* name) { return g_in.count(name) ? g_in[name].size() : 0; }

#define OK(call)                                                                              \
    do {                                                                                      \
        if ((call) != SLAMGPU_OK) {                                                           \
            std::fprintf(stderr, "call_latency: %s failed: %s\n", #call, slamgpu_last_error()); \
            std::exit(1);                                                                     \
        }                                                                                     \
    } while (0)

struct Stat {
    double median_us, p10_us, p90_us, min_us;
};

static Stat time_call(const std::function<void()>& fn, int reps)
{
    for (int i = 0; i < (reps < 20 ? 2 : 20); i++) fn();
    std::vector<double> t(reps);
    for (int i = 0; i < reps; i++) {
        const auto a = std::chrono::steady_clock::now();
        fn();
        const auto b = std::chrono::steady_clock::now();
        t[i] = std::chrono::duration<double, std::micro>(b - a).count();
    }
    std::sort(t.begin(), t.end());
    return {t[reps / 2], t[reps / 10], t[reps * 9 / 10], t[0]};
}

static bool g_first = true;
static void emit(const char* name, const Stat& s, const char* extra = nullptr)
{
    std::printf("%s\"%s\": {\"us_per_call\": %.2f, \"p10_us\": %.2f, \"p90_us\": %.2f, \"min_us\": %.2f%s%s}", g_first ? "" : ", ",
                name, s.median_us, s.p10_us, s.p90_us, s.min_us, extra ? ", " : "", extra ? extra : "");
    g_first = false;
}

template <class T> static T* pinned(size_t n)
{
    void* p = nullptr;
    OK(slamgpu_host_alloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
    return static_cast<T*>(p);
}

int main(int argc, char** argv)
{
    if (argc < 2 || !load(argv[1])) {
        std::fprintf(stderr, "usage: call_latency <inputs.blob> [reps]\n");
        return 2;
    }
    const int reps = argc > 2 ? std::atoi(argv[2]) : 300;
    slamgpu_ctx* ctx = nullptr;
    OK(slamgpu_create(0, &ctx));
    std::printf("{");
    char extra[256];

    // ---- C1: one 2000 x 2000 pair, pageable and pinned caller buffers
    {
        const int n = static_cast<int>(in_bytes("c1_q") / 32), m = static_cast<int>(in_bytes("c1_t") / 32);
        const uint8_t *q = in<uint8_t>("c1_q"), *t = in<uint8_t>("c1_t");
        std::vector<int32_t> oq(n), ot(n);
        int count = 0;
        Stat s = time_call([&] { OK(slamgpu_match_descriptors(ctx, q, n, t, m, 64.0f, 0.75f, oq.data(), ot.data(), &count)); }, reps);
        std::snprintf(extra, sizeof extra, "\"matches\": %d, \"buffers\": \"pageable\"", count);
        emit("c1_match_descriptors_2000x2000", s, extra);
        uint8_t *pq = pinned<uint8_t>(size_t(n) * 32), *pt = pinned<uint8_t>(size_t(m) * 32);
        std::memcpy(pq, q, size_t(n) * 32);
        std::memcpy(pt, t, size_t(m) * 32);
        s = time_call([&] { OK(slamgpu_match_descriptors(ctx, pq, n, pt, m, 64.0f, 0.75f, oq.data(), ot.data(), &count)); }, reps);
        std::snprintf(extra, sizeof extra, "\"matches\": %d, \"buffers\": \"pinned (slamgpu_host_alloc)\"", count);
        emit("c1_match_descriptors_2000x2000_pinned", s, extra);
        s = time_call([&] { OK(slamgpu_match_features(ctx, q, n, t, m, 64.0f, ot.data(), oq.data(), &count)); }, reps);
        std::snprintf(extra, sizeof extra, "\"matches\": %d, \"buffers\": \"pageable\"", count);
        emit("c1_match_features_2000x2000", s, extra);
        slamgpu_host_free(pq);
        slamgpu_host_free(pt);
    }

    // ---- C3's frame-to-frame step: 8000 x 8000 cross-check
    if (in_bytes("c3_prev")) {
        const int n = static_cast<int>(in_bytes("c3_prev") / 32), m = static_cast<int>(in_bytes("c3_cur") / 32);
        std::vector<int32_t> oq(m), ot(m);
        int count = 0;
        Stat s = time_call(
            [&] { OK(slamgpu_match_features(ctx, in<uint8_t>("c3_prev"), n, in<uint8_t>("c3_cur"), m, 64.0f, ot.data(), oq.data(), &count)); },
            std::max(reps / 3, 30));
        std::snprintf(extra, sizeof extra, "\"matches\": %d", count);
        emit("c3_match_features_8000x8000", s, extra);
    }

    // ---- C2: resident frame and point set; per call: pose and skip mask up, proposals down
    {
        const int nkp = static_cast<int>(in_bytes("c2_kp_xy") / 8), np = static_cast<int>(in_bytes("c2_pos") / 12);
        OK(slamgpu_frame_upload(ctx, 1, nkp, in<float>("c2_kp_xy"), in<uint8_t>("c2_kp_desc"), nullptr, in<uint8_t>("c2_kp_matched")));
        OK(slamgpu_points_upload(ctx, 1, np, in<float>("c2_pos"), in<float>("c2_normal"), in<float>("c2_range"),
                                 in<int32_t>("c2_obs_offset"), in<uint8_t>("c2_obs_desc"), nullptr));
        slamgpu_proj_params prm;
        std::memcpy(&prm, in<uint8_t>("c2_params"), sizeof prm);
        std::vector<int32_t> op(nkp), od(nkp);
        for (float radius : {15.0f, 20.0f}) {
            prm.radius = radius;
            Stat s = time_call(
                [&] { OK(slamgpu_match_projection(ctx, 1, 1, in<float>("c2_pose"), &prm, in<uint8_t>("c2_skip"), nullptr, 0, op.data(), od.data())); },
                reps);
            int matches = 0;
            for (int k = 0; k < nkp; k++) matches += op[k] >= 0;
            std::snprintf(extra, sizeof extra, "\"matches\": %d, \"points\": %d, \"keypoints\": %d", matches, np, nkp);
            emit(radius == 15.0f ? "c2_projection_r15" : "c2_projection_r20", s, extra);
        }
        // the sparse form the adapter's resident path uses: is_matched as a list
        std::vector<int32_t> matched_points;
        const uint8_t* skip = in<uint8_t>("c2_skip");
        for (int i = 0; i < np; i++)
            if (skip[i]) matched_points.push_back(i);
        Stat s = time_call(
            [&] {
                OK(slamgpu_match_projection_lists(ctx, 1, 1, in<float>("c2_pose"), &prm, matched_points.data(),
                                                  static_cast<int>(matched_points.size()), nullptr, 0, op.data(), od.data()));
            },
            reps);
        std::snprintf(extra, sizeof extra, "\"matched_points\": %d", static_cast<int>(matched_points.size()));
        emit("c2_projection_lists_r20", s, extra);
        OK(slamgpu_frame_release(ctx, 1));
        OK(slamgpu_points_release(ctx, 1));
    }

    // ---- loop query: the newest key frame against 1 / 5 / 20 / 64 candidates of the resident database
    if (in_bytes("db_desc")) {
        const int rows = *in<int32_t>("db_rows");
        const int n_kf = static_cast<int>(in_bytes("db_desc") / 32 / rows);
        for (int k = 0; k < n_kf; k++) {
            int frame = -1;
            OK(slamgpu_kfdb_append(ctx, 0, in<uint8_t>("db_desc") + size_t(k) * rows * 32, rows, &frame));
        }
        const int64_t cap = int64_t(64) * rows;
        int32_t *oq = pinned<int32_t>(cap), *ot = pinned<int32_t>(cap);
        for (int n_cand : {1, 5, 20, 64}) {
            if (n_cand >= n_kf) continue;
            std::vector<int32_t> pairs, off(n_cand + 1);
            for (int c = 0; c < n_cand; c++) {
                pairs.push_back(n_kf - 1);
                pairs.push_back((c * 7) % (n_kf - 1));
            }
            int64_t total = 0;
            Stat s = time_call(
                [&] { OK(slamgpu_kfdb_match_pairs(ctx, 0, pairs.data(), n_cand, 64.0f, 0.75f, off.data(), oq, ot, cap, &total)); },
                std::max(reps / 2, 30));
            std::snprintf(extra, sizeof extra, "\"matches\": %lld, \"rows_per_key_frame\": %d", static_cast<long long>(total), rows);
            char name[64];
            std::snprintf(name, sizeof name, "loop_query_%d_candidates", n_cand);
            emit(name, s, extra);
        }
        OK(slamgpu_kfdb_release(ctx, 0));
        slamgpu_host_free(oq);
        slamgpu_host_free(ot);
    }
    std::printf("}\n");
    slamgpu_destroy(ctx);
    return 0;
}
