// adapter_driver.cpp -- TEST INFRASTRUCTURE.  Builds a world with the mock data model and calls slam::MapMatcherGpu
// exactly as the reference's owners call MapMatcher (src/Tracker.cpp:232-248, src/Mapper.cpp:208-219,
// src/LoopDetector.cpp:181).
//   adapter_driver <world.bin> <out.txt>      static checks, then the world's mutation script replayed three times
//                                             (per-call uploads / resident mirror with hooks / resident mirror by
//                                             scanning); every CHECK step dumps the host's map and the results
//   adapter_driver --bench <points> <keypoints> <frames> <out.json>
//                                             a C3-sized mock world: wall time of MapMatcherGpu::track per frame
//                                             with a mutation mix, resident against per-call uploads
#include <array>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <random>
#include <sstream>
#include <string>
#include <vector>

#include "Camera.h"
#include "Frame.h"
#include "Map.h"
#include "LoopDetectorGpu.h"
#include "MapMatcherGpu.h"
#include "MapPoint.h"

using namespace slam;

template <class T> T rd(std::istream& f)
{
    T v;
    f.read(reinterpret_cast<char*>(&v), sizeof(T));
    return v;
}

static ExtractedFeatures read_features(std::istream& f, int n)
{
    ExtractedFeatures e;
    e.keypoints.resize(n);
    for (int i = 0; i < n; i++) {
        e.keypoints[i].pt.x = rd<float>(f);
        e.keypoints[i].pt.y = rd<float>(f);
    }
    e.descriptors = cv::Mat(n, 32);
    if (n) f.read(reinterpret_cast<char*>(e.descriptors.ptr<uint8_t>(0)), size_t(n) * 32);
    return e;
}

static Eigen::Matrix4f read_pose(std::istream& f)
{
    Eigen::Matrix4f p;
    for (int i = 0; i < 16; i++) p.m[i] = rd<float>(f);   // column-major
    return p;
}

static uint32_t bits(float f)
{
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return u;
}

struct Op {
    int code;
    int a = 0, b = 0, c = 0;
    float x = 0, y = 0, z = 0;
    Eigen::Matrix4f pose;
};

struct World {
    std::unique_ptr<Camera> camera;
    std::vector<std::unique_ptr<KeyFrame>> kfs;
    Map map;
    std::vector<MapPoint*> by_id;             // creation id -> point (nullptr once removed)
    ExtractedFeatures frame_feats;
    Eigen::Matrix4f frame_pose;
    std::vector<std::pair<int, int>> frame_matches;   // (keypoint, creation id)
    std::vector<int> fuse;                    // creation ids, -1 = nullptr
    int required = 0;
    std::vector<Op> ops;

    MapPoint& new_point(const Eigen::Vector3f& x)
    {
        MapPoint& p = map.add(x);
        p.creation_id = static_cast<int>(by_id.size());
        by_id.push_back(&p);
        return p;
    }
};

static bool load_world(const char* path, World& w)
{
    std::ifstream f(path, std::ios::binary);
    const float fx = rd<float>(f), fy = rd<float>(f), cx = rd<float>(f), cy = rd<float>(f);
    const int width = rd<int32_t>(f), height = rd<int32_t>(f);
    w.camera = std::make_unique<Camera>(fx, fy, cx, cy, width, height);
    const int n_kf = rd<int32_t>(f);
    for (int k = 0; k < n_kf; k++) {
        const int nkp = rd<int32_t>(f);
        auto feats = read_features(f, nkp);
        w.kfs.push_back(std::make_unique<KeyFrame>(k, feats));
        w.kfs.back()->set_pose(read_pose(f));
    }
    const int n_points = rd<int32_t>(f);
    for (int i = 0; i < n_points; i++) {
        const float x = rd<float>(f), y = rd<float>(f), z = rd<float>(f);
        MapPoint& p = w.new_point(Eigen::Vector3f(x, y, z));
        const int n_obs = rd<int32_t>(f);
        for (int o = 0; o < n_obs; o++) {
            const int k = rd<int32_t>(f), idx = rd<int32_t>(f);
            p.add_observation(w.kfs[k].get(), static_cast<size_t>(idx));
        }
    }
    const int n_kfm = rd<int32_t>(f);
    for (int i = 0; i < n_kfm; i++) {
        const int k = rd<int32_t>(f), kp = rd<int32_t>(f), pi = rd<int32_t>(f);
        w.kfs[k]->add_map_match(MapPointMatch{*w.by_id[pi], static_cast<size_t>(kp)});
    }
    const int nkp = rd<int32_t>(f);
    w.frame_feats = read_features(f, nkp);
    w.frame_pose = read_pose(f);
    const int n_fm = rd<int32_t>(f);
    for (int i = 0; i < n_fm; i++) {
        const int kp = rd<int32_t>(f), pi = rd<int32_t>(f);
        w.frame_matches.push_back({kp, pi});
    }
    const int n_fuse = rd<int32_t>(f);
    for (int i = 0; i < n_fuse; i++) w.fuse.push_back(rd<int32_t>(f));
    w.required = rd<int32_t>(f);
    if (!f) return false;
    const int n_ops = rd<int32_t>(f);
    if (!f) return true;                      // a world without a script
    for (int i = 0; i < n_ops; i++) {
        Op op;
        op.code = rd<int32_t>(f);
        switch (op.code) {
        case 0: op.a = rd<int32_t>(f); op.b = rd<int32_t>(f); op.c = rd<int32_t>(f); break;   // associate kf, point, keypoint
        case 1: op.a = rd<int32_t>(f); op.b = rd<int32_t>(f); break;                          // disassociate kf, point
        case 2: op.a = rd<int32_t>(f); break;                                                 // remove point
        case 3: op.a = rd<int32_t>(f); op.b = rd<int32_t>(f); break;                          // fuse kept, discarded
        case 4: op.a = rd<int32_t>(f); op.x = rd<float>(f); op.y = rd<float>(f); op.z = rd<float>(f); break;   // set_position
        case 5: op.a = rd<int32_t>(f); op.pose = read_pose(f); break;                         // set_pose kf
        case 6: op.x = rd<float>(f); op.y = rd<float>(f); op.z = rd<float>(f); op.a = rd<int32_t>(f); op.b = rd<int32_t>(f); break;   // create_point at, kf, keypoint
        default: break;                                                                       // 7 = check
        }
        w.ops.push_back(op);
    }
    return static_cast<bool>(f);
}

static Frame make_frame(World& w)
{
    Frame frame(1000, w.frame_feats);
    frame.set_pose(w.frame_pose);
    for (auto [kp, pi] : w.frame_matches)
        if (w.by_id[pi] != nullptr) frame.add_map_match(MapPointMatch{*w.by_id[pi], static_cast<size_t>(kp)});
    return frame;
}

static void dump(std::ostream& out, const char* name, const std::vector<MapPointMatch>& ms)
{
    out << name << " " << ms.size() << "\n";
    for (const auto& m : ms) out << m.point.creation_id << " " << m.keypoint_index << "\n";
}

// the host's map as the oracle must see it: live points in map order, observations in the host's iteration order
static void dump_map(std::ostream& out, World& w)
{
    out << "map " << w.map.size() << "\n";
    for (auto& p : w.map) {
        out << p.creation_id << " " << bits(p.position().x()) << " " << bits(p.position().y()) << " " << bits(p.position().z())
            << " " << p.observations().size();
        for (const auto& [kf, idx] : p.observations()) {
            size_t k = 0;
            while (w.kfs[k].get() != kf) k++;
            out << " " << k << " " << idx;
        }
        out << "\n";
    }
    out << "poses " << w.kfs.size() << "\n";
    for (auto& kf : w.kfs) {
        for (int i = 0; i < 16; i++) out << bits(kf->pose().m[i]) << (i == 15 ? "\n" : " ");
    }
}

// mode 0: per-call uploads, 1: resident mirror with hooks, 2: resident mirror by scanning
static std::string run_script(const char* path, int mode, bool with_map_dump)
{
    World w;
    if (!load_world(path, w)) throw std::runtime_error("short world file");
    std::ostringstream out;
    MapMatcherGpu matcher(*w.camera, 64.0F, cv::NORM_HAMMING);
    if (mode == 1) {
        w.map.listener = &matcher;
        matcher.attach(w.map, MirrorMode::Hooks);
    } else if (mode == 2) {
        matcher.attach(w.map, MirrorMode::Scan);
    }
    int step = 0;
    for (const Op& op : w.ops) {
        switch (op.code) {
        case 0: if (w.by_id[op.b]) w.map.associate(*w.kfs[op.a], *w.by_id[op.b], static_cast<size_t>(op.c)); break;
        case 1: if (w.by_id[op.b]) w.map.disassociate(*w.kfs[op.a], *w.by_id[op.b]); break;
        case 2:
            if (w.by_id[op.a]) {
                w.map.remove_point(w.by_id[op.a]);
                w.by_id[op.a] = nullptr;
            }
            break;
        case 3:
            if (w.by_id[op.a] && w.by_id[op.b] && op.a != op.b) {
                w.map.fuse(*w.by_id[op.a], *w.by_id[op.b]);
                w.by_id[op.b] = nullptr;
            }
            break;
        case 4: if (w.by_id[op.a]) w.map.set_position(*w.by_id[op.a], Eigen::Vector3f(op.x, op.y, op.z)); break;
        case 5: w.map.set_pose(*w.kfs[op.a], op.pose); break;
        case 6: {
            MapPoint& p = w.new_point(Eigen::Vector3f(op.x, op.y, op.z));
            w.map.associate(*w.kfs[op.a], p, static_cast<size_t>(op.b));
            break;
        }
        default: {
            out << "check " << step++ << "\n";
            if (with_map_dump) dump_map(out, w);
            {
                Frame frame = make_frame(w);
                dump(out, "match_map", matcher.match_map(frame, w.map));
                dump(out, "match_key_frame", matcher.match_key_frame(frame, w.map, w.kfs[w.required].get()));
                std::vector<MapPoint*> fuse;
                for (int pi : w.fuse) fuse.push_back(pi < 0 ? nullptr : w.by_id[pi]);
                dump(out, "match_for_fuse", matcher.match_for_fuse(frame, fuse));
                auto tr = matcher.track(frame, w.map, *w.kfs[w.required]);
                dump(out, "track_key_frame", tr.first);
                dump(out, "track_map", tr.second);
            }
            if (mode != 0) out << "audit " << matcher.audit() << "\n";
            break;
        }
        }
    }
    if (mode != 0) {
        const auto& s = matcher.mirror_stats();
        std::cerr << "mode " << mode << ": syncs " << s.syncs << " appended " << s.appended << " removed " << s.removed
                  << " observation updates " << s.observation_updates << " position updates " << s.position_updates
                  << " observer updates " << s.observer_updates << " compactions " << s.compactions << " full uploads "
                  << s.full_uploads << " per-call uploads " << s.legacy_uploads << "\n";
        out << "legacy_uploads " << s.legacy_uploads << "\n";
        out << "full_uploads " << s.full_uploads << "\n";
    }
    return out.str();
}

static int run_world(const char* path, const char* out_path)
{
    World w;
    if (!load_world(path, w)) {
        std::cerr << "short world file\n";
        return 3;
    }
    std::ofstream out(out_path);
    const size_t n_points = w.map.size();
    // the host's observation iteration order (pointer-keyed unordered_map): the oracle must sum in it
    out << "obs_order " << n_points << "\n";
    for (auto& p : w.map) {
        out << p.observations().size();
        for (const auto& [kf, idx] : p.observations()) {
            size_t k = 0;
            while (w.kfs[k].get() != kf) k++;
            out << " " << k << " " << idx;
        }
        out << "\n";
    }
    try {
        MapMatcherGpu matcher(*w.camera, 64.0F, cv::NORM_HAMMING);
        Frame frame = make_frame(w);
        dump(out, "match_map", matcher.match_map(frame, w.map));
        dump(out, "match_key_frame", matcher.match_key_frame(frame, w.map, w.kfs[w.required].get()));
        std::vector<MapPoint*> fuse;
        for (int pi : w.fuse) fuse.push_back(pi < 0 ? nullptr : w.by_id[pi]);
        dump(out, "match_for_fuse", matcher.match_for_fuse(frame, fuse));
        dump(out, "match_descriptors", matcher.match_descriptors(frame, *w.kfs[w.required]));
        {   // every key frame as a loop candidate, one batched call
            std::vector<const KeyFrame*> candidates;
            for (auto& kf : w.kfs) candidates.push_back(kf.get());
            const auto lists = matcher.match_descriptors_batch(frame, candidates);
            for (size_t c = 0; c < lists.size(); c++) dump(out, ("mdb" + std::to_string(c)).c_str(), lists[c]);
        }
        auto fm = matcher.match_features(w.kfs[0]->features(), frame.features(), 64.0F);
        out << "match_features " << fm.size() << "\n";
        for (const auto& m : fm) out << m.train_index << " " << m.query_index << "\n";
        // the float order the library assumes against the data model's own project() (here the mock's; in the
        // reference tree: Eigen's)
        out << "self_check " << matcher.self_check() << "\n";
        // last: it applies its matches to the frame (Tracker::match_with_last_key_frame + match_with_map)
        auto tr = matcher.track(frame, w.map, *w.kfs[w.required]);
        dump(out, "track_key_frame", tr.first);
        dump(out, "track_map", tr.second);
        out << "frame_matches_after_track " << frame.num_map_matches() << "\n";

        if (!w.ops.empty()) {
            const std::string hooks = run_script(path, 1, true);
            const std::string scan = run_script(path, 2, true);
            const std::string percall = run_script(path, 0, true);
            // identical apart from the audit / upload-count lines of the resident modes
            auto strip = [](const std::string& s) {
                std::istringstream in(s);
                std::string line, kept;
                while (std::getline(in, line))
                    if (line.rfind("audit", 0) != 0 && line.rfind("legacy_uploads", 0) != 0 && line.rfind("full_uploads", 0) != 0)
                        kept += line + "\n";
                return kept;
            };
            // The three replays run on three separately built worlds: MapPoint::observations() is an unordered_map
            // keyed by KeyFrame POINTERS (src/MapPoint.h:38), so its iteration order -- the order of the float sum
            // in avg_viewing_normal -- may differ between them, exactly as it may between two runs of the reference.
            // Each replay is therefore checked against the oracle fed ITS OWN dumped map; "modes_agree" is information.
            out << "modes_agree " << (strip(hooks) == percall && strip(scan) == percall ? 1 : 0) << "\n";
            std::ofstream(std::string(out_path) + ".hooks") << hooks;
            std::ofstream(std::string(out_path) + ".scan") << scan;
            std::ofstream(std::string(out_path) + ".percall") << percall;
            out << "script\n" << hooks;
        }
    } catch (const std::exception& e) {
        std::cerr << e.what() << "\n";
        return 1;
    }
    return 0;
}

// ---- bench: a C3-sized mock world -----------------------------------------------------------------

static double now_ms()
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static int run_bench(int n_points, int nkp, int n_frames, const char* out_path)
{
    std::mt19937 rng(7);
    auto uni = [&](float lo, float hi) { return std::uniform_real_distribution<float>(lo, hi)(rng); };
    const float fx = 1400.0F, fy = 1400.0F, cx = 960.0F, cy = 540.0F;
    const int width = 1920, height = 1080;
    Camera camera(fx, fy, cx, cy, width, height);
    auto random_desc = [&](uint8_t* d) {
        for (int i = 0; i < 32; i += 4) {
            const uint32_t v = rng();
            std::memcpy(d + i, &v, 4);
        }
    };
    auto make_features = [&](int n) {
        ExtractedFeatures e;
        e.keypoints.resize(n);
        e.descriptors = cv::Mat(n, 32);
        for (int i = 0; i < n; i++) {
            e.keypoints[i].pt.x = std::floor(uni(0, width));
            e.keypoints[i].pt.y = std::floor(uni(0, height));
            random_desc(e.descriptors.ptr<uint8_t>(i));
        }
        return e;
    };
    auto small_pose = [&]() {
        Eigen::Matrix4f p;
        const float a = uni(-0.02F, 0.02F), b = uni(-0.02F, 0.02F);
        p(0, 0) = std::cos(a); p(0, 1) = -std::sin(a); p(1, 0) = std::sin(a); p(1, 1) = std::cos(a); p(2, 2) = 1.0F; p(3, 3) = 1.0F;
        p(0, 2) = b; p(2, 0) = -b;
        p(0, 3) = uni(-0.3F, 0.3F); p(1, 3) = uni(-0.3F, 0.3F); p(2, 3) = uni(-0.3F, 0.3F);
        return p;
    };
    const int n_kf = 200, kf_kp = 2000;
    std::vector<std::unique_ptr<KeyFrame>> kfs;
    for (int k = 0; k < n_kf; k++) {
        kfs.push_back(std::make_unique<KeyFrame>(k, make_features(kf_kp)));
        kfs.back()->set_pose(small_pose());
    }
    Map map;
    auto add_point = [&]() -> MapPoint& {
        const float z = uni(5.0F, 50.0F), u = uni(0, width), v = uni(0, height);
        MapPoint& p = map.add(Eigen::Vector3f((u - cx) / fx * z, (v - cy) / fy * z, z));
        const int n_obs = 2 + static_cast<int>(rng() % 5);
        for (int o = 0; o < n_obs; o++) p.add_observation(kfs[rng() % n_kf].get(), rng() % kf_kp);
        return p;
    };
    for (int i = 0; i < n_points; i++) add_point();
    KeyFrame* last_kf = kfs.back().get();
    for (int i = 0; i < 3000; i++) map.associate(*last_kf, map.at(rng() % map.size()), static_cast<size_t>(i % kf_kp));

    // the tracked frame: a share of its keypoints sits on point projections with an observation's descriptor
    ExtractedFeatures feats = make_features(nkp);
    Eigen::Matrix4f pose = small_pose();
    for (int i = 0; i < nkp / 3; i++) {
        MapPoint& p = map.at(rng() % map.size());
        const Eigen::Vector2f uv = camera.project(pose, p.position());
        if (!(uv.x() >= 0 && uv.x() < width && uv.y() >= 0 && uv.y() < height)) continue;
        feats.keypoints[i].pt.x = std::floor(uv.x());
        feats.keypoints[i].pt.y = std::floor(uv.y());
        const auto& obs = *p.observations().begin();
        std::memcpy(feats.descriptors.ptr<uint8_t>(i), obs.first->descriptor(obs.second).ptr<uint8_t>(0), 32);
    }

    auto mutate = [&](int frame_no) {   // a key-frame insertion every 5th frame: the reference's Mapper::insert
        if (frame_no % 5 != 0) return;
        KeyFrame* kf = kfs[static_cast<size_t>(frame_no / 5) % n_kf].get();
        for (int i = 0; i < 200; i++) {
            MapPoint& p = add_point();
            map.associate(*kf, p, rng() % kf_kp);
        }
        for (int i = 0; i < 300; i++) map.associate(*kf, map.at(rng() % map.size()), rng() % kf_kp);
        for (int i = 0; i < 50; i++) map.remove_point(&map.at(rng() % map.size()));
        for (int i = 0; i < 500; i++) {
            MapPoint& p = map.at(rng() % map.size());
            map.set_position(p, Eigen::Vector3f(p.position().x() + uni(-0.01F, 0.01F), p.position().y(), p.position().z()));
        }
        for (int i = 0; i < 5; i++) {
            KeyFrame* moved = kfs[rng() % n_kf].get();
            Eigen::Matrix4f q = moved->pose();
            q(0, 3) += uni(-0.01F, 0.01F);
            map.set_pose(*moved, q);
        }
    };

    std::ofstream out(out_path);
    out << "{\"points\": " << n_points << ", \"keypoints\": " << nkp << ", \"frames\": " << n_frames;
    size_t matches_resident = 0, matches_percall = 0;
    try {
        for (int mode = 1; mode <= 2; mode++) {      // 1 = hooks, 2 = scan
            MapMatcherGpu matcher(camera, 64.0F, cv::NORM_HAMMING);
            map.listener = mode == 1 ? &matcher : nullptr;
            const double t_attach0 = now_ms();
            matcher.attach(map, mode == 1 ? MirrorMode::Hooks : MirrorMode::Scan);
            const double t_attach = now_ms() - t_attach0;
            std::vector<double> quiet, busy, apply;
            for (int f = 0; f < n_frames + 3; f++) {
                mutate(f);
                Frame frame(f, feats);
                frame.set_pose(pose);
                const double t0 = now_ms();
                auto tr = matcher.track(frame, map, *last_kf);
                // the add_map_match loops are the reference's own host code (O(matches x keypoints)): reported apart
                const double dt = now_ms() - t0 - matcher.mirror_stats().last_apply_ms;
                if (f < 3) continue;                  // warm-up
                (f % 5 == 0 ? busy : quiet).push_back(dt);
                apply.push_back(matcher.mirror_stats().last_apply_ms);
                matches_resident = tr.first.size() + tr.second.size();
            }
            auto mean = [](const std::vector<double>& v) {
                double s = 0;
                for (double x : v) s += x;
                return v.empty() ? 0.0 : s / static_cast<double>(v.size());
            };
            const auto& s = matcher.mirror_stats();
            out << ", \"" << (mode == 1 ? "hooks" : "scan") << "\": {\"attach_ms\": " << t_attach
                << ", \"track_ms_quiet_frame\": " << mean(quiet) << ", \"track_ms_key_frame_insert\": " << mean(busy)
                << ", \"add_map_match_loops_ms\": " << mean(apply)
                << ", \"audit\": " << matcher.audit() << ", \"appended\": " << s.appended << ", \"removed\": " << s.removed
                << ", \"observation_updates\": " << s.observation_updates << ", \"position_updates\": " << s.position_updates
                << ", \"observer_updates\": " << s.observer_updates << ", \"per_call_uploads\": " << s.legacy_uploads << "}";
            map.listener = nullptr;
        }
        {   // per-call uploads: what round 1's adapter did on every call
            MapMatcherGpu matcher(camera, 64.0F, cv::NORM_HAMMING);
            std::vector<double> t;
            for (int f = 0; f < 5; f++) {
                Frame frame(f, feats);
                frame.set_pose(pose);
                const double t0 = now_ms();
                auto tr = matcher.track(frame, map, *last_kf);
                if (f >= 2) t.push_back(now_ms() - t0 - matcher.mirror_stats().last_apply_ms);
                matches_percall = tr.first.size() + tr.second.size();
            }
            double s = 0;
            for (double x : t) s += x;
            out << ", \"per_call_upload_track_ms\": " << s / static_cast<double>(t.size());
        }
    } catch (const std::exception& e) {
        std::cerr << e.what() << "\n";
        return 1;
    }
    out << ", \"final_map_points\": " << map.size() << ", \"matches_resident\": " << matches_resident
        << ", \"matches_per_call\": " << matches_percall << "}\n";
    return 0;
}

// ---- loop detector helper: vocabulary lookups, scoring, PnP model scoring ---------------------------

static uint64_t bits64(double d)
{
    uint64_t u;
    std::memcpy(&u, &d, 8);
    return u;
}

static int run_loop(const char* path, const char* out_path)
{
    std::ifstream f(path, std::ios::binary);
    std::ofstream out(out_path);
    const int n_nodes = rd<int32_t>(f);
    std::vector<int32_t> child_offset(n_nodes + 1);
    f.read(reinterpret_cast<char*>(child_offset.data()), sizeof(int32_t) * child_offset.size());
    std::vector<int32_t> child_ids(static_cast<size_t>(child_offset[n_nodes]));
    f.read(reinterpret_cast<char*>(child_ids.data()), sizeof(int32_t) * child_ids.size());
    std::vector<uint8_t> node_desc(size_t(n_nodes) * 32);
    f.read(reinterpret_cast<char*>(node_desc.data()), node_desc.size());
    std::vector<double> weight(n_nodes);
    f.read(reinterpret_cast<char*>(weight.data()), sizeof(double) * weight.size());
    std::vector<int32_t> word_id(n_nodes);
    f.read(reinterpret_cast<char*>(word_id.data()), sizeof(int32_t) * word_id.size());
    const int n_kf = rd<int32_t>(f);
    std::vector<std::unique_ptr<KeyFrame>> kfs;
    for (int k = 0; k < n_kf; k++) {
        const int nkp = rd<int32_t>(f);
        kfs.push_back(std::make_unique<KeyFrame>(k, read_features(f, nkp)));
    }
    const int n_corr = rd<int32_t>(f);
    std::vector<cv::Point3f> object(n_corr);
    std::vector<cv::Point2f> image(n_corr);
    for (auto& p : object) { p.x = rd<float>(f); p.y = rd<float>(f); p.z = rd<float>(f); }
    for (auto& p : image) { p.x = rd<float>(f); p.y = rd<float>(f); }
    const int n_models = rd<int32_t>(f);
    std::vector<std::array<double, 12>> models(n_models);
    for (auto& m : models)
        for (double& v : m) v = rd<double>(f);
    double K4[4];
    for (double& v : K4) v = rd<double>(f);
    if (!f) {
        std::cerr << "short loop file\n";
        return 3;
    }
    try {
        // shares the matcher's context, as LoopDetector::Impl holds both (src/LoopDetector.cpp:314-330)
        Camera camera(static_cast<float>(K4[0]), static_cast<float>(K4[1]), static_cast<float>(K4[2]), static_cast<float>(K4[3]), 1241, 376);
        MapMatcherGpu matcher(camera, 64.0F, cv::NORM_HAMMING);
        LoopDetectorGpu loop(matcher.context());
        loop.set_vocabulary(n_nodes, child_offset.data(), child_ids.data(), node_desc.data(), weight.data(), word_id.data());
        std::vector<DBoW2::BowVector> bows;
        for (int k = 0; k < n_kf; k++) {
            bows.push_back(loop.bow_of(kfs[k]->features()));            // bow_of, src/LoopDetector.cpp:85-92
            out << "bow " << bows.back().size() << "\n";
            for (const auto& [word, value] : bows.back()) out << word << " " << bits64(value) << "\n";
            if (loop.add_key_frame(bows.back()) != static_cast<size_t>(k)) return 4;
        }
        const std::vector<double> scores = loop.score_against_all(bows.back());   // score_candidates, :346-373
        out << "scores " << scores.size() << "\n";
        for (double s : scores) out << bits64(s) << "\n";
        const auto pnp = loop.score_pnp_models(object, image, models, K4[0], K4[1], K4[2], K4[3]);   // verify_pnp, :203-214
        out << "pnp_counts " << pnp.inlier_counts.size() << "\n";
        for (int32_t c : pnp.inlier_counts) out << c << "\n";
        out << "pnp_best 1\n" << pnp.best << " " << pnp.inliers << "\n";
        out << "pnp_mask " << pnp.inlier_mask.size() << "\n";
        for (uint8_t m : pnp.inlier_mask) out << int(m) << "\n";
        // an empty frame gives an empty vector (:88) and scores 0 against everything
        ExtractedFeatures none;
        out << "empty_bow " << loop.bow_of(none).size() << "\n";
    } catch (const std::exception& e) {
        std::cerr << e.what() << "\n";
        return 1;
    }
    return 0;
}

int main(int argc, char** argv)
{
    if (argc >= 4 && std::strcmp(argv[1], "--loop") == 0) return run_loop(argv[2], argv[3]);
    if (argc >= 6 && std::strcmp(argv[1], "--bench") == 0)
        return run_bench(std::atoi(argv[2]), std::atoi(argv[3]), std::atoi(argv[4]), argv[5]);
    if (argc < 3) return 2;
    return run_world(argv[1], argv[2]);
}
