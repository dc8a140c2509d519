// adapter_driver.cpp -- TEST INFRASTRUCTURE.  Reads a world written by tests/test_gpu_adapter.py, builds it
// with the mock data model, calls slam::MapMatcherGpu exactly as the reference's owners call MapMatcher
// (src/Tracker.cpp:232-248, src/Mapper.cpp:208-219, src/LoopDetector.cpp:181) and writes the results.
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <memory>
#include <string>
#include <vector>

#include "Camera.h"
#include "Frame.h"
#include "Map.h"
#include "MapMatcherGpu.h"
#include "MapPoint.h"

using namespace slam;

template <class T> T rd(std::ifstream& f)
{
    T v;
    f.read(reinterpret_cast<char*>(&v), sizeof(T));
    return v;
}

static ExtractedFeatures read_features(std::ifstream& f, int n)
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

static Eigen::Matrix4f read_pose(std::ifstream& f)
{
    Eigen::Matrix4f p;
    for (int i = 0; i < 16; i++) p.m[i] = rd<float>(f);   // column-major
    return p;
}

int main(int argc, char** argv)
{
    if (argc < 3) return 2;
    std::ifstream f(argv[1], std::ios::binary);
    std::ofstream out(argv[2]);
    const float fx = rd<float>(f), fy = rd<float>(f), cx = rd<float>(f), cy = rd<float>(f);
    const int width = rd<int32_t>(f), height = rd<int32_t>(f);
    Camera camera(fx, fy, cx, cy, width, height);

    const int n_kf = rd<int32_t>(f);
    std::vector<std::unique_ptr<KeyFrame>> kfs;
    for (int k = 0; k < n_kf; k++) {
        const int nkp = rd<int32_t>(f);
        auto feats = read_features(f, nkp);
        kfs.push_back(std::make_unique<KeyFrame>(k, feats));
        kfs.back()->set_pose(read_pose(f));
    }
    Map map;
    const int n_points = rd<int32_t>(f);
    for (int i = 0; i < n_points; i++) {
        const float x = rd<float>(f), y = rd<float>(f), z = rd<float>(f);
        MapPoint& p = map.add(Eigen::Vector3f(x, y, z));
        const int n_obs = rd<int32_t>(f);
        for (int o = 0; o < n_obs; o++) {
            const int k = rd<int32_t>(f), idx = rd<int32_t>(f);
            p.add_observation(kfs[k].get(), static_cast<size_t>(idx));
        }
    }
    const int n_kfm = rd<int32_t>(f);
    for (int i = 0; i < n_kfm; i++) {
        const int k = rd<int32_t>(f), kp = rd<int32_t>(f), pi = rd<int32_t>(f);
        kfs[k]->add_map_match(MapPointMatch{map.at(pi), static_cast<size_t>(kp)});
    }
    const int nkp = rd<int32_t>(f);
    auto feats = read_features(f, nkp);
    Frame frame(1000, feats);
    frame.set_pose(read_pose(f));
    const int n_fm = rd<int32_t>(f);
    for (int i = 0; i < n_fm; i++) {
        const int kp = rd<int32_t>(f), pi = rd<int32_t>(f);
        frame.add_map_match(MapPointMatch{map.at(pi), static_cast<size_t>(kp)});
    }
    const int n_fuse = rd<int32_t>(f);
    std::vector<MapPoint*> fuse;
    for (int i = 0; i < n_fuse; i++) {
        const int pi = rd<int32_t>(f);
        fuse.push_back(pi < 0 ? nullptr : &map.at(pi));
    }
    const int required = rd<int32_t>(f);
    if (!f) {
        std::cerr << "short world file\n";
        return 3;
    }

    // point -> index, for printing
    std::unordered_map<const MapPoint*, int> index;
    {
        int i = 0;
        for (auto& p : map) index[&p] = i++;
    }
    // the host's observation iteration order (pointer-keyed unordered_map): the oracle must sum in it
    out << "obs_order " << n_points << "\n";
    for (auto& p : map) {
        out << p.observations().size();
        for (const auto& [kf, idx] : p.observations()) {
            int k = 0;
            while (kfs[k].get() != kf) k++;
            out << " " << k << " " << idx;
        }
        out << "\n";
    }

    try {
        MapMatcherGpu matcher(camera, 64.0F, cv::NORM_HAMMING);
        auto dump = [&](const char* name, const std::vector<MapPointMatch>& ms) {
            out << name << " " << ms.size() << "\n";
            for (const auto& m : ms) out << index[&m.point] << " " << m.keypoint_index << "\n";
        };
        dump("match_map", matcher.match_map(frame, map));
        dump("match_key_frame", matcher.match_key_frame(frame, map, kfs[required].get()));
        dump("match_for_fuse", matcher.match_for_fuse(frame, fuse));
        dump("match_descriptors", matcher.match_descriptors(frame, *kfs[required]));
        {   // every key frame as a loop candidate, one batched call
            std::vector<const KeyFrame*> candidates;
            for (auto& kf : kfs) candidates.push_back(kf.get());
            const auto lists = matcher.match_descriptors_batch(frame, candidates);
            for (size_t c = 0; c < lists.size(); c++) dump(("mdb" + std::to_string(c)).c_str(), lists[c]);
        }
        auto fm = matcher.match_features(kfs[0]->features(), frame.features(), 64.0F);
        out << "match_features " << fm.size() << "\n";
        for (const auto& m : fm) out << m.train_index << " " << m.query_index << "\n";
        // last: it applies its matches to the frame (Tracker::match_with_last_key_frame + match_with_map)
        auto tr = matcher.track(frame, map, *kfs[required]);
        dump("track_key_frame", tr.first);
        dump("track_map", tr.second);
        out << "frame_matches_after_track " << frame.num_map_matches() << "\n";
    } catch (const std::exception& e) {
        std::cerr << e.what() << "\n";
        return 1;
    }
    return 0;
}
