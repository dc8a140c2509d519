// extern "C" handle around the REFERENCE's slam::KDTree2D (compiled from
// /root/reference/src/KDTree.cpp by oracle/Makefile).  Test infrastructure only: used to pin
// oracle/slam_oracle.cpp's implicit-permutation KD-tree against the real thing.
#include <cstdint>
#include <vector>

#include "KDTree.h"

extern "C" {

void* ref_kd_create(const float* xy, int n)
{
    std::vector<Eigen::Vector2f> pts;
    pts.reserve(n);
    for (int i = 0; i < n; i++) pts.emplace_back(xy[2 * i], xy[2 * i + 1]);
    auto* tree = new slam::KDTree2D();
    tree->build(pts);
    return tree;
}

int ref_kd_radius_search(void* handle, float tx, float ty, float radius, int32_t* out, int cap)
{
    auto* tree = static_cast<slam::KDTree2D*>(handle);
    std::vector<size_t> r = tree->radius_search(Eigen::Vector2f(tx, ty), radius);
    const int n = static_cast<int>(r.size());
    for (int i = 0; i < n && i < cap; i++) out[i] = static_cast<int32_t>(r[i]);
    return n;
}

void ref_kd_destroy(void* handle) { delete static_cast<slam::KDTree2D*>(handle); }
}
