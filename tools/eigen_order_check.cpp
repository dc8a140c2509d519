// eigen_order_check.cpp -- the pin for the float OPERATION ORDER of the projection path.
//
// libslamgpu restates four Eigen expressions of the reference in a fixed order of roundings (explicit _rn
// intrinsics on the device, -ffp-contract=off on the host); the GPU results are bit-identical to that
// restatement (tests/test_gpu_proj.py).  Whether the restatement is the order YOUR Eigen evaluates cannot be
// checked in the build container (Eigen is not installed there), so this program does it: it evaluates the
// reference's expressions VERBATIM with real Eigen and compares bit patterns with the library's host
// restatements (slamgpu_host_project / _camera_center / _point_stats, no GPU needed) on seeded inputs,
// including projections at u >= 1024 px where one float ulp (1.2e-4 px) exceeds the 1e-4 px tolerance.
//
//   g++ -O2 -std=c++17 -ffp-contract=off -I/path/to/eigen3 -I<repo>/include tools/eigen_order_check.cpp
//       -L<repo>/racing-slam_b200 -lslamgpu -Wl,-rpath,<repo>/racing-slam_b200 -o eigen_order_check
//   ./eigen_order_check [n = 200000] [seed = 1]
//
// Build it with the flags of the reference build (its CMakeLists.txt sets no -march and no -ffast-math; GCC on
// x86-64 then emits no FMA; on aarch64 GCC contracts by default: pass -ffp-contract=off there, to both).
// Exit code 0 and "0 mismatches" on every line = the derived order is Eigen's; anything else names the
// expression whose order differs, with the first offending input, and parity of the projection path
//  is then NOT bit-exact on that toolchain.
//
// The expressions, verbatim from the reference tree:
//   src/Camera.cpp:27-31    Eigen::Vector3f uv = m_K * pose.block<3, 4>(0, 0) * point.homogeneous();
//                           if (uv.z() < 0) return Eigen::Vector2f::Ones() * -1;  return uv.hnormalized();
//   src/Frame.cpp:41        return -m_pose.block<3, 3>(0, 0).transpose() * m_pose.block<3, 1>(0, 3);
//   src/MapPoint.cpp:27-32  for (obs) { C = frame->camera_center(); v = (m_position - C).normalized();
//                           normal += v.normalized(); }  return normal.normalized();
//   src/MapPoint.cpp:37-44  distance = (m_position - frame->camera_center()).norm(); min / max
//   src/MapMatcher.cpp:62-70  ray = position - camera_center; viewing_normal.dot(ray.normalized()); ray.norm()
#include <Eigen/Dense>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <random>
#include <vector>

#include "slamgpu.h"

static uint32_t bits(float f)
{
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return u;
}

// ---- the reference's expressions, verbatim ----------------------------------------------------------
static Eigen::Vector2f ref_project(const Eigen::Matrix3f& m_K, const Eigen::Matrix4f& pose, const Eigen::Vector3f& point)
{
    Eigen::Vector3f uv = m_K * pose.block<3, 4>(0, 0) * point.homogeneous();
    if (uv.z() < 0) {
        return Eigen::Vector2f::Ones() * -1; // Return invalid point
    }
    return uv.hnormalized();
}

static Eigen::Vector3f ref_camera_center(const Eigen::Matrix4f& m_pose)
{
    return -m_pose.block<3, 3>(0, 0).transpose() * m_pose.block<3, 1>(0, 3);
}

static Eigen::Vector3f ref_avg_viewing_normal(const Eigen::Vector3f& m_position, const std::vector<Eigen::Vector3f>& centers)
{
    Eigen::Vector3f normal = Eigen::Vector3f::Zero();
    for (const auto& center : centers) {
        Eigen::Vector3f C = center;
        Eigen::Vector3f v = (m_position - C).normalized();
        normal += v.normalized();
    }
    return normal.normalized();
}

static std::pair<float, float> ref_observed_distance_range(const Eigen::Vector3f& m_position,
                                                           const std::vector<Eigen::Vector3f>& centers)
{
    float nearest = std::numeric_limits<float>::max();
    float furthest = 0.0F;
    for (const auto& center : centers) {
        float distance = (m_position - center).norm();
        nearest = std::min(nearest, distance);
        furthest = std::max(furthest, distance);
    }
    return {nearest, furthest};
}

int main(int argc, char** argv)
{
    const int n = argc > 1 ? std::atoi(argv[1]) : 200000;
    const unsigned seed = argc > 2 ? static_cast<unsigned>(std::atoi(argv[2])) : 1u;
    std::mt19937 rng(seed);
    auto uni = [&](float lo, float hi) { return std::uniform_real_distribution<float>(lo, hi)(rng); };

    // two cameras: KITTI-sized (u < 1241) and 1080p (u up to 1920: one ulp = 1.2e-4 px above 1024)
    const float cams[2][6] = {{718.856F, 718.856F, 607.19F, 185.22F, 1241, 376}, {1400.0F, 1400.0F, 960.0F, 540.0F, 1920, 1080}};
    long bad_project = 0, bad_center = 0, bad_normal = 0, bad_range = 0, bad_gate = 0, above_1024 = 0;
    for (int i = 0; i < n; i++) {
        const float* cam = cams[i & 1];
        Eigen::Matrix3f K = Eigen::Matrix3f::Identity();
        K(0, 0) = cam[0];
        K(1, 1) = cam[1];
        K(0, 2) = cam[2];
        K(1, 2) = cam[3];
        slamgpu_proj_params prm;
        std::memset(&prm, 0, sizeof(prm));
        prm.fx = cam[0]; prm.fy = cam[1]; prm.cx = cam[2]; prm.cy = cam[3];
        prm.width = static_cast<int32_t>(cam[4]); prm.height = static_cast<int32_t>(cam[5]);
        // pose: a rotation (angle-axis through Eigen, rounded to float) and a translation
        const Eigen::Vector3f axis = Eigen::Vector3f(uni(-1, 1), uni(-1, 1), uni(-1, 1)).normalized();
        const Eigen::Matrix3f R = Eigen::AngleAxisf(uni(-0.6F, 0.6F), axis).toRotationMatrix();
        Eigen::Matrix4f pose = Eigen::Matrix4f::Identity();
        pose.block<3, 3>(0, 0) = R;
        pose.block<3, 1>(0, 3) = Eigen::Vector3f(uni(-3, 3), uni(-3, 3), uni(-3, 3));
        // a point that projects somewhere in (and a little around) the image; one in 16 behind the camera
        const float z = (i % 16 == 0 ? -1.0F : 1.0F) * uni(2.0F, 60.0F);
        const float u = uni(-50.0F, cam[4] + 50.0F), v = uni(-50.0F, cam[5] + 50.0F);
        const Eigen::Vector3f Xc((u - cam[2]) / cam[0] * z, (v - cam[3]) / cam[1] * z, z);
        const Eigen::Vector3f X = R.transpose() * (Xc - pose.block<3, 1>(0, 3));

        // P1  Camera::project
        const Eigen::Vector2f uv = ref_project(K, pose, X);
        float lib_uv[2];
        slamgpu_host_project(pose.data(), &prm, X.data(), lib_uv);
        if (bits(uv.x()) != bits(lib_uv[0]) || bits(uv.y()) != bits(lib_uv[1])) {
            if (bad_project++ == 0)
                std::printf("  first project mismatch at i=%d: Eigen (%.9g, %.9g) library (%.9g, %.9g)\n", i, uv.x(), uv.y(),
                            lib_uv[0], lib_uv[1]);
        }
        above_1024 += uv.x() >= 1024.0F;

        // camera centre (P3 / P4 inputs)
        const Eigen::Vector3f C = ref_camera_center(pose);
        float lib_c[3];
        slamgpu_host_camera_center(pose.data(), lib_c);
        if (bits(C.x()) != bits(lib_c[0]) || bits(C.y()) != bits(lib_c[1]) || bits(C.z()) != bits(lib_c[2])) {
            if (bad_center++ == 0)
                std::printf("  first camera_center mismatch at i=%d: Eigen (%.9g, %.9g, %.9g) library (%.9g, %.9g, %.9g)\n", i,
                            C.x(), C.y(), C.z(), lib_c[0], lib_c[1], lib_c[2]);
        }

        // avg_viewing_normal / observed_distance_range over 1..8 observers
        const int n_obs = 1 + static_cast<int>(rng() % 8);
        std::vector<Eigen::Vector3f> centers;
        std::vector<float> flat;
        for (int o = 0; o < n_obs; o++) {
            centers.emplace_back(uni(-2, 2), uni(-2, 2), uni(-2, 2));
            flat.insert(flat.end(), {centers.back().x(), centers.back().y(), centers.back().z()});
        }
        const Eigen::Vector3f normal = ref_avg_viewing_normal(X, centers);
        const std::pair<float, float> range = ref_observed_distance_range(X, centers);
        float lib_n[3], lib_r[2];
        slamgpu_host_point_stats(X.data(), n_obs, flat.data(), lib_n, lib_r);
        if (bits(normal.x()) != bits(lib_n[0]) || bits(normal.y()) != bits(lib_n[1]) || bits(normal.z()) != bits(lib_n[2])) {
            if (bad_normal++ == 0)
                std::printf("  first avg_viewing_normal mismatch at i=%d (%d observers): Eigen (%.9g, %.9g, %.9g) library (%.9g, %.9g, %.9g)\n",
                            i, n_obs, normal.x(), normal.y(), normal.z(), lib_n[0], lib_n[1], lib_n[2]);
        }
        if (bits(range.first) != bits(lib_r[0]) || bits(range.second) != bits(lib_r[1])) {
            if (bad_range++ == 0)
                std::printf("  first observed_distance_range mismatch at i=%d: Eigen (%.9g, %.9g) library (%.9g, %.9g)\n", i,
                            range.first, range.second, lib_r[0], lib_r[1]);
        }

        // the gate expressions of match_via_reproject (src/MapMatcher.cpp:62-70): dot(normal, ray.normalized()), ray.norm().
        // The library's kernels evaluate them as point_stats does for ONE observer at the camera centre: the
        // normal of a single observation is ray.normalized().normalized() and the range is (norm, norm).
        const Eigen::Vector3f ray = X - C;
        const float cosine = normal.dot(ray.normalized());
        const float distance = ray.norm();
        float one_n[3], one_r[2];
        slamgpu_host_point_stats(X.data(), 1, lib_c, one_n, one_r);
        const Eigen::Vector3f rn = ray.normalized();
        // dot product order: a0 + (a1 + a2)?  Compare against both the library's assumption and report
        const float p0 = normal.x() * rn.x(), p1 = normal.y() * rn.y(), p2 = normal.z() * rn.z();
        volatile float t = p1 + p2;
        volatile float lib_cos = p0 + t;
        if (bits(distance) != bits(one_r[0]) || bits(cosine) != bits(lib_cos)) {
            if (bad_gate++ == 0)
                std::printf("  first gate mismatch at i=%d: norm Eigen %.9g library %.9g; dot Eigen %.9g library %.9g\n", i, distance,
                            one_r[0], cosine, static_cast<float>(lib_cos));
        }
    }
    std::printf("Camera::project            %ld mismatches of %d (%ld projections at u >= 1024 px)\n", bad_project, n, above_1024);
    std::printf("Frame::camera_center       %ld mismatches of %d\n", bad_center, n);
    std::printf("avg_viewing_normal         %ld mismatches of %d\n", bad_normal, n);
    std::printf("observed_distance_range    %ld mismatches of %d\n", bad_range, n);
    std::printf("gate dot / norm            %ld mismatches of %d\n", bad_gate, n);
    const long bad = bad_project + bad_center + bad_normal + bad_range + bad_gate;
    std::printf("%s\n", bad == 0 ? "OK: the operation order libslamgpu assumes is the one this Eigen build evaluates"
                                 : "MISMATCH: see above; the projection path is not bit-exact on this toolchain");
    return bad == 0 ? 0 : 1;
}
