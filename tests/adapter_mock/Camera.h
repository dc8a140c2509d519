#pragma once
#include <Eigen/Dense>
#include "mock_math.h"
namespace slam {
class Camera {   // accessors of reference src/Camera.h:13-16 and project / is_in_image (src/Camera.cpp:25-37)
  public:
    Camera(float fx, float fy, float cx, float cy, int width, int height) : m_width(width), m_height(height)
    {
        m_K(0, 0) = fx; m_K(1, 1) = fy; m_K(0, 2) = cx; m_K(1, 2) = cy; m_K(2, 2) = 1.0F;
    }
    const Eigen::Matrix3f& get_intrinsic_matrix() const { return m_K; }
    int get_width() const { return m_width; }
    int get_height() const { return m_height; }
    Eigen::Vector2f project(const Eigen::Matrix4f& pose, const Eigen::Vector3f& X) const
    {   // uv = K * pose[0:3, 0:4] * [X; 1]: M = K * P34 first, then M[:, :3] * X + M[:, 3]
        float h[3];
        for (int i = 0; i < 3; i++) {
            float M[4];
            for (int j = 0; j < 4; j++) {
                volatile float p0 = m_K(i, 0) * pose(0, j), p1 = m_K(i, 1) * pose(1, j), p2 = m_K(i, 2) * pose(2, j);
                M[j] = mockmath::sum3(p0, p1, p2);
            }
            volatile float q0 = M[0] * X.x(), q1 = M[1] * X.y(), q2 = M[2] * X.z();
            volatile float s = mockmath::sum3(q0, q1, q2);
            h[i] = s + M[3];
        }
        if (h[2] < 0.0F) {
            return Eigen::Vector2f(-1.0F, -1.0F);
        }
        return Eigen::Vector2f(h[0] / h[2], h[1] / h[2]);
    }
  private:
    Eigen::Matrix3f m_K;
    int m_width, m_height;
};
} // namespace slam
