#pragma once
#include <Eigen/Dense>
namespace slam {
class Camera {   // accessors of reference src/Camera.h:13-16
  public:
    Camera(float fx, float fy, float cx, float cy, int width, int height) : m_width(width), m_height(height)
    {
        m_K(0, 0) = fx; m_K(1, 1) = fy; m_K(0, 2) = cx; m_K(1, 2) = cy; m_K(2, 2) = 1.0F;
    }
    const Eigen::Matrix3f& get_intrinsic_matrix() const { return m_K; }
    int get_width() const { return m_width; }
    int get_height() const { return m_height; }
  private:
    Eigen::Matrix3f m_K;
    int m_width, m_height;
};
} // namespace slam
