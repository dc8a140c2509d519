// Minimal OpenCV stand-in for the adapter test: cv::Mat as an N x 32 byte matrix, cv::KeyPoint, NormTypes.
#pragma once
#include <cstdint>
#include <memory>
#include <vector>
namespace cv {
enum NormTypes { NORM_L2 = 4, NORM_HAMMING = 6 };
struct Point2f { float x = 0, y = 0; Point2f() = default; Point2f(float a, float b) : x(a), y(b) {} };
struct Point3f { float x = 0, y = 0, z = 0; Point3f() = default; Point3f(float a, float b, float c) : x(a), y(b), z(c) {} };
struct KeyPoint { Point2f pt; };
class Mat {
  public:
    int rows = 0, cols = 0;
    Mat() = default;
    Mat(int r, int c) : rows(r), cols(c), m_data(std::make_shared<std::vector<uint8_t>>(size_t(r) * c)), m_off(0) {}
    bool empty() const { return rows == 0; }
    template <class T> const T* ptr(int r) const { return reinterpret_cast<const T*>(m_data->data() + m_off + size_t(r) * cols); }
    template <class T> T* ptr(int r) { return reinterpret_cast<T*>(m_data->data() + m_off + size_t(r) * cols); }
    Mat row(int r) const
    {
        Mat m;
        m.rows = 1; m.cols = cols; m.m_data = m_data; m.m_off = m_off + size_t(r) * cols;
        return m;
    }
  private:
    std::shared_ptr<std::vector<uint8_t>> m_data;
    size_t m_off = 0;
};
} // namespace cv
