// Minimal stand-in for <opencv2/core/core.hpp>: just enough of cv::Mat / cv::FileStorage for the
// reference's vendored DBoW2 (thirdparty/DBoW2) to compile UNMODIFIED into oracle/_ref/libref_dbow2.so.
// Test infrastructure only (OpenCV's C++ headers are absent from this image).  cv::Mat here owns its
// bytes and copies deeply; DBoW2 never relies on cv::Mat's shared-buffer aliasing.  FileStorage is a
// stub that is never opened: vocabularies go in and out through the wrapper's plain arrays.
#pragma once

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <cstring>
#include <string>
#include <vector>

#define CV_8U 0
#define CV_32F 5

namespace cv {

class Mat {
  public:
    int rows = 0, cols = 0;
    Mat() = default;
    Mat(int r, int c, int type) { create(r, c, type); }
    static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }
    void create(int r, int c, int type)
    {
        rows = r;
        cols = c;
        m_elem = type == CV_32F ? 4 : 1;
        m_data.assign(static_cast<size_t>(r) * c * m_elem, 0);
    }
    void release()
    {
        rows = cols = 0;
        m_data.clear();
    }
    bool empty() const { return m_data.empty(); }
    Mat clone() const { return *this; }
    Mat row(int i) const
    {
        Mat m;
        m.rows = 1;
        m.cols = cols;
        m.m_elem = m_elem;
        const size_t stride = static_cast<size_t>(cols) * m_elem;
        m.m_data.assign(m_data.begin() + i * stride, m_data.begin() + (i + 1) * stride);
        return m;
    }
    template <class T> T* ptr(int r = 0) { return reinterpret_cast<T*>(m_data.data() + static_cast<size_t>(r) * cols * m_elem); }
    // An EMPTY Mat reads as 64 zero bytes here.  DBoW2's k-means can leave a node without a descriptor when a
    // cluster ends up empty (FORB::meanValue releases the mean); with the real cv::Mat the next
    // FORB::distance on it dereferences NULL.  The stand-in keeps such vocabularies usable (the wrapper
    // exports the same zero descriptor), which only matters for the randomly seeded live test.
    template <class T> const T* ptr(int r = 0) const
    {
        static const uint8_t zeros[64] = {0};
        if (m_data.empty()) return reinterpret_cast<const T*>(zeros);
        return reinterpret_cast<const T*>(m_data.data() + static_cast<size_t>(r) * cols * m_elem);
    }

  private:
    int m_elem = 1;
    std::vector<uint8_t> m_data;
};

class FileNode {
  public:
    enum { SEQ = 5 };
    FileNode operator[](const std::string&) const { return FileNode(); }
    FileNode operator[](const char*) const { return FileNode(); }
    FileNode operator[](int) const { return FileNode(); }
    size_t size() const { return 0; }
    int type() const { return SEQ; }
    bool empty() const { return true; }
    operator int() const { return 0; }
    operator float() const { return 0.0f; }
    operator double() const { return 0.0; }
    operator std::string() const { return std::string(); }
};

class FileStorage {
  public:
    enum { READ = 0, WRITE = 1 };
    FileStorage() = default;
    FileStorage(const std::string&, int) {}
    bool isOpened() const { return false; }
    FileNode operator[](const std::string&) const { return FileNode(); }
    FileNode operator[](const char*) const { return FileNode(); }
    void release() {}
};

template <class T> inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }

} // namespace cv

// the real header brings these into scope for DBoW2's unqualified calls
using std::log;
using std::pow;
using std::sqrt;
