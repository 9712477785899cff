// TEST INFRASTRUCTURE ONLY. Minimal stand-in for the few OpenCV core types include/cvUtils/Conversion.hh uses
// (cv::Mat single-channel, cv::KeyPoint, cv::DMatch), because this image has no OpenCV C++ development package.
// Layout and constants follow OpenCV's public API (CV_8UC1 = 0, CV_32FC1 = 5, CV_64FC1 = 6).
#pragma once
#include <cstddef>
#include <cstdint>
#include <memory>
#include <vector>

#define CV_8UC1 0
#define CV_32FC1 5
#define CV_64FC1 6
typedef unsigned char uchar;

namespace cv {
struct Point2f { float x = 0, y = 0; };
struct KeyPoint {
    Point2f pt; float size = 0, angle = -1, response = 0; int octave = 0, class_id = -1;
};
struct DMatch {
    int queryIdx = -1, trainIdx = -1, imgIdx = -1; float distance = 0;
    DMatch() {}
    DMatch(int q, int t, float d) : queryIdx(q), trainIdx(t), distance(d) {}
};
class Mat {
  public:
    int rows = 0, cols = 0;
    uchar* data = nullptr;
    Mat() {}
    Mat(int r, int c, int type) : rows(r), cols(c), m_type(type), m_store(std::make_shared<std::vector<uchar>>(size_t(r) * c * esz(type))) {
        data = m_store->data();
    }
    int type() const { return m_type; }
    int channels() const { return 1; }
    bool empty() const { return rows == 0 || cols == 0; }
    size_t elemSize() const { return esz(m_type); }
    uchar* ptr(int r = 0) { return data + size_t(r) * cols * elemSize(); }
    const uchar* ptr(int r = 0) const { return data + size_t(r) * cols * elemSize(); }
    template <typename T> T* ptr(int r = 0) { return reinterpret_cast<T*>(ptr(r)); }
    template <typename T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(ptr(r)); }
    template <typename T> T& at(int r, int c) { return ptr<T>(r)[c]; }
    template <typename T> const T* begin() const { return reinterpret_cast<const T*>(data); }
    template <typename T> const T* end() const { return begin<T>() + size_t(rows) * cols; }
  private:
    static size_t esz(int t) { return t == CV_8UC1 ? 1 : t == CV_32FC1 ? 4 : 8; }
    int m_type = 0;
    std::shared_ptr<std::vector<uchar>> m_store;
};
}  // namespace cv
