// Minimal stand-in for <opencv2/core.hpp> (OpenCV C++ is not in this image): just the types and calls that
// include/orbx/ORBextractor.h and include/orbx/ImageGrabber.h use under ORBX_WITH_OPENCV, with OpenCV's names, layouts (cv::KeyPoint = 28 bytes)
// and semantics (Mat headers share memory unless created; ROI = same buffer, same step).  Test infrastructure only:
// it exists so that the cv::Mat / cv::KeyPoint overload is compiled and run by tests/test_gpu_cpp_api.py.
#ifndef ORBX_OPENCV_STUB_CORE_HPP
#define ORBX_OPENCV_STUB_CORE_HPP
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <memory>
#include <stdexcept>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_32FC1 5
#define CV_8UC3 16
#define CV_Assert(expr) do { if (!(expr)) throw std::runtime_error("CV_Assert failed: " #expr); } while (0)

namespace cv {
struct Point2f { float x = 0, y = 0; };
struct Size { int width = 0, height = 0; Size() {} Size(int w, int h) : width(w), height(h) {} };
struct Rect { int x, y, width, height; Rect(int x_, int y_, int w_, int h_) : x(x_), y(y_), width(w_), height(h_) {} };
struct Range { int start, end; };
class KeyPoint {
public:
    Point2f pt;
    float size = 0, angle = -1, response = 0;
    int octave = 0, class_id = -1;
};
class Mat {
public:
    int rows = 0, cols = 0;
    uint8_t* data = nullptr;
    size_t step = 0;
    Mat() {}
    Mat(int r, int c, int type, void* d, size_t s = 0) : rows(r), cols(c), data((uint8_t*)d), step(s ? s : (size_t)c * esz(type)), type_(type) {}
    Mat(int r, int c, int type) { create(r, c, type); }
    void create(int r, int c, int type) {
        if (r == rows && c == cols && type == type_ && own_) return;
        own_ = std::shared_ptr<uint8_t>((uint8_t*)malloc((size_t)r * c * esz(type) + 1), free);
        rows = r; cols = c; data = own_.get(); step = (size_t)c * esz(type); type_ = type;
    }
    static size_t esz(int type) { return type == CV_8UC3 ? 3 : (type == CV_32FC1 ? 4 : 1); }
    size_t elemSize() const { return esz(type_); }
    void release() { own_.reset(); rows = cols = 0; data = nullptr; step = 0; }
    bool empty() const { return !data || rows == 0 || cols == 0; }
    int type() const { return type_; }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }
    Mat operator()(const Rect& r) const { Mat m = *this; m.data = data + (size_t)r.y * step + r.x * elemSize(); m.rows = r.height; m.cols = r.width; return m; }
    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
    Mat clone() const { Mat m(rows, cols, CV_8UC1); for (int y = 0; y < rows; ++y) for (int x = 0; x < cols; ++x) m.data[y * m.step + x] = data[y * step + x]; return m; }
    template <class T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + x * sizeof(T)); }
    template <class T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + x * sizeof(T)); }
    uint8_t* ptr(int y = 0) { return data + (size_t)y * step; }
    const uint8_t* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <class T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
    template <class T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
private:
    std::shared_ptr<uint8_t> own_;
    int type_ = CV_8UC1;
};
class _InputArray {
public:
    _InputArray() {}
    _InputArray(const Mat& m) : m_(&m) {}
    Mat getMat() const { return m_ ? *m_ : Mat(); }
    bool empty() const { return !m_ || m_->empty(); }
private:
    const Mat* m_ = nullptr;
};
class _OutputArray {
public:
    _OutputArray(Mat& m) : m_(&m) {}
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    void release() const { m_->release(); }
    Mat getMat() const { return *m_; }
private:
    Mat* m_;
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
inline InputArray noArray() { static _InputArray none; return none; }
}  // namespace cv
#endif
