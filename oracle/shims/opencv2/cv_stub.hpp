/*
 * oracle/shims/opencv2/cv_stub.hpp — TEST INFRASTRUCTURE.
 *
 * Minimal stand-in for the slice of OpenCV that Source.cpp touches so the file compiles
 * unmodified without OpenCV C++ (not installed in this image).  The blur pipelines only
 * use Mat::data, Mat::size[0..1], Mat::channels() (Source.cpp:149,163,272,...).  The I/O
 * and cv::GaussianBlur / cv::blur calls of Test()/main() (Source.cpp:599-602,623,630,635)
 * are declared but abort if reached: they are the out-of-scope comparison line.
 */
#ifndef SHIM_CV_STUB_HPP
#define SHIM_CV_STUB_HPP
#include <cstdint>
#include <cstdio>
#include <cstdlib>

namespace cv {
struct MatSize {
    int p[2];
    int &operator[](int i) { return p[i]; }
    const int &operator[](int i) const { return p[i]; }
};
struct Size {
    int width, height;
    Size(int w, int h) : width(w), height(h) {}
};
struct Mat {
    uint8_t *data = nullptr;
    MatSize size{{0, 0}};
    int nch = 3;
    int channels() const { return nch; }
    void release() {}
};
inline Mat imread(const char *) { std::fprintf(stderr, "cv stub: imread unavailable\n"); std::abort(); }
inline bool imwrite(const char *, const Mat &) { std::fprintf(stderr, "cv stub: imwrite unavailable\n"); std::abort(); }
inline void resize(const Mat &, Mat &, Size) { std::fprintf(stderr, "cv stub: resize unavailable\n"); std::abort(); }
inline void GaussianBlur(const Mat &, Mat &, Size, double) { std::fprintf(stderr, "cv stub: GaussianBlur unavailable\n"); std::abort(); }
inline void blur(const Mat &, Mat &, Size) { std::fprintf(stderr, "cv stub: blur unavailable\n"); std::abort(); }
}  // namespace cv
#endif
