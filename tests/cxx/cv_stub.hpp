// Test stand-ins for the few OpenCV / Boost / reference types include/ugdet_face_detector.hpp touches, so that the
// shim can be compiled and run in an image without OpenCV headers.  Layouts follow the real ones where the shim
// relies on them: cv::Rect = {x, y, width, height} ints; cv::Mat exposes data / cols / rows / step / empty();
// cv::Scalar indexes doubles; Configuration::DetectionSection has the reference's four fields
// (/root/reference/src/structure.h:43-52).
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>

namespace cv {
struct Rect {
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int x_, int y_, int w_, int h_) : x(x_), y(y_), width(w_), height(h_) {}
};
struct Scalar {
    double val[4];
    Scalar(double a = 0, double b = 0, double c = 0, double d = 0) : val{a, b, c, d} {}
    double operator[](int i) const { return val[i]; }
};
struct Mat {
    unsigned char* data = nullptr;
    int cols = 0, rows = 0;
    size_t step = 0;
    bool empty() const { return !data || cols == 0 || rows == 0; }
};
class CascadeClassifier {};
}  // namespace cv

namespace boost { namespace filesystem {
class path {
  public:
    path() {}
    path(const std::string& s) : s_(s) {}
    const std::string& native() const { return s_; }
  private:
    std::string s_;
};
}}  // namespace boost::filesystem

namespace ugproj {
struct Configuration {
    struct DetectionSection {
        cv::Scalar skin_lower;
        cv::Scalar skin_upper;
        boost::filesystem::path cascade_filepath;
        double scale;
    };
    DetectionSection detection;
};
}  // namespace ugproj
