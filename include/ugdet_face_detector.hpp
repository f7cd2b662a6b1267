// ugdet_face_detector.hpp -- drop-in replacement for the reference's src/tracker/detector.h + detector.cc.
//
// Same namespace, class name, constructor and method signatures as /root/reference/src/tracker/detector.h:10-19:
//
//     FaceDetector(cv::CascadeClassifier& cascade, const Configuration::DetectionSection& cfg);
//     void detectFaces(const cv::Mat& frame, std::vector<cv::Rect>& rects);
//
// so the two call sites -- `FaceDetector detector(cascade, this->cfg_->detection);` (face_tracker.cc:43) and
// `detector->detectFaces(curr_frame, rects);` (face_tracker.cc:295) -- compile unchanged.  A maintainer replaces the
// body of src/tracker/detector.h with `#include "ugdet_face_detector.hpp"`, deletes detector.cc from
// src/tracker/CMakeLists.txt and links libugdet.so.
//
// What it does differently inside: the cv::CascadeClassifier& is ignored (OpenCV offers no way to read a loaded
// cascade back); the model is loaded from cfg.cascade_filepath, the same file TrackerFileInput::open loaded it from
// (file_io.cc:11).  Everything detector.cc:10-43 computes happens in ugd_detect_bgr on the GPU.  Errors surface as
// std::runtime_error -- the reference's detectFaces can only fail by a cv::Exception, which no caller handles either.
//
// Header-only; needs <opencv2/core.hpp>-style cv::Mat / cv::Rect / cv::Scalar and the reference's structure.h
// (both already included by detector.h), plus include/ugdet.h.
#ifndef UGDET_FACE_DETECTOR_HPP
#define UGDET_FACE_DETECTOR_HPP

#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

#include "ugdet.h"

#ifndef UGDET_SHIM_NO_REFERENCE_INCLUDES      // tests/cxx supplies stand-ins for these two
#include <opencv2/opencv.hpp>
#include "../structure.h"
#endif

namespace ugproj {

class FaceDetector {
  public:
    // detector.h:12-14.  `cascade` is accepted for source compatibility and not used (see above).
    FaceDetector(cv::CascadeClassifier& cascade, const Configuration::DetectionSection& cfg)
        : cascade(cascade), cfg_(cfg) {
        const std::string xml = cfg_.cascade_filepath.native();
        if (int rc = ugd_cascade_load_xml(xml.c_str(), &model_))
            throw std::runtime_error("ugdet: cannot load cascade " + xml + ": " + ugd_strerror(rc));
    }
    ~FaceDetector() {
        if (det_) ugd_destroy(det_);
        if (model_) ugd_cascade_free(model_);
    }
    FaceDetector(const FaceDetector&) = delete;
    FaceDetector& operator=(const FaceDetector&) = delete;

    // detector.h:15 / detector.cc:10-43.  `frame`: BGR 8UC3, any row step; `rects` is APPENDED to (detector.cc:41).
    void detectFaces(const cv::Mat& frame, std::vector<cv::Rect>& rects) {
        if (frame.empty()) return;
        ensure(frame.cols, frame.rows);
        int n = 0;
        for (;;) {                                   // a list longer than the buffer is fetched again, never cut
            const int cap = static_cast<int>(buf_.size());
            const int rc = ugd_detect_bgr(det_, frame.data, frame.cols, frame.rows, static_cast<size_t>(frame.step),
                                          buf_.data(), cap, &n);
            if (rc == UGD_ERR_TRUNCATED && n > cap) { buf_.resize(static_cast<size_t>(n)); continue; }
            if (rc != UGD_OK) throw std::runtime_error(std::string("ugdet: detectFaces: ") + ugd_last_error(det_));
            break;
        }
        for (int i = 0; i < n; ++i) rects.push_back(cv::Rect(buf_[i].x, buf_[i].y, buf_[i].width, buf_[i].height));
    }

  private:
    // The detector is sized for the first frame it sees (the tracker scans every frame at one size,
    // face_tracker.cc:112) and rebuilt if a larger one ever arrives.
    void ensure(int w, int h) {
        if (det_ && w <= max_w_ && h <= max_h_) return;
        if (det_) { ugd_destroy(det_); det_ = nullptr; }
        ugd_params p;
        ugd_params_default(&p);                      // minNeighbors 2, minSize 30x30: the literals of detector.cc:29-31
        p.scale_factor = cfg_.scale;                 // structure.cc:83-85
        for (int i = 0; i < 3; ++i) {                // Y, Cr, Cb bounds: structure.cc:75-82
            p.skin_lower[i] = clamp8(cfg_.skin_lower[i]);
            p.skin_upper[i] = clamp8(cfg_.skin_upper[i]);
        }
        max_w_ = w > max_w_ ? w : max_w_;
        max_h_ = h > max_h_ ? h : max_h_;
        int dev = 0;
        if (const char* e = std::getenv("UGDET_DEVICE")) dev = std::atoi(e);
        if (int rc = ugd_create(model_, &p, dev, max_w_, max_h_, /*max_batch*/ 1, &det_))
            throw std::runtime_error(std::string("ugdet: ugd_create: ") + ugd_strerror(rc));
        if (buf_.empty()) buf_.resize(256);
    }
    static uint8_t clamp8(double v) { return static_cast<uint8_t>(v < 0 ? 0 : v > 255 ? 255 : v + 0.5); }

    cv::CascadeClassifier& cascade;                  // kept so the class layout mirrors detector.h:17-18
    const Configuration::DetectionSection& cfg_;
    ugd_cascade* model_ = nullptr;
    ugd_detector* det_ = nullptr;
    int max_w_ = 0, max_h_ = 0;
    std::vector<ugd_rect> buf_;
};

}  // namespace ugproj

#endif  // UGDET_FACE_DETECTOR_HPP
