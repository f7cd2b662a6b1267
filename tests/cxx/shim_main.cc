// A C++ caller of the drop-in shim, written like the reference's own call sites:
//   FaceDetector detector(cascade, cfg->detection);          (face_tracker.cc:43)
//   detector->detectFaces(curr_frame, rects);                (face_tracker.cc:295)
// usage: shim_main <cascade.xml> <frame.bgr> <w> <h> <row step bytes> [scale]
// Reads a raw BGR frame (rows `step` bytes apart), runs detectFaces twice (the second call must APPEND), prints
// "n" then one "x y w h" line per rectangle.  tests/test_shim.py compares the output with the oracle.
#include <cstdio>
#include <cstdlib>
#include <vector>

#define UGDET_SHIM_NO_REFERENCE_INCLUDES
#include "cv_stub.hpp"
#include "ugdet_face_detector.hpp"

int main(int argc, char** argv) {
    if (argc < 6) { std::fprintf(stderr, "usage: %s cascade.xml frame.bgr w h step [scale]\n", argv[0]); return 2; }
    const int w = std::atoi(argv[3]), h = std::atoi(argv[4]);
    const size_t step = (size_t)std::atol(argv[5]);
    std::vector<unsigned char> buf(step * (size_t)h);
    FILE* f = std::fopen(argv[2], "rb");
    if (!f || std::fread(buf.data(), 1, buf.size(), f) != buf.size()) { std::fprintf(stderr, "cannot read %s\n", argv[2]); return 2; }
    std::fclose(f);

    ugproj::Configuration cfg;
    cfg.detection.skin_lower = cv::Scalar(0, 133, 77);            // structure.cc:75-82 defaults
    cfg.detection.skin_upper = cv::Scalar(255, 173, 127);
    cfg.detection.cascade_filepath = boost::filesystem::path(argv[1]);
    cfg.detection.scale = argc > 6 ? std::atof(argv[6]) : 1.1;
    cv::CascadeClassifier cascade;                                 // what TrackerFileInput would hold
    try {
        ugproj::FaceDetector detector(cascade, cfg.detection);
        cv::Mat frame;
        frame.data = buf.data(); frame.cols = w; frame.rows = h; frame.step = step;
        std::vector<cv::Rect> rects;
        rects.push_back(cv::Rect(-1, -1, -1, -1));                 // detectFaces appends: this one must survive
        detector.detectFaces(frame, rects);
        const size_t once = rects.size() - 1;
        detector.detectFaces(frame, rects);
        if (rects.size() != 1 + 2 * once || rects[0].x != -1) { std::fprintf(stderr, "detectFaces did not append\n"); return 1; }
        std::printf("%zu\n", once);
        for (size_t i = 1; i <= once; ++i) std::printf("%d %d %d %d\n", rects[i].x, rects[i].y, rects[i].width, rects[i].height);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    return 0;
}
