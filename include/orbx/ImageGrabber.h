/*
 * orbx/ImageGrabber.h -- C++ face of the pre-processing entries of the C ABI (orbx.h: orbx_preproc_*), named after the
 * image half of the reference's ros2_orbslam3::ImageGrabber
 * (/root/reference ros2_ws/src/mono-inertial/include/image_grabber.hpp:38-49 constructor, :96-110 ConvertImageToGPU):
 * it holds the two undistortion maps and the tracking image size and turns a BGR camera frame into the u8 gray image
 * System::TrackMonocular takes -- remap(INTER_CUBIC, BORDER_CONSTANT) -> resize(INTER_LINEAR) -> BGR2GRAY.
 * The ROS subscription / IMU synchronisation half of that class is host logic and stays in the reference.
 * Header-only; throws std::runtime_error on failure (no CPU fallback).
 */
#ifndef ORBX_IMAGEGRABBER_H
#define ORBX_IMAGEGRABBER_H
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../orbx.h"
#include "ORBextractor.h"

namespace ros2_orbslam3 {

class ImageGrabber {
public:
    /* undistortion_map_1 / _2: the two CV_32FC1 planes of initUndistortRectifyMap (map_w x map_h, dense);
     * width x height: the size ConvertImageToGPU resizes to (cam.new_width x cam.new_height) */
    ImageGrabber(int src_width, int src_height, const float* undistortion_map_1, const float* undistortion_map_2, int map_width,
                 int map_height, int width, int height, int device = 0)
        : m_width(width), m_height(height) {
        check(orbx_preproc_create(src_width, src_height, undistortion_map_1, undistortion_map_2, map_width, map_height, width, height, 1,
                                  device, &h_));
        grey_.resize((size_t)width * height);
    }
#ifdef ORBX_WITH_OPENCV
    ImageGrabber(cv::Size src, const cv::Mat& undistortion_map_1, const cv::Mat& undistortion_map_2, int width, int height, int device = 0)
        : ImageGrabber(src.width, src.height, checked(undistortion_map_1).ptr<float>(), checked(undistortion_map_2).ptr<float>(),
                       undistortion_map_1.cols, undistortion_map_1.rows, width, height, device) {}
    /* image_grabber.hpp:96-110; the returned Mat views this object's buffer, as the reference returns its member */
    cv::Mat ConvertImageToGPU(const cv::Mat& bgr) {
        CV_Assert(bgr.type() == CV_8UC3);
        ConvertImageToGPU(bgr.data, bgr.step);
        return cv::Mat(m_height, m_width, CV_8UC1, grey_.data(), (size_t)m_width);
    }
#endif
    ~ImageGrabber() { if (h_) orbx_preproc_destroy(h_); }
    ImageGrabber(const ImageGrabber&) = delete;
    ImageGrabber& operator=(const ImageGrabber&) = delete;

    /* one BGR frame (rows `step` bytes apart) -> m_height x m_width gray, dense */
    const std::vector<uint8_t>& ConvertImageToGPU(const uint8_t* bgr, size_t step) {
        check(orbx_preproc_host(h_, bgr, 0, (int)step, 1, grey_.data()));
        return grey_;
    }
    /* ConvertImageToGPU -> System::TrackMonocular -> ORBextractor::operator() with the gray image staying on the
     * device; returns what operator() returns */
    int ConvertAndExtract(const uint8_t* bgr, size_t step, ORB_SLAM3::ORBextractor& extractor, std::vector<orbx_keypoint>& keypoints,
                          std::vector<uint8_t>& descriptors, const std::vector<int>& vLappingArea) {
        const int lap[2] = {vLappingArea.size() > 0 ? vLappingArea[0] : 0, vLappingArea.size() > 1 ? vLappingArea[1] : 0};
        check(orbx_preproc_extract_host(h_, extractor.prepare(m_width, m_height), bgr, 0, (int)step, 1, lap));
        return extractor.fetch_last(keypoints, descriptors);
    }
    orbx_preproc* handle() { return h_; }
    const int m_width, m_height;

private:
    static void check(orbx_status st) { if (st != ORBX_OK) throw std::runtime_error(std::string("orbx: ") + orbx_last_error()); }
#ifdef ORBX_WITH_OPENCV
    static const cv::Mat& checked(const cv::Mat& m) { CV_Assert(m.type() == CV_32FC1 && m.isContinuous()); return m; }
#endif
    orbx_preproc* h_ = nullptr;
    std::vector<uint8_t> grey_;
};

}  // namespace ros2_orbslam3
#endif
