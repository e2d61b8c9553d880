// Drop-in for src/WindowSearch/LinearSearch.h + LinearSearch.cpp (LinearSearch.h:9-20).
#ifndef STEREO_RECONSTRUCTION_LINEARSEARCH_H
#define STEREO_RECONSTRUCTION_LINEARSEARCH_H
#include <iostream>
#include <opencv2/core.hpp>
#include "../sr_shim_common.h"

class LinearSearch {
public:
    LinearSearch(cv::Mat& leftImage, cv::Mat& rightImage) : leftImage_(leftImage), rightImage_(rightImage) {}
    cv::Mat computeDisparityMap(double smoothFactor) {
        std::cout << "h1:" << leftImage_.rows << "\n";                                           // LinearSearch.cpp:16-17
        std::cout << "h2:" << rightImage_.rows << "\n";
        sr_search_params p{SR_VARIANT_LINEAR, 1, 0, 200, smoothFactor, 0, 19.0, SR_COST_SSD_L2, 0, 0};
        cv::Mat out(rightImage_.rows, rightImage_.cols, CV_64F);
        sr_image l = sr_shim::image(leftImage_), r = sr_shim::image(rightImage_);
        sr_shim::check(sr_block_search(sr_shim::ctx(), &l, &r, &p, out.ptr<double>(), out.step, nullptr));
        return out;
    }
private:
    cv::Mat leftImage_;
    cv::Mat rightImage_;
};
#endif
