// Drop-in for src/WindowSearch/BlockSearch.h + BlockSearch.cpp of the reference (same class, same
// signatures, same defaults, BlockSearch.h:9-46); the work is done by libsr_b200.so on the GPU.
#ifndef STEREO_RECONSTRUCTION_BLOCKSEARCH_H
#define STEREO_RECONSTRUCTION_BLOCKSEARCH_H
#include <iostream>
#include <vector>
#include <opencv2/core.hpp>
#include "../sr_shim_common.h"

class BlockSearch {
public:
    BlockSearch(cv::Mat& leftImage, cv::Mat& rightImage, int blockSize, int minDisparity, int maxDisparity)
        : leftImage_(leftImage), rightImage_(rightImage), blockSize_(blockSize), maxDisparity_(maxDisparity),
          minDisparity_(minDisparity) { blockSizes_.push_back(blockSize); }                       // BlockSearch.cpp:6-13
    BlockSearch(cv::Mat& leftImage, cv::Mat& rightImage, std::vector<int>& blockSizes, int minDisparity, int maxDisparity)
        : leftImage_(leftImage), rightImage_(rightImage), blockSize_(blockSizes.at(0)), maxDisparity_(maxDisparity),
          minDisparity_(minDisparity), blockSizes_(blockSizes) {}                                 // :15-22 (blockSize_ = blockSizes[0])

    cv::Mat computeDisparityMapLeft(double smoothFactor) { return run(SR_VARIANT_LEFT, smoothFactor, false, 19.0, leftImage_); }
    cv::Mat computeDisparityMapRight(double smoothFactor, bool varBlock = false, double thres = 19.0) {
        return run(SR_VARIANT_RIGHT, smoothFactor, varBlock, thres, rightImage_);
    }

private:
    cv::Mat run(int variant, double s, bool varBlock, double thres, const cv::Mat& ref) {
        sr_search_params p{variant, blockSize_, minDisparity_, maxDisparity_, s, varBlock ? 1 : 0, thres, SR_COST_SSD_L2, 0, 0};
        cv::Mat out(ref.rows, ref.cols, CV_64F);                                                 // BlockSearch.cpp:33,97
        sr_image l = sr_shim::image(leftImage_), r = sr_shim::image(rightImage_);
        int maxBlock = blockSize_;
        sr_shim::check(sr_block_search(sr_shim::ctx(), &l, &r, &p, out.ptr<double>(), out.step, &maxBlock));
        if (variant == SR_VARIANT_RIGHT) std::cout << "max block size: " << maxBlock << "\n";   // :177
        return out;
    }
    cv::Mat leftImage_;   // shallow copies, never written (BlockSearch.h:40-41)
    cv::Mat rightImage_;
    int blockSize_;
    int maxDisparity_;
    int minDisparity_;
    std::vector<int> blockSizes_;
};
#endif
