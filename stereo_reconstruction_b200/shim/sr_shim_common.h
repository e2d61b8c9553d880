// sr_shim_common.h -- shared plumbing of the header-compatible shims: one process-wide sr_ctx per
// host thread, conversion of non-zero C-ABI status into C++ exceptions (the reference's error
// convention is exceptions: cv::Exception for bad ROIs, std::runtime_error elsewhere).
#pragma once
#include <stdexcept>
#include <string>
#include <opencv2/core.hpp>
#include "sr_capi.h"   // copied next to this file by INTEGRATION.md 2 (in this repository: -I include)

namespace sr_shim {
inline sr_ctx* ctx() {
    thread_local sr_ctx* c = nullptr;
    if (!c && sr_ctx_create(0, &c) != SR_OK) throw std::runtime_error(std::string("libsr_b200: ") + sr_last_error(nullptr));
    return c;
}
inline void check(int rc) {
    if (rc == SR_OK) return;
    std::string msg = std::string("libsr_b200: ") + sr_last_error(ctx());
    if (rc == SR_E_SHAPE) CV_Error(cv::Error::StsBadSize, msg);   // where the reference throws cv::Exception
    throw std::runtime_error(msg);
}
inline sr_image image(const cv::Mat& m) {
    CV_Assert(m.type() == CV_8UC3);
    return sr_image{m.data, m.cols, m.rows, m.step};
}
}  // namespace sr_shim
