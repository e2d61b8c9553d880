// Drop-in for src/Reconstruction/reconstruction.h + reconstruction.cpp (reconstruction.h:13-34): same
// struct Vertex, same free functions, same defaults; arithmetic on the GPU through libsr_b200.so.
#ifndef STEREO_RECONSTRUCTION_RECONSTRUCTION_H
#define STEREO_RECONSTRUCTION_RECONSTRUCTION_H
#include <iostream>
#include <string>
#include <vector>
#include <opencv2/core.hpp>
#include <cmath>
#include "../Eigen.h"             // the reference's own src/Eigen.h: Vector4f, Vector4uc, Matrix3f, MINF
#include "../sr_shim_common.h"

struct Vertex {
    EIGEN_MAKE_ALIGNED_OPERATOR_NEW
    Vector4f position;   // 16 B
    Vector4uc color;     // 4 B (+12 B padding): sizeof(Vertex) == sizeof(sr_vertex) == 32
};
static_assert(sizeof(Vertex) == sizeof(sr_vertex), "Vertex layout differs from sr_vertex");

inline bool CheckTriangularValidity(Vertex* vertices, unsigned int one, unsigned int two, unsigned int three, float threshold) {
    // reconstruction.cpp:46-69 -- nine subtractions and three square roots: computed here, a GPU round trip per triangle
    // would cost more than the whole mesh (WriteMesh / sr_mesh_faces evaluate all cells on the device in one pass)
    const Vector4f &a = vertices[one].position, &b = vertices[two].position, &c = vertices[three].position;
    if (a.x() == MINF || b.x() == MINF || c.x() == MINF) return false;
    auto len = [](const Vector4f& p, const Vector4f& q) {
        return sqrtf(powf(p.x() - q.x(), 2) + powf(p.y() - q.y(), 2) + powf(p.z() - q.z(), 2));
    };
    return !(len(a, b) > threshold || len(a, c) > threshold || len(b, c) > threshold);
}

inline bool WriteMesh(Vertex* vertices, unsigned int width, unsigned int height, const std::string& filename, float edgeThreshold) {
    int rc = sr_write_mesh(sr_shim::ctx(), reinterpret_cast<sr_vertex*>(vertices), (int)width, (int)height, filename.c_str(), edgeThreshold);
    if (rc == SR_E_IO) return false;                                                             // reconstruction.cpp:117
    sr_shim::check(rc);
    return true;
}

inline void removeDisparityOutliers(cv::Mat& disparityMap, int kernelSize = 500, float thrFront = 1.5, float thrBack = 0.8) {
    CV_Assert(disparityMap.type() == CV_32FC1);
    sr_shim::check(sr_remove_outliers(sr_shim::ctx(), disparityMap.ptr<float>(), disparityMap.cols, disparityMap.rows,
                                      disparityMap.step, kernelSize, thrFront, thrBack));
}

inline void scaleDisparityMap(cv::Mat& disparityMap, float scalingFactor) {
    CV_Assert(disparityMap.type() == CV_32FC1);
    sr_shim::check(sr_scale(sr_shim::ctx(), disparityMap.ptr<float>(), disparityMap.cols, disparityMap.rows, disparityMap.step, scalingFactor));
}

inline cv::Mat convertDisparityToDepth(const cv::Mat& dispImage, float focalLength, float baseline = 1.0f) {
    cv::Mat d = dispImage.isContinuous() ? dispImage : dispImage.clone();
    CV_Assert(d.type() == CV_32FC1);
    cv::Mat depth(d.rows, d.cols, CV_32FC1);
    sr_shim::check(sr_disparity_to_depth(sr_shim::ctx(), d.ptr<float>(), depth.ptr<float>(), d.cols, d.rows, focalLength, baseline));
    return depth;
}

inline void reconstruction(cv::Mat bgrImage, cv::Mat depthValues, Matrix3f intrinsics, float thrMesh = 1.0f) {
    cv::Mat depth = depthValues.isContinuous() ? depthValues : depthValues.clone();
    std::vector<sr_vertex> v((size_t)depth.cols * depth.rows);
    // Eigen Matrix3f is column-major: K(0,0), K(1,1), K(0,2), K(1,2)   (reconstruction.cpp:154-157)
    sr_shim::check(sr_backproject(sr_shim::ctx(), bgrImage.data, bgrImage.step, depth.ptr<float>(), depth.cols, depth.rows,
                                  intrinsics(0, 0), intrinsics(1, 1), intrinsics(0, 2), intrinsics(1, 2), v.data()));
    int rc = sr_write_mesh(sr_shim::ctx(), v.data(), depth.cols, depth.rows, "../../results/reconstruction_mesh.off", thrMesh);
    if (rc != SR_OK) std::cout << "Failed to write mesh!\nCheck file path!" << std::endl;        // :201-204
}
#endif
