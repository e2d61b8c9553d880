// dropin_caller.cpp -- a caller of the hot path shaped like the reference's own call sites, compiled TWICE from this one
// source: (a) against the reference's headers + its BlockSearch.cpp / LinearSearch.cpp / reconstruction.cpp (the
// reference build, oracle/_ref/dropin_ref) and (b) against the swapped tree of INTEGRATION.md 2 -- the shim headers in
// place of the reference's, linked to libsr_b200.so.  Both read the same raw images and write the same files; the test
// compares them byte for byte (tests/test_shim_dropin.py).
//   ImageRectifier::computeDisparityMapRight / Left   Rectification/rectification.cpp:66-88  (without the warp: host OpenCV)
//   main.cpp:41-65                                    8-bit hand-off, removeDisparityOutliers, convertDisparityToDepth, reconstruction
//   rectification_main.cpp:194-195                    LinearSearch(left, right).computeDisparityMap(1.0)
//   Tests/test_reconstruction.cpp:120                 scaleDisparityMap
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "WindowSearch/BlockSearch.h"
#include "WindowSearch/LinearSearch.h"
#include "Reconstruction/reconstruction.h"

static cv::Mat read_bgr(const std::string& path, int w, int h) {
    cv::Mat m(h, w, CV_8UC3);
    std::ifstream f(path, std::ios::binary);
    f.read(reinterpret_cast<char*>(m.data), (std::streamsize)w * h * 3);
    if (!f) { std::cerr << "cannot read " << path << std::endl; std::exit(2); }
    return m;
}
template <typename T>
static void dump(const std::string& path, const cv::Mat& m) {
    std::ofstream f(path, std::ios::binary);
    for (int y = 0; y < m.rows; y++) f.write(reinterpret_cast<const char*>(&m.at<T>(y, 0)), (std::streamsize)m.cols * sizeof(T));
}

// rectification.cpp:78-81 (ImageRectifier::computeDisparityMapRight before the inverse warp)
static cv::Mat computeDisparityMapRight(cv::Mat& leftRectifiedImage_, cv::Mat& rightRectifiedImage_, int blockSize, int minDisparity,
                                        int maxDisparity, double smoothFactor, bool varBlock = false, double thres = 10.0) {
    BlockSearch blockSearch = BlockSearch(leftRectifiedImage_, rightRectifiedImage_, blockSize, minDisparity, maxDisparity);
    cv::Mat disparityMapRight = blockSearch.computeDisparityMapRight(smoothFactor, varBlock, thres);
    return disparityMapRight;
}
// rectification.cpp:66-70
static cv::Mat computeDisparityMapLeft(cv::Mat& l, cv::Mat& r, int blockSize, int maxDisparity, double smoothFactor) {
    BlockSearch blockSearch = BlockSearch(l, r, blockSize, 0, maxDisparity);
    return blockSearch.computeDisparityMapLeft(smoothFactor);
}

int main(int argc, char** argv) {
    if (argc < 7) { std::cerr << "usage: dropin_caller left.raw right.raw w h maxD outdir" << std::endl; return 2; }
    const int w = std::atoi(argv[3]), h = std::atoi(argv[4]), maxD = std::atoi(argv[5]);
    const std::string out = argv[6];
    cv::Mat left = read_bgr(argv[1], w, h), right = read_bgr(argv[2], w, h);

    cv::Mat disparityMapRight = computeDisparityMapRight(left, right, 17, 0, maxD, 0.9);          // main.cpp:40
    dump<double>(out + "/right_b17_s09.bin", disparityMapRight);
    dump<double>(out + "/right_var.bin", computeDisparityMapRight(left, right, 5, 0, maxD, 1.0, true, 40.0));
    dump<double>(out + "/left_b9.bin", computeDisparityMapLeft(left, right, 9, maxD, 0.9));
    std::vector<int> blockSizes{7};
    // the vector constructor leaves blockSize_ uninitialised in the reference (BlockSearch.cpp:15-22): constructed, not used
    BlockSearch unused(left, right, blockSizes, 0, maxD);
    (void)unused;
    // LinearSearch reads left(i, k) up to k = j + 199 without a bound (LinearSearch.cpp:32-33): give it a left image that
    // is 199 columns wider (the image repeated) so that the reference build stays inside its buffer
    cv::Mat leftWide(h, w + 199, CV_8UC3);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w + 199; x++) leftWide.at<cv::Vec3b>(y, x) = left.at<cv::Vec3b>(y, x % w);
    LinearSearch linearSearch(leftWide, right);                                                  // rectification_main.cpp:194
    dump<double>(out + "/linear.bin", linearSearch.computeDisparityMap(1.0));

    // main.cpp:41-50: the 8-bit PNG hand-off (imwrite / readGrayscaleImageAsDisparityMap are host OpenCV: the same
    // rounding -- rint half to even, saturate -- written out here)
    cv::Mat disparityImage(h, w, CV_32FC1);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            double v = std::nearbyint(disparityMapRight.at<double>(y, x));
            disparityImage.at<float>(y, x) = (float)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    removeDisparityOutliers(disparityImage, 500, 1.5, 0.8);                                      // main.cpp:53
    dump<float>(out + "/disp_filtered.bin", disparityImage);
    float focalLength = 1733.74f, baseline = 1.f;
    cv::Mat depthValues = convertDisparityToDepth(disparityImage, focalLength, baseline);        // main.cpp:56-58
    dump<float>(out + "/depth.bin", depthValues);
    Matrix3f K = Matrix3f::Zero();
    K(0, 0) = 1733.74f; K(1, 1) = 1733.68f; K(0, 2) = w / 2.f; K(1, 2) = h / 2.f; K(2, 2) = 1.f;
    reconstruction(right, depthValues, K, 1.0f);                                                 // main.cpp:61-65 -> ../../results/reconstruction_mesh.off
    cv::Mat scaled = disparityImage.clone();
    scaleDisparityMap(scaled, 0.37f);                                                            // test_reconstruction.cpp:120
    dump<float>(out + "/scaled.bin", scaled);

    // WriteMesh / CheckTriangularValidity on caller-made vertices (reconstruction.h:24-26)
    std::vector<Vertex> vertices((size_t)6 * 4);
    for (int i = 0; i < 24; i++) {
        float z = 1.f + 0.3f * (float)((i * 7) % 5);
        vertices[i].position = (i == 9) ? Vector4f(MINF, MINF, MINF, MINF) : Vector4f((float)(i % 6) * 0.4f, (float)(i / 6) * 0.4f, z, 1.f);
        vertices[i].color = Vector4uc((unsigned char)(i * 10), (unsigned char)(255 - i), 7, 255);
    }
    std::ofstream tv(out + "/triangles.txt");
    for (unsigned i = 0; i + 7 < 24; i++) tv << CheckTriangularValidity(vertices.data(), i, i + 6, i + 1, 0.7f) << "\n";
    tv << WriteMesh(vertices.data(), 6, 4, out + "/small_mesh.off", 0.7f) << "\n";
    tv << WriteMesh(vertices.data(), 6, 4, out + "/no/such/dir/mesh.off", 0.7f) << "\n";
    return 0;
}
