"""The drop-in boundary, compiled and run.  INTEGRATION.md 2 is carried out in a scratch directory -- the shim headers
take the place of the reference's BlockSearch.h / LinearSearch.h / reconstruction.h, sr_shim_common.h and sr_capi.h sit
next to them -- and tests/cpp/dropin_caller.cpp (call sites shaped like rectification.cpp:66-88, main.cpp:40-65,
rectification_main.cpp:194-195) is compiled against that tree:
  * CPU: the swapped tree compiles and links (OpenCV / Eigen facilities through the test-only stand-ins of
    oracle/refstub; in the build container also against the reference's REAL src/Eigen.h, syntax only);
  * GPU: the shim-built caller (libsr_b200.so) and the reference-built caller (oracle/_ref/dropin_ref = the same source
    linked with the reference's own .cpp files) run on the same images and must write byte-identical files -- disparity
    maps, filtered map, depth, scaled map, both mesh files, the validity list -- and print the same lines."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "stereo_reconstruction_b200", "shim")
STUB = os.path.join(ROOT, "oracle", "refstub")
CALLER = os.path.join(ROOT, "tests", "cpp", "dropin_caller.cpp")
LIBDIR = os.path.join(ROOT, "stereo_reconstruction_b200")
REF_SRC = "/root/reference/src"
GXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"


def swapped_tree(dst, eigen_h):
    """INTEGRATION.md 2, verbatim: cp shim/WindowSearch/*.h src/WindowSearch/; cp shim/Reconstruction/*.h
    src/Reconstruction/; cp shim/sr_shim_common.h include/sr_capi.h src/   (src/Eigen.h is the reference's own)."""
    src = os.path.join(dst, "src")
    for d in ("WindowSearch", "Reconstruction"):
        os.makedirs(os.path.join(src, d))
        for f in os.listdir(os.path.join(SHIM, d)):
            shutil.copy(os.path.join(SHIM, d, f), os.path.join(src, d, f))
    shutil.copy(os.path.join(SHIM, "sr_shim_common.h"), src)
    shutil.copy(os.path.join(ROOT, "include", "sr_capi.h"), src)
    shutil.copy(eigen_h, os.path.join(src, "Eigen.h"))
    return src


def build_shim_caller(tmp, eigen_h=os.path.join(ROOT, "tests", "cpp", "standin_src", "Eigen.h"), syntax_only=False):
    src = swapped_tree(str(tmp), eigen_h)
    exe = os.path.join(str(tmp), "dropin_b200")
    cmd = [GXX, "-std=c++17", "-O2", "-ffp-contract=off", "-Wall", "-I", src, "-I", STUB, CALLER]
    cmd += ["-fsyntax-only"] if syntax_only else ["-o", exe, "-L", LIBDIR, "-lsr_b200", f"-Wl,-rpath,{LIBDIR}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]
    return exe


def test_swapped_tree_compiles_and_links(tmp_path):
    exe = build_shim_caller(tmp_path)
    assert os.path.exists(exe)
    # every hot-path symbol the shims bind comes from libsr_b200.so, none from the reference's .cpp files
    syms = subprocess.run(["nm", "-u", exe], capture_output=True, text=True).stdout
    for s in ("sr_block_search", "sr_remove_outliers", "sr_disparity_to_depth", "sr_backproject", "sr_write_mesh", "sr_scale"):
        assert s in syms


@pytest.mark.skipif(not os.path.isdir(REF_SRC), reason="needs the reference tree (build container)")
def test_swapped_tree_compiles_with_the_references_own_eigen_h(tmp_path):
    build_shim_caller(tmp_path, eigen_h=os.path.join(REF_SRC, "Eigen.h"), syntax_only=True)


@pytest.mark.skipif(not os.path.isdir(REF_SRC), reason="needs the reference tree (build container)")
def test_reference_call_sites_compile_against_the_shims(tmp_path):
    """The reference's real call sites: a TU that includes the swapped headers the way Rectification/rectification.hpp
    and main.cpp do and repeats rectification.cpp:66-88 / main.cpp:53-65 literally (those files themselves also need
    OpenCV calib3d / Ceres / the Middlebury SDK, which no stand-in provides)."""
    src = swapped_tree(str(tmp_path), os.path.join(REF_SRC, "Eigen.h"))
    tu = os.path.join(str(tmp_path), "callsites.cpp")
    open(tu, "w").write('''
#include "WindowSearch/BlockSearch.h"
#include "Reconstruction/reconstruction.h"
struct ImageRectifier {
    cv::Mat leftRectifiedImage_, rightRectifiedImage_;
    cv::Mat computeDisparityMapLeft(int blockSize, int maxDisparity, double smoothFactor) {
        BlockSearch blockSearch = BlockSearch(leftRectifiedImage_, rightRectifiedImage_, blockSize, 0, maxDisparity);
        cv::Mat disparityMapLeft = blockSearch.computeDisparityMapLeft(smoothFactor);
        return disparityMapLeft;
    }
    cv::Mat computeDisparityMapRight(int blockSize, int minDisparity, int maxDisparity, double smoothFactor, bool varBlock = false, double thres = 10.0) {
        BlockSearch blockSearch = BlockSearch(leftRectifiedImage_, rightRectifiedImage_, blockSize, minDisparity, maxDisparity);
        cv::Mat disparityMapRight = blockSearch.computeDisparityMapRight(smoothFactor, varBlock, thres);
        return disparityMapRight;
    }
};
void main_cpp_53_65(cv::Mat disparityImage, cv::Mat imageRight, Matrix3f intrinsics) {
    removeDisparityOutliers(disparityImage, 500, 1.5, 0.8);
    float focalLength = intrinsics(0, 0);
    float baseline = 1.f;
    cv::Mat depthValues = convertDisparityToDepth(disparityImage, focalLength, baseline);
    float thrMesh = 1.f;
    reconstruction(imageRight, depthValues, intrinsics, thrMesh);
}
''')
    r = subprocess.run([GXX, "-std=c++17", "-fsyntax-only", "-Wall", "-I", src, "-I", STUB, tu], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]


@pytest.mark.gpu
def test_shim_caller_equals_reference_caller(tmp_path, orc):
    ref_exe = os.path.join(ROOT, "oracle", "_ref", "dropin_ref")
    if not os.path.exists(ref_exe):
        pytest.skip("oracle/_ref/dropin_ref not built (needs the reference tree at build time)")
    exe = build_shim_caller(tmp_path / "build")
    w, h, D = 236, 120, 40
    L, R = orc.synth_pair(0xB200, 5, w, h, 0, D)
    R[30:36, 50:70] = 0                                            # black pixels are skipped
    L.tofile(str(tmp_path / "left.raw")); R.tofile(str(tmp_path / "right.raw"))
    outs, logs = {}, {}
    for tag, binary in (("ref", ref_exe), ("b200", exe)):
        root = tmp_path / tag
        cwd = root / "a" / "b"                                     # reconstruction() writes ../../results/reconstruction_mesh.off
        os.makedirs(cwd); os.makedirs(root / "results"); os.makedirs(root / "out")
        r = subprocess.run([binary, str(tmp_path / "left.raw"), str(tmp_path / "right.raw"), str(w), str(h), str(D), str(root / "out")],
                           cwd=str(cwd), capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, (tag, r.stderr[-2000:])
        logs[tag] = r.stdout
        outs[tag] = {f: open(os.path.join(root, "out", f), "rb").read() for f in sorted(os.listdir(root / "out")) if os.path.isfile(root / "out" / f)}
        outs[tag]["reconstruction_mesh.off"] = open(root / "results" / "reconstruction_mesh.off", "rb").read()
    assert sorted(outs["ref"]) == sorted(outs["b200"]) and len(outs["ref"]) == 10
    for f in outs["ref"]:
        assert outs["ref"][f] == outs["b200"][f], f"{f} differs between the reference build and the shim build"
    assert logs["ref"] == logs["b200"]                              # "max block size: ..", "h1:..", "h2:..", "Failed to write mesh!" lines
    assert "max block size: 17" in logs["b200"] and f"h1:{h}" in logs["b200"]
    d = np.frombuffer(outs["b200"]["right_b17_s09.bin"], np.float64).reshape(h, w)
    np.testing.assert_array_equal(d, orc.block_search_right_fast(L, R, 17, 0, D, 0.9))
