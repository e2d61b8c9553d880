// TEST-ONLY stand-in for the reference's src/Eigen.h (which cannot be copied into this repository and does not exist on
// the GPU box): the names the shim headers and the drop-in caller take from it -- MINF, Vector4f, Vector4uc, Matrix3f,
// `using namespace Eigen` -- over the Eigen storage stand-in of oracle/refstub.  In the build container the same
// sources are ALSO compiled against the reference's real src/Eigen.h (tests/test_shim_dropin.py).
#pragma once
#include <limits>
#ifndef MINF
#define MINF -std::numeric_limits<float>::infinity()
#endif
#include <Eigen/Dense>
typedef Eigen::Matrix<unsigned char, 4, 1> Vector4uc;
using namespace Eigen;
