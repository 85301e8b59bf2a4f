// Oracle shim (test infrastructure only): an empty mkl.h.  The reference headers do
// `#define EIGEN_USE_MKL_ALL` + `#include "mkl.h"` (e.g. iLQR.h:3-4).  With no INTEL_MKL_VERSION
// defined, Eigen's MKL_support.h disables its MKL back-end by itself, so the reference compiles
// against plain Eigen kernels.  oneMKL is not installable in this image.
#pragma once
