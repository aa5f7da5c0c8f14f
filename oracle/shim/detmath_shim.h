// Force-included (-include) into every C++ file of the reference's opende when the
// PARITY oracle is built.  It redirects the three libm calls the quickstep path makes
// (dAtan2 / dSin / dCos -> atan2 / sin / cos, include/gazebo/ode/common.h:189-192) to
// the reproducible implementations of include/gz_detmath.h, so that the CPU oracle and
// the GPU kernels evaluate bit-identical transcendentals.  No reference source is
// modified; this is a choice of math library for the oracle build.
// The timed baseline flavour ("fast") is built WITHOUT this header and uses glibc's libm.
#pragma once
#include <math.h>
#ifdef __cplusplus
#include <cmath>
#include <complex>
#include <valarray>
#endif
#include "gz_detmath.h"
#define atan2(y, x) gzdet_atan2((y), (x))
#define sin(x) gzdet_sin((x))
#define cos(x) gzdet_cos((x))
