// TEST INFRASTRUCTURE (oracle).  Force-included (-include) when building
// libmpenv_ref_det.so: routes the five transcendental functions the
// reference's simulation code calls (sim.cpp / utils.cpp / brt/math.inl:
// sinf, cosf, logf, atan2f, asinf) to the project's bit-reproducible
// implementations in madrona-mp-env_b200/csrc/mp_math.h, so the reference
// code and the CUDA kernels execute the same IEEE operation sequence.
// tanf (sim.cpp:5871, one-time frustum setup) stays libm on both sides.
#pragma once
#include <cmath>
#include <math.h>
#include <cstdlib>
#include <algorithm>
#include <array>
#include <bit>
#include "../madrona-mp-env_b200/csrc/mp_math.h"
#define sinf(x) ::mpenv::mp_sinf(x)
#define cosf(x) ::mpenv::mp_cosf(x)
#define logf(x) ::mpenv::mp_logf(x)
#define atan2f(y, x) ::mpenv::mp_atan2f(y, x)
#define asinf(x) ::mpenv::mp_asinf(x)
