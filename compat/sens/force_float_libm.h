// Sensitivity study only (tools/oracle_sensitivity.py, -include'd when the reference is compiled as a variant):
// choose the overload of the libm calls the reference makes unqualified on FLOAT arguments
// (frustum_culling.cpp:91-95 tan, voxel_grid_occlusion_estimation.cpp:560 sqrt).  Calls on double arguments keep the
// double function.
#pragma once
#include <cmath>  // NOT <math.h>: its C++ wrapper pulls the float overloads into the global namespace, which is the very
                  // thing being varied here
#if defined(ASSCPP_SENS_TANF) || defined(ASSCPP_SENS_TAN_DOUBLE)
inline double asscpp_sens_tan(double x) { return __builtin_tan(x); }
#ifdef ASSCPP_SENS_TANF
inline float asscpp_sens_tan(float x) { return __builtin_tanf(x); }
#else
inline double asscpp_sens_tan(float x) { return __builtin_tan((double)x); }
#endif
namespace std { using ::asscpp_sens_tan; }
#define tan asscpp_sens_tan
#endif
#if defined(ASSCPP_SENS_SQRTF) || defined(ASSCPP_SENS_SQRT_DOUBLE)
inline double asscpp_sens_sqrt(double x) { return __builtin_sqrt(x); }
#ifdef ASSCPP_SENS_SQRTF
inline float asscpp_sens_sqrt(float x) { return __builtin_sqrtf(x); }
#else
inline double asscpp_sens_sqrt(float x) { return __builtin_sqrt((double)x); }
#endif
namespace std { using ::asscpp_sens_sqrt; }
#define sqrt asscpp_sens_sqrt
#endif
