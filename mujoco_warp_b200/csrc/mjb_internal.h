// mjb_internal.h -- what the translation units of libmjb200.so share besides the public C ABI (not exported to users).
//   mjb_api.cu      the step pipeline (heavy: every stage-group kernel), handle definitions
//   mjb_support.cu  state packing, masked reset, contact forces (light: compiles in seconds)
//   mjb_rk4.cu      Runge-Kutta 4 integrator around the step pipeline (light)
#pragma once
#include "mjb200.h"
#include "mjb_types.cuh"

// Fills the kernel parameter block of a (model, data) pair with a zeroed option block; returns 0 or a negative error code.
int mjb_internal_params(const MjbModel* m, const MjbData* d, KParams* K);
// Records the thread's last error message (mjb_last_error) and returns `code`.
int mjb_internal_fail(int code, const char* msg);
// Counts one kernel launch of this library (mjb_launch_count).
void mjb_internal_count_launch(void);
// Runge-Kutta scratch owned by the data handle: a device buffer of at least `floats` floats (cudaMalloc on first use or growth, freed by
// mjb_data_destroy), or nullptr with the error recorded.
float* mjb_internal_rk_scratch(MjbData* d, size_t floats);
