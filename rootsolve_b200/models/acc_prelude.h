// Included by the accessor-form device models before model_math.h: the point functions then take the SDK's rs_state
// (y[i] carries the finite-difference increment of the evaluation) and are device-only.
#pragma once
#include "rootsolve_b200_model.cuh"
#define RSM_STATE rs_state
#define RSM_HD __device__ __forceinline__
#include "model_math.h"
