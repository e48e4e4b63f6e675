// adjoint.cuh -- path-replay adjoint kernels (prbvolpath / prb_volpathaos). Filled in below.
#pragma once
#include "kernels.cuh"
