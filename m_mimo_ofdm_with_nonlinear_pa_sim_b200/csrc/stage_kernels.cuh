// Stage-level kernels behind the drop-in building-block classes (filled in below).
#pragma once
#include "prep_kernels.cuh"
namespace mmimo {}
static int stage_kernels_init() { return 0; }
