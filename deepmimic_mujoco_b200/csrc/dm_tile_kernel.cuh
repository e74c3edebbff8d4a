// K1b: shared-memory tile kernel (TMA bulk copies) -- see below.
#pragma once
#include "dm_kernels.cuh"
