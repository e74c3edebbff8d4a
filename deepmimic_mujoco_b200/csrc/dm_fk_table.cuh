// K3: float64 forward kinematics of clip key frames (load time) -- shared by libdmhot's dm_fk_tables and the
// device clip compiler (dm_clip_compiler.cu).
#pragma once
#include <cuda_runtime.h>

#include "dm_math.cuh"

namespace dm {

// ---- K3: float64 FK of key frames --------------------------------------------------------------
struct TableIO {
  const double* qp;
  const double* qv;
  __device__ __forceinline__ double qpos(int j) const { return qp[j]; }
  __device__ __forceinline__ double qvel(int j) const { return qv[j]; }
};
struct TableSink {
  static constexpr bool kWantsGeoms = true;
  __device__ static constexpr bool contact_on() { return false; }
  double* bx;
  double* gx;
  __device__ __forceinline__ void body(int i, double x, double y, double z) {
    bx[3 * i] = x; bx[3 * i + 1] = y; bx[3 * i + 2] = z;
  }
  __device__ __forceinline__ void geom(int g, double x, double y, double z) {
    gx[3 * g] = x; gx[3 * g + 1] = y; gx[3 * g + 2] = z;
  }
};

template <class Rb>
__global__ void fk_table_kernel(int n, const double* qpos, const double* qvel, double* body_xpos, double* geom_xpos) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n) return;
  TableIO in{qpos + (size_t)f * Rb::NQ, qvel + (size_t)f * Rb::NV};
  TableSink sink{body_xpos + (size_t)f * Rb::NBODY * 3, geom_xpos + (size_t)f * Rb::NGEOM * 3};
  // world body and its geoms (the floor plane at the origin)
  sink.body(0, 0.0, 0.0, 0.0);
  for (int g = 0; g < Rb::NGEOM; ++g) sink.geom(g, 0.0, 0.0, 0.0);
  FkOut<double> o;
  Rb::template fk<double>(in, o, sink);
}

}  // namespace dm
