// Device clip compiler (K5): raw mocap "Frames" -> the tables MocapDM.read_raw_data builds
// (reference src/mujoco/mocap_v2.py:33-336, src/mujoco/mocap_util.py:31-48), computed on the GPU in float64,
// plus the fp32 reward / state rows the step kernel indexes.  Implemented in dm_clip_compiler.cu, which is
// compiled with -fmad=false so that a*b+c is evaluated as numpy evaluates it (two roundings); the libm calls
// (sin, cos, atan2, sqrt) are CUDA's, 1-2 ulp from glibc's.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace dm {

struct CompiledClip {  // device pointers, owned by the caller of compile_clip_device (cudaFree each)
  int len = 0;         // frames after resampling = get_length()
  int n_key = 0;       // key frames in the file
  int k = 1;           // resampling factor int(dt / 0.01666)
  double dt = 0;       // MocapDM.dt after load
  double* qpos = nullptr;       // [len, nq]
  double* qvel = nullptr;       // [len, nv]
  double* body_xpos = nullptr;  // [len, nbody, 3]
  double* geom_xpos = nullptr;  // [len, ngeom, 3]
};

// d_frames: DEVICE [n_frames, n_cols] float64 rows of the "Frames" array.  Returns NULL or a static message.
// *launches is incremented by the number of kernels enqueued.
const char* compile_clip_device(int robot, const double* d_frames, int n_frames, int n_cols, int direct_qpos, int getup,
                                double dt, CompiledClip* out, long long* launches, cudaStream_t stream);
void free_compiled_clip(CompiledClip* c);

// fp32 rows of one clip (dm_step.cuh RowLayout) and its [nq + nv] state rows, written at row `row0` of the tables.
// Packs rows [first, first + n_rows) of the compiled clip.
const char* pack_rows_device(int robot, const CompiledClip& c, int first, int n_rows, int row0, float* d_rows,
                             float* d_state_rows, long long* launches, cudaStream_t stream);

}  // namespace dm
