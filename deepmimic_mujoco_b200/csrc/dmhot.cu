// libdmhot.so -- C ABI over the kernels (see include/dmhot.h for the contract and the reference
// code each entry point replaces).  Host side is plain C++: argument checking, packing of clip
// tables, launches.  No CPU implementation of the step exists in this library.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/dmhot.h"
#include "dm_kernels.cuh"
#include "dm_tile_launch.cuh"
#include "dm_team_launch.cuh"
#include "dm_soa_kernel.cuh"
#include "dm_pack.h"
#include <chrono>
#include "dm_copy_pool.h"
#include "dm_clip_compiler.h"
#include "dm_policy.cuh"
#include "dm_policy_tc.cuh"

using namespace dm;

static thread_local std::string g_last_error;

struct Stage {  // one pinned host area + its device mirror
  void* h_pin = nullptr;
  void* d_mem = nullptr;
  size_t bytes = 0;
};

struct dm_handle {
  int device = 0, robot = 0, kind = 0, num_sms = 148;
  dm_dims_t dims{};
  std::string err;
  // clip tables
  float* d_rows = nullptr;        // packed reward rows [total_rows, row_stride]
  float* d_state_rows = nullptr;  // [total_rows, nq + nv]
  std::vector<float> h_rows;
  int total_rows = 0;
  std::vector<CompiledClip> compiled;  // float64 device tables of dm_compile_clips (entries of same_as clips are empty)
  std::vector<int> compiled_src;       // clip -> index into `compiled` that holds its tables
  ClipSet cs{};
  // statistics
  double* d_stats = nullptr;
  // ... their collective (dm_peer.cuh): snapshot, global totals, sequence number, error word, peer-mapped buffers
  double* d_snap = nullptr;
  double* d_total = nullptr;
  PeerState* d_peer_state = nullptr;
  PeerComm* peer_comm[kPeerMaxWorld] = {};  // [peer_rank] = own cudaMalloc'd buffer, the others cudaIpc mappings
  int peer_rank = 0, peer_world = 1;        // world 1 until dm_stats_peer_import has connected the ranks
  int auto_every = 0;                       // dm_stats_set_auto: an exchange with every auto_every-th dm_step with statistics
  long long auto_count = 0;
  struct SoaMapEntry { const void *q, *v, *o; long long ld; int n; SoaMaps maps; };
  std::vector<SoaMapEntry> soa_maps;  // tensor maps of the SoA buffer sets seen lately (encoding costs a few us per call)
  int exch_where = 0;                        // where the auto mode's last exchange went: 0 nowhere / ordered, 1 the step's stream, 2 the side stream
  bool fuse_now = false, fused_done = false;  // this dm_step is due an exchange / the tile kernel carried it (exchange CTA)
  cudaStream_t coll_stream = nullptr;       // side stream of the auto mode
  cudaEvent_t coll_event = nullptr;
  // policy (K6)
  float* d_pol_packed = nullptr;  // weight chunks in consumption order
  uint8_t* d_pol_packed16 = nullptr;  // the same stream as fp16 hi / lo pairs for the tensor-core kernel
  int pol_variant = 1;            // 0 = fp32 CUDA cores (K6a), 1 = tcgen05 (K6b)
  bool pol_tc_smem_set = false;
  int pol_trace_cta = 0;
  long long* pol_trace = nullptr;  // debug: device buffer [64][8] of clock64 marks (dm_policy_set_trace)
  float* d_pol_bias = nullptr;    // [h0 | h1 | PHEADC]
  float* d_pol_norm = nullptr;    // [mean | inv_std | reward scale]
  PolicyParams pol{};             // dims and chunk counts of the uploaded network (pointers filled per launch)
  float pol_clip_obs = 10.0f, pol_clip_r = 10.0f;
  bool pol_smem_set = false;
  // host-mode staging (the step's area doubles as the caller-visible pinned buffers of dm_host_buffers)
  Stage step_stage, reset_stage;
  cudaStream_t host_stream = nullptr;
  cudaStream_t host_stream2 = nullptr;  // mapped host step: chunks alternate between the two streams
  cudaEvent_t obs_event = nullptr;
  cudaEvent_t chunk_event[16] = {};     // mapped host step: one per chunk in flight
  int host_mode = DM_HOST_MAPPED;       // DM_HOST_*; DM_HOST_MODE overrides
  int row_stage_override = -1;          // DM_ROW_STAGE=0: never stage per-environment rows
  int wide_override = -1;               // DM_WIDE=0: never use the two-group instantiation, 2: always
  int pdl = 1;                          // programmatic dependent launch of the tile kernel (DM_PDL=0 turns it off)
  int host_trace = 0;                   // DM_HOST_TRACE: print the mean host time of every phase of the mapped step at destroy
  double ht[5] = {0, 0, 0, 0, 0};       // copy-in, launch, wait, copy-out, calls
  bool contact_slot_dirty = true, clip_slot_dirty = true;  // the pinned contact / clip slots may hold non-zero words
  size_t slots_n = 0;                   // batch size the slots were last carved for (another n moves them)
  int host_chunk = 0;                   // environments per chunk of the mapped host step (0 = auto; DM_HOST_CHUNK)
  std::unique_ptr<CopyPool> pool;  // helper threads for the staging copies of big host batches (created on first use)
  bool pool_small = false;         // the pool was created for a small batch (no helper threads)
  int copy_threads = -1;           // -1 = auto (3 helpers when the machine has the cores), DM_COPY_THREADS overrides
  // bookkeeping
  long long launches = 0;
  int variant = -1;  // -1 auto (= 1), 0 strided only (K1a), 1 tile kernel (K1b), 2 team kernel (K1c; opt-in)
  unsigned long long* d_team_trace = nullptr;  // DM_TEAM_TRACE: device buffer [32] of %globaltimer marks (printed at destroy)
  int team_maxt = 0, team_nin = 0, team_teams = 0, team_table = -1;  // tuning knobs of the team kernel (DM_TEAM_*; 0 / -1 = auto)
  int contact_mode = 0;  // DM_CONTACT_*
  int table_smem_override = -1, max_groups_override = 0, nbuf_override = 0, skip_compute = 0;  // tuning knobs (DM_TABLE_SMEM, DM_MAX_GROUPS)
  bool timing = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev;
  size_t ev_used = 0;
  double t_ms = 0;
  long long t_n = 0;
};

static int fail(dm_handle* h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_last_error = buf;
  if (h) h->err = buf;
  return code;
}
#define CU(h, call)                                                                                  \
  do {                                                                                               \
    cudaError_t e_ = (call);                                                                         \
    if (e_ != cudaSuccess) return fail(h, DM_E_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

template <class F>
static int dispatch(const dm_handle* h, F&& f) {  // robot x env kind -> template instantiation
  if (h->robot == DM_ROBOT_HUMANOID3D && h->kind == DM_ENV_SINGLE) return f(Robot_humanoid3d{}, std::integral_constant<int, KIND_SINGLE>{});
  if (h->robot == DM_ROBOT_UNITREE_G1 && h->kind == DM_ENV_SINGLE) return f(Robot_unitree_g1{}, std::integral_constant<int, KIND_SINGLE>{});
  if (h->robot == DM_ROBOT_UNITREE_G1 && h->kind == DM_ENV_COMBINED) return f(Robot_unitree_g1{}, std::integral_constant<int, KIND_COMBINED>{});
  return DM_E_ARG;
}

extern "C" int dm_abi_version(void) { return DM_ABI_VERSION; }

extern "C" const char* dm_last_error(const dm_handle* h) { return h ? h->err.c_str() : g_last_error.c_str(); }

extern "C" int dm_create(int device, int robot, int env_kind, dm_handle** out) {
  if (!out) return fail(nullptr, DM_E_ARG, "dm_create: out is NULL");
  *out = nullptr;
  if (robot != DM_ROBOT_HUMANOID3D && robot != DM_ROBOT_UNITREE_G1) return fail(nullptr, DM_E_ARG, "dm_create: unknown robot %d", robot);
  if (env_kind != DM_ENV_SINGLE && env_kind != DM_ENV_COMBINED) return fail(nullptr, DM_E_ARG, "dm_create: unknown env kind %d", env_kind);
  if (env_kind == DM_ENV_COMBINED && robot != DM_ROBOT_UNITREE_G1)
    return fail(nullptr, DM_E_ARG, "dm_create: DPCombinedEnv exists for unitree_g1 only (combined_env.py:164)");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(nullptr, DM_E_CUDA, "dm_create: no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return fail(nullptr, DM_E_ARG, "dm_create: device %d out of range (%d devices)", device, ndev);
  CU(nullptr, cudaSetDevice(device));
  dm_handle* h = new dm_handle();
  cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device);
  h->device = device;
  h->robot = robot;
  h->kind = env_kind;
  if (const char* e = getenv("DM_TABLE_SMEM")) h->table_smem_override = atoi(e);
  if (const char* e = getenv("DM_SKIP_COMPUTE")) h->skip_compute = atoi(e);
  if (const char* e = getenv("DM_NBUF")) h->nbuf_override = atoi(e);
  if (const char* e = getenv("DM_MAX_GROUPS")) h->max_groups_override = atoi(e);
  if (const char* e = getenv("DM_POLICY_VARIANT")) h->pol_variant = atoi(e) ? 1 : 0;
  if (const char* e = getenv("DM_COPY_THREADS")) h->copy_threads = atoi(e);
  if (const char* e = getenv("DM_HOST_MODE")) h->host_mode = atoi(e) ? DM_HOST_MAPPED : DM_HOST_STAGED;
  if (const char* e = getenv("DM_HOST_CHUNK")) h->host_chunk = atoi(e);
  if (const char* e = getenv("DM_HOST_TRACE")) h->host_trace = atoi(e);
  if (const char* e = getenv("DM_PDL")) h->pdl = atoi(e);
  if (const char* e = getenv("DM_WIDE")) h->wide_override = atoi(e);
  if (const char* e = getenv("DM_ROW_STAGE")) h->row_stage_override = atoi(e);
  if (const char* e = getenv("DM_VARIANT")) h->variant = atoi(e);
  if (const char* e = getenv("DM_TEAM_MAXT")) h->team_maxt = atoi(e);
  if (const char* e = getenv("DM_TEAM_NIN")) h->team_nin = atoi(e);
  if (const char* e = getenv("DM_TEAM_TEAMS")) h->team_teams = atoi(e);
  if (const char* e = getenv("DM_TEAM_TABLE")) h->team_table = atoi(e);
  if (getenv("DM_TEAM_TRACE")) {
    cudaMalloc(&h->d_team_trace, 32 * sizeof(unsigned long long));
    cudaMemset(h->d_team_trace, 0, 32 * sizeof(unsigned long long));
  }
  dispatch(h, [&](auto rb, auto kind) -> int {
    using Rb = decltype(rb);
    using OL = ObsLayout<decltype(kind)::value, Rb>;
    dm_dims_t& d = h->dims;
    d.nq = Rb::NQ; d.nv = Rb::NV; d.nu = Rb::NU; d.na = Rb::NA; d.obs_dim = OL::DIM;
    d.nbody = Rb::NBODY; d.ngeom = Rb::NGEOM; d.nsel = Rb::NSEL; d.row_stride = RowLayout<Rb>::STRIDE;
    d.n_stats = 6 + 2 * OL::DIM; d.robot = robot; d.env_kind = env_kind;
    return 0;
  });
  if (cudaMalloc(&h->d_stats, sizeof(double) * h->dims.n_stats) != cudaSuccess ||
      cudaMemset(h->d_stats, 0, sizeof(double) * h->dims.n_stats) != cudaSuccess ||
      cudaStreamCreateWithFlags(&h->host_stream, cudaStreamNonBlocking) != cudaSuccess) {
    int rc = fail(nullptr, DM_E_CUDA, "dm_create: %s", cudaGetErrorString(cudaGetLastError()));
    delete h;
    return rc;
  }
  *out = h;
  return DM_OK;
}

extern "C" int dm_destroy(dm_handle* h) {
  if (!h) return DM_OK;
  cudaSetDevice(h->device);
  if (h->d_team_trace) {
    unsigned long long t[32];
    cudaDeviceSynchronize();
    cudaMemcpy(t, h->d_team_trace, sizeof t, cudaMemcpyDeviceToHost);
    const char* nm[10] = {"start", "dep-wait done", "tile+table landed", "part done", "barrier A", "combine done", "stats done",
                          "loop end", "store read", "end"};
    for (int w = 0; w < 2; ++w)
      for (int k = 1; k < 10; ++k)
        if (t[w * 16 + k]) fprintf(stderr, "[dmhot] team trace warp %d  %-18s +%6.2f us\n", w, nm[k], (double)(t[w * 16 + k] - t[0]) * 1e-3);
    cudaFree(h->d_team_trace);
  }
  // nothing may still be reading or writing the pinned blocks freed below
  if (h->host_stream) cudaStreamSynchronize(h->host_stream);
  if (h->host_stream2) cudaStreamSynchronize(h->host_stream2);
  cudaFree(h->d_rows);
  cudaFree(h->d_state_rows);
  cudaFree(h->d_stats);
  for (int r = 0; r < kPeerMaxWorld; ++r)
    if (h->peer_comm[r]) {
      if (r == h->peer_rank) cudaFree(h->peer_comm[r]); else cudaIpcCloseMemHandle(h->peer_comm[r]);
    }
  if (h->coll_stream) { cudaStreamSynchronize(h->coll_stream); cudaStreamDestroy(h->coll_stream); }
  if (h->coll_event) cudaEventDestroy(h->coll_event);
  cudaFree(h->d_snap);
  cudaFree(h->d_total);
  cudaFree(h->d_peer_state);
  cudaFree(h->d_pol_packed);
  cudaFree(h->d_pol_packed16);
  cudaFree(h->d_pol_bias);
  cudaFree(h->d_pol_norm);
  for (auto& c : h->compiled) free_compiled_clip(&c);
  for (Stage* st : {&h->step_stage, &h->reset_stage}) {
    cudaFree(st->d_mem);
    if (st->h_pin) cudaFreeHost(st->h_pin);
  }
  for (auto& p : h->ev) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
  if (h->obs_event) cudaEventDestroy(h->obs_event);
  if (h->host_trace && h->ht[4] > 0)
    fprintf(stderr, "[dmhot] mapped host step, mean over %.0f calls: copy-in %.1f us, launch %.1f us, wait %.1f us, copy-out %.1f us\n",
            h->ht[4], h->ht[0] / h->ht[4], h->ht[1] / h->ht[4], h->ht[2] / h->ht[4], h->ht[3] / h->ht[4]);
  for (cudaEvent_t e : h->chunk_event)
    if (e) cudaEventDestroy(e);
  if (h->host_stream2) cudaStreamDestroy(h->host_stream2);
  if (h->host_stream) cudaStreamDestroy(h->host_stream);
  delete h;
  return DM_OK;
}

extern "C" int dm_get_dims(const dm_handle* h, dm_dims_t* out) {
  if (!h || !out) return fail(nullptr, DM_E_ARG, "dm_get_dims: NULL argument");
  *out = h->dims;
  return DM_OK;
}

extern "C" int dm_get_model(const dm_handle* h, double* body_mass, float* lim_lo, float* lim_hi, float* gear, float* ctrl_lo,
                            float* ctrl_hi, int32_t* ee_geom) {
  if (!h) return fail(nullptr, DM_E_ARG, "dm_get_model: NULL handle");
  return dispatch(h, [&](auto rb, auto) -> int {
    using Rb = decltype(rb);
    for (int i = 0; i < Rb::NBODY; ++i) if (body_mass) body_mass[i] = Rb::body_mass(i);
    for (int k = 0; k < Rb::NSEL; ++k) {
      if (lim_lo) lim_lo[k] = Rb::lim_lo(k);
      if (lim_hi) lim_hi[k] = Rb::lim_hi(k);
    }
    for (int k = 0; k < Rb::NU; ++k) {
      if (gear) gear[k] = Rb::gear(k);
      if (ctrl_lo) ctrl_lo[k] = Rb::ctrl_lo(k);
      if (ctrl_hi) ctrl_hi[k] = Rb::ctrl_hi(k);
    }
    for (int k = 0; k < 4; ++k) if (ee_geom) ee_geom[k] = Rb::ee_geom(k);
    return DM_OK;
  });
}

extern "C" int dm_fk_tables(dm_handle* h, int32_t n, const double* qpos, const double* qvel, double* body_xpos, double* geom_xpos) {
  if (!h || n <= 0 || !qpos || !qvel || !body_xpos || !geom_xpos) return fail(h, DM_E_ARG, "dm_fk_tables: bad argument");
  CU(h, cudaSetDevice(h->device));
  const dm_dims_t& d = h->dims;
  const size_t bq = sizeof(double) * n * d.nq, bv = sizeof(double) * n * d.nv, bb = sizeof(double) * n * d.nbody * 3,
               bg = sizeof(double) * n * d.ngeom * 3;
  double *dq = nullptr, *dv = nullptr, *db = nullptr, *dg = nullptr;
  int rc = DM_OK;
  auto cleanup = [&] { cudaFree(dq); cudaFree(dv); cudaFree(db); cudaFree(dg); };
#define CUF(call)                                                                         \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) { rc = fail(h, DM_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); cleanup(); return rc; } \
  } while (0)
  CUF(cudaMalloc(&dq, bq)); CUF(cudaMalloc(&dv, bv)); CUF(cudaMalloc(&db, bb)); CUF(cudaMalloc(&dg, bg));
  CUF(cudaMemcpy(dq, qpos, bq, cudaMemcpyHostToDevice));
  CUF(cudaMemcpy(dv, qvel, bv, cudaMemcpyHostToDevice));
  dispatch(h, [&](auto rb, auto) -> int {
    using Rb = decltype(rb);
    fk_table_kernel<Rb><<<(n + 63) / 64, 64>>>(n, dq, dv, db, dg);
    return 0;
  });
  h->launches++;
  CUF(cudaGetLastError());
  CUF(cudaMemcpy(body_xpos, db, bb, cudaMemcpyDeviceToHost));
  CUF(cudaMemcpy(geom_xpos, dg, bg, cudaMemcpyDeviceToHost));
#undef CUF
  cleanup();
  return DM_OK;
}

// Packs the reward rows on the host in float64 and uploads them as fp32 (dm_step.cuh: RowLayout).
extern "C" int dm_upload_clips(dm_handle* h, int32_t n_clips, const dm_clip_t* clips) {
  if (!h || !clips || n_clips <= 0 || n_clips > DM_MAX_CLIPS) return fail(h, DM_E_ARG, "dm_upload_clips: need 1..%d clips", DM_MAX_CLIPS);
  if (h->kind == DM_ENV_COMBINED && n_clips != 4)
    return fail(h, DM_E_ARG, "dm_upload_clips: DPCombinedEnv needs exactly 4 clips (walk, run, getup, to_getup)");
  CU(h, cudaSetDevice(h->device));
  return dispatch(h, [&](auto rb, auto) -> int {
    using Rb = decltype(rb);
    using RL = RowLayout<Rb>;
    ClipSet cs{};
    std::vector<float> rows, srows;
    int total = 0;
    const char* bad = pack_clip_rows<Rb>(clips, n_clips, cs, rows, srows, total);
    if (bad) return fail(h, DM_E_ARG, "dm_upload_clips: %s", bad);
    cudaFree(h->d_rows);
    cudaFree(h->d_state_rows);
    h->d_rows = h->d_state_rows = nullptr;
    CU(h, cudaMalloc(&h->d_rows, rows.size() * sizeof(float)));
    CU(h, cudaMalloc(&h->d_state_rows, srows.size() * sizeof(float)));
    CU(h, cudaMemcpy(h->d_rows, rows.data(), rows.size() * sizeof(float), cudaMemcpyHostToDevice));
    CU(h, cudaMemcpy(h->d_state_rows, srows.data(), srows.size() * sizeof(float), cudaMemcpyHostToDevice));
    cs.rows = h->d_rows;
    h->cs = cs;
    h->total_rows = total;
    h->h_rows.swap(rows);
    for (auto& c : h->compiled) free_compiled_clip(&c);
    h->compiled.clear();
    h->compiled_src.clear();
    return DM_OK;
  });
}

// The same tables, built on the device from the raw "Frames" rows (dm_clip_compiler.cu).
extern "C" int dm_compile_clips(dm_handle* h, int32_t n_clips, const dm_raw_clip_t* raw) {
  if (!h || !raw || n_clips <= 0 || n_clips > DM_MAX_CLIPS) return fail(h, DM_E_ARG, "dm_compile_clips: need 1..%d clips", DM_MAX_CLIPS);
  if (h->kind == DM_ENV_COMBINED && n_clips != 4)
    return fail(h, DM_E_ARG, "dm_compile_clips: DPCombinedEnv needs exactly 4 clips (walk, run, getup, to_getup)");
  CU(h, cudaSetDevice(h->device));
  std::vector<CompiledClip> comp(n_clips);
  std::vector<int> src(n_clips);
  auto drop = [&] { for (auto& c : comp) free_compiled_clip(&c); };
  for (int c = 0; c < n_clips; ++c) {
    const dm_raw_clip_t& r = raw[c];
    if (r.same_as >= 0) {
      if (r.same_as >= c || raw[r.same_as].same_as >= 0) { drop(); return fail(h, DM_E_ARG, "dm_compile_clips: clip %d: same_as must name an earlier compiled clip", c); }
      src[c] = r.same_as;
      continue;
    }
    src[c] = c;
    if (!r.frames || r.n_frames <= 0 || r.n_cols < 2) { drop(); return fail(h, DM_E_ARG, "dm_compile_clips: clip %d has no frames", c); }
    double* d_frames = nullptr;
    const size_t bytes = sizeof(double) * (size_t)r.n_frames * r.n_cols;
    if (cudaMalloc(&d_frames, bytes) != cudaSuccess || cudaMemcpy(d_frames, r.frames, bytes, cudaMemcpyHostToDevice) != cudaSuccess) {
      cudaFree(d_frames); drop();
      return fail(h, DM_E_CUDA, "dm_compile_clips: upload of clip %d: %s", c, cudaGetErrorString(cudaGetLastError()));
    }
    const char* bad = compile_clip_device(h->robot, d_frames, r.n_frames, r.n_cols, r.direct_qpos, r.getup, r.frames[0], &comp[c],
                                          &h->launches, h->host_stream);
    cudaFree(d_frames);
    if (bad) { drop(); return fail(h, DM_E_ARG, "dm_compile_clips: clip %d: %s", c, bad); }
  }
  ClipSet cs{};
  cs.n_clips = n_clips;
  int total = 0;
  for (int c = 0; c < n_clips; ++c) {
    const dm_raw_clip_t& r = raw[c];
    const int n_rows = comp[src[c]].len;
    if (r.fixed_row >= n_rows || (r.fixed_row >= 0 && r.virtual_len <= 0)) { drop(); return fail(h, DM_E_ARG, "dm_compile_clips: clip %d: fixed_row / virtual_len inconsistent", c); }
    // a clip that shares another's tables (MotionTransition) indexes that clip's rows: nothing is stored twice
    cs.clip[c] = ClipDesc{r.same_as >= 0 ? cs.clip[r.same_as].row0 : total, r.fixed_row >= 0 ? r.virtual_len : n_rows, r.flags, r.fixed_row};
    if (r.same_as < 0) total += n_rows;
  }
  const size_t stride = h->dims.row_stride, sstride = h->dims.nq + h->dims.nv;
  float *d_rows = nullptr, *d_srows = nullptr;
  if (cudaMalloc(&d_rows, sizeof(float) * total * stride) != cudaSuccess || cudaMalloc(&d_srows, sizeof(float) * total * sstride) != cudaSuccess) {
    cudaFree(d_rows); cudaFree(d_srows); drop();
    return fail(h, DM_E_NOMEM, "dm_compile_clips: %s", cudaGetErrorString(cudaGetLastError()));
  }
  for (int c = 0; c < n_clips; ++c) {
    if (raw[c].same_as >= 0) continue;
    const CompiledClip& k = comp[src[c]];
    const char* bad = pack_rows_device(h->robot, k, 0, k.len, cs.clip[c].row0, d_rows, d_srows, &h->launches, h->host_stream);
    if (bad) { cudaFree(d_rows); cudaFree(d_srows); drop(); return fail(h, DM_E_CUDA, "dm_compile_clips: %s", bad); }
  }
  std::vector<float> rows((size_t)total * stride);
  cudaError_t e = cudaMemcpyAsync(rows.data(), d_rows, rows.size() * sizeof(float), cudaMemcpyDeviceToHost, h->host_stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->host_stream);
  if (e != cudaSuccess) { cudaFree(d_rows); cudaFree(d_srows); drop(); return fail(h, DM_E_CUDA, "dm_compile_clips: %s", cudaGetErrorString(e)); }
  cudaFree(h->d_rows);
  cudaFree(h->d_state_rows);
  for (auto& c : h->compiled) free_compiled_clip(&c);
  h->d_rows = d_rows;
  h->d_state_rows = d_srows;
  cs.rows = d_rows;
  h->cs = cs;
  h->total_rows = total;
  h->h_rows.swap(rows);
  h->compiled.swap(comp);
  h->compiled_src.swap(src);
  return DM_OK;
}

extern "C" int dm_get_compiled_clip(const dm_handle* h, int32_t clip, int32_t* len, double* dt, double* qpos, double* qvel,
                                    double* body_xpos, double* geom_xpos) {
  if (!h) return fail(nullptr, DM_E_ARG, "dm_get_compiled_clip: NULL handle");
  dm_handle* hm = const_cast<dm_handle*>(h);
  if (clip < 0 || clip >= (int)h->compiled_src.size()) return fail(hm, DM_E_STATE, "dm_get_compiled_clip: clip %d was not built by dm_compile_clips", clip);
  const CompiledClip& c = h->compiled[h->compiled_src[clip]];
  if (len) *len = c.len;
  if (dt) *dt = c.dt;
  CU(hm, cudaSetDevice(h->device));
  const dm_dims_t& d = h->dims;
  const size_t n = c.len;
  if (qpos) CU(hm, cudaMemcpy(qpos, c.qpos, sizeof(double) * n * d.nq, cudaMemcpyDeviceToHost));
  if (qvel) CU(hm, cudaMemcpy(qvel, c.qvel, sizeof(double) * n * d.nv, cudaMemcpyDeviceToHost));
  if (body_xpos) CU(hm, cudaMemcpy(body_xpos, c.body_xpos, sizeof(double) * n * d.nbody * 3, cudaMemcpyDeviceToHost));
  if (geom_xpos) CU(hm, cudaMemcpy(geom_xpos, c.geom_xpos, sizeof(double) * n * d.ngeom * 3, cudaMemcpyDeviceToHost));
  return DM_OK;
}

extern "C" int dm_get_packed_rows(const dm_handle* h, float* rows, int32_t* total_rows) {
  if (!h || !total_rows) return fail(nullptr, DM_E_ARG, "dm_get_packed_rows: NULL argument");
  *total_rows = h->total_rows;
  if (rows) memcpy(rows, h->h_rows.data(), h->h_rows.size() * sizeof(float));
  return DM_OK;
}

// ---------------------------------------------------------------------------------------------
static int check_step_args(dm_handle* h, const dm_step_args* a, const char* who) {
  if (!h || !a) return fail(h, DM_E_ARG, "%s: NULL argument", who);
  if (!h->d_rows) return fail(h, DM_E_STATE, "%s: no clips uploaded (call dm_upload_clips first)", who);
  if (a->n < 0) return fail(h, DM_E_ARG, "%s: n = %d", who, a->n);
  if (a->n == 0) return DM_OK;  // an empty batch is a no-op; its buffers may be NULL
  if (a->layout != DM_LAYOUT_AOS && a->layout != DM_LAYOUT_SOA) return fail(h, DM_E_ARG, "%s: unknown layout %d", who, a->layout);
  if (a->layout == DM_LAYOUT_SOA && a->ld < a->n) return fail(h, DM_E_ARG, "%s: SoA ld %lld < n %d", who, (long long)a->ld, a->n);
  if (!a->qpos || !a->qvel || !a->action || !a->idx || !a->ep_len || !a->obs || !a->reward || !a->done)
    return fail(h, DM_E_ARG, "%s: a required buffer is NULL (qpos qvel action idx ep_len obs reward done)", who);
  if (h->kind == DM_ENV_COMBINED && !a->clip) return fail(h, DM_E_ARG, "%s: DPCombinedEnv needs the clip (motion state) buffer", who);
  return DM_OK;
}

static void timing_begin(dm_handle* h, cudaStream_t s) {
  if (!h->timing) return;
  if (h->ev_used == h->ev.size()) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    h->ev.emplace_back(a, b);
  }
  cudaEventRecord(h->ev[h->ev_used].first, s);
}
static void timing_end(dm_handle* h, cudaStream_t s) {
  if (!h->timing) return;
  cudaEventRecord(h->ev[h->ev_used].second, s);
  h->ev_used++;
}

template <class T>
static T* off(T* p, long long n) { return p ? p + n : p; }

static PeerParams peer_params(const dm_handle* h, int enable);

// 2-D tensor map of a row-major [rows, ld] fp32 array whose boxes are `box_rows` rows of 128 consecutive environments
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static int soa_encode(dm_handle* h, CUtensorMap* tm, const float* base, int rows, long long ld, int n, int box_rows) {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess || !p)
      return fail(h, DM_E_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    fn = (EncodeTiledFn)p;
  }
  const cuuint64_t dims[2] = {(cuuint64_t)n, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(float)};
  const cuuint32_t box[2] = {128u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(h, DM_E_CUDA, "cuTensorMapEncodeTiled failed (%d) for a [%d, %lld] array", (int)r, rows, ld);
  return DM_OK;
}

static int launch_step(dm_handle* h, const dm_step_args* a, cudaStream_t s) {
  if (a->n == 0) return DM_OK;
  return dispatch(h, [&](auto rb, auto kind) -> int {
    using Rb = decltype(rb);
    constexpr int KIND = decltype(kind)::value;
    using OL = ObsLayout<KIND, Rb>;
    using TC = TileCfg<KIND, Rb>;
    StepParams p{};
    p.n = a->n;
    p.aos = a->layout == DM_LAYOUT_AOS;
    p.es = p.aos ? 1 : a->ld;
    p.qpos = a->qpos; p.qvel = a->qvel; p.action = a->action; p.contact = a->contact;
    p.idx = a->idx; p.ep_len = a->ep_len; p.clip = a->clip;
    p.obs = a->obs; p.reward = a->reward; p.done = a->done; p.reason = a->reason;
    p.torque = a->torque; p.terms = a->terms; p.ep_return = a->ep_return;
    p.cs = h->cs;
    p.cs.contact_from_fk = h->contact_mode;
    // K1b (shared-memory tiles, TMA in / TMA out) for the full 128-env tiles of a row-major batch, K1a for the rest
    const bool aligned = ((reinterpret_cast<uintptr_t>(a->qpos) | reinterpret_cast<uintptr_t>(a->qvel) |
                           reinterpret_cast<uintptr_t>(a->obs) | reinterpret_cast<uintptr_t>(a->action) |
                           reinterpret_cast<uintptr_t>(a->torque)) & 15u) == 0;
    int tiled = 0;
    bool soa_tiled = false;
    const bool team = h->variant == 2;
    if (p.aos && team && aligned && a->n >= 32) {
      // K1c: tiles of 32 environments, one team of four warps per tile (dm_team_kernel.cuh)
      using MC = TeamCfg<KIND, Rb>;
      tiled = a->n / MC::TE * MC::TE;
      const int n_tiles = tiled / MC::TE;
      const size_t table_bytes = sizeof(float) * (size_t)h->total_rows * RowLayout<Rb>::STRIDE;
      const size_t avail = MC::SMEM_BUDGET - MC::FIXED_BYTES;
      const int want = (n_tiles + h->num_sms - 1) / h->num_sms;  // teams per CTA that cover the batch in one wave
      // instantiation by register budget: up to 4 teams per CTA get 128 registers, up to 6 get 80, 8 get 64
      int maxt = want <= kTeamMaxtWide ? kTeamMaxtWide : want <= kTeamMaxtMid ? kTeamMaxtMid : kTeamMaxtFull;
      if (a->terms) maxt = kTeamMaxtFull;
      if (h->team_maxt > 0) maxt = h->team_maxt;
      auto fit = [&](bool table, int nin) {  // teams that fit beside the table with nin input slots each
        if (table && table_bytes > avail) return 0;
        return (int)((avail - (table ? table_bytes : 0)) / (sizeof(float) * MC::team_floats(nin)));
      };
      int T = want < maxt ? (want < 1 ? 1 : want) : maxt;
      if (h->team_teams > 0 && h->team_teams < T) T = h->team_teams;
      // the clip table goes to shared memory when that does not cost a team; a second input slot when it costs neither
      bool table_smem = fit(true, 1) >= T;
      if (h->team_table >= 0) table_smem = h->team_table != 0 && fit(true, 1) >= 1;
      if (fit(table_smem, 1) < T) T = fit(table_smem, 1);
      const int tiles_per_team = (n_tiles + h->num_sms * T - 1) / (h->num_sms * T);
      int nin = (tiles_per_team >= 2 && fit(table_smem, 2) >= T) ? 2 : 1;
      if (h->team_nin > 0 && fit(table_smem, h->team_nin) >= T) nin = h->team_nin;
      const int grid = (n_tiles + T - 1) / T < h->num_sms ? (n_tiles + T - 1) / T : h->num_sms;
      const size_t smem = MC::FIXED_BYTES + (table_smem ? table_bytes : 0) + (size_t)T * sizeof(float) * MC::team_floats(nin);
      StepParams q = p;
      q.n = tiled;
      TeamExtra ex{a->accumulate_stats ? h->d_stats : nullptr, n_tiles, (int)(table_bytes / sizeof(float)), nin, h->skip_compute,
                   h->d_team_trace};
      if (h->d_team_trace && h->launches % 64 == 0)
        fprintf(stderr, "[dmhot] team launch: n_tiles %d grid %d teams %d maxt %d nin %d table_smem %d smem %zu\n", n_tiles, grid, T, maxt,
                nin, (int)table_smem, smem);
      TeamLaunch tl{grid, T, maxt, smem, s, h->pdl != 0, a->terms != nullptr, a->accumulate_stats != 0, table_smem};
      timing_begin(h, s);
      cudaError_t e = launch_team<KIND, Rb>(tl, q, ex);
      CU(h, e);
      h->launches++;
    } else if (p.aos && !team && h->variant != 0 && aligned && a->n >= TC::TE) {
      tiled = a->n / TC::TE * TC::TE;
      timing_begin(h, s);
    } else if (!p.aos && h->variant != 0 && aligned && a->ld % 4 == 0 && a->n >= 128) {
      // K1s: structure-of-arrays batches, tiles of 128 environments as one 512-byte bulk copy per row (dm_soa_kernel.cuh)
      using SC = SoaCfg<KIND, Rb>;
      tiled = a->n / SC::TE * SC::TE;
      soa_tiled = true;
      const int n_tiles = tiled / SC::TE;
      const size_t table_bytes = sizeof(float) * (size_t)h->total_rows * RowLayout<Rb>::STRIDE;
      const size_t acc_bytes = a->accumulate_stats ? SC::ACC_BYTES : 0;
      const size_t avail = SC::SMEM_BUDGET - SC::FIXED_BYTES;
      int G = SoaGroups<KIND, Rb>::G;
      while (G > 1 && G * SC::SLOT_BYTES + acc_bytes > avail) --G;
      const int want = (n_tiles + h->num_sms - 1) / h->num_sms;
      G = want < G ? (want < 1 ? 1 : want) : G;
      const size_t table_pad = (table_bytes + 127) / 128 * 128;  // tensor-map copies want 128-byte aligned shared memory
      bool table_smem = table_pad + G * SC::SLOT_BYTES + acc_bytes <= avail;
      if (h->table_smem_override >= 0) table_smem = table_smem && h->table_smem_override != 0;
      const size_t smem = SC::FIXED_BYTES + (table_smem ? table_pad : 0) + G * SC::SLOT_BYTES + acc_bytes;
      const int grid = (n_tiles + G - 1) / G < h->num_sms ? (n_tiles + G - 1) / G : h->num_sms;
      StepParams q = p;
      q.n = tiled;
      SoaExtra ex{a->accumulate_stats ? h->d_stats : nullptr, n_tiles, (int)(table_bytes / sizeof(float)), (int)(table_pad / sizeof(float)),
                  getenv("DM_PREFETCH_NEXT") ? atoi(getenv("DM_PREFETCH_NEXT")) : 1};
      const SoaMaps* maps = nullptr;
      for (const auto& m : h->soa_maps)
        if (m.q == a->qpos && m.v == a->qvel && m.o == a->obs && m.ld == a->ld && m.n == tiled) { maps = &m.maps; break; }
      if (!maps) {
        if (h->soa_maps.size() >= 128) h->soa_maps.clear();
        h->soa_maps.reserve(128);  // entries never move while they are referenced below
        dm_handle::SoaMapEntry m{a->qpos, a->qvel, a->obs, a->ld, tiled, {}};
        int rc2 = soa_encode(h, &m.maps.q, a->qpos, Rb::NQ, a->ld, tiled, Rb::NQ);
        if (!rc2) rc2 = soa_encode(h, &m.maps.v, a->qvel, Rb::NV, a->ld, tiled, Rb::NV);
        if (!rc2) rc2 = soa_encode(h, &m.maps.oj, a->obs, OL::DIM, a->ld, tiled, Rb::NJ);
        if (!rc2) rc2 = soa_encode(h, &m.maps.oc, a->obs, OL::DIM, a->ld, tiled, SC::NCOMP);
        if (rc2) return rc2;
        h->soa_maps.push_back(m);
        maps = &h->soa_maps.back().maps;
      }
      SoaLaunch sl{grid, G, smem, s, h->pdl != 0, a->terms != nullptr, a->accumulate_stats != 0, table_smem, maps};
      timing_begin(h, s);
      cudaError_t e = launch_soa<KIND, Rb>(sl, q, ex);
      CU(h, e);
      h->launches++;
    } else {
      timing_begin(h, s);
    }
    if (tiled && !team && !soa_tiled) {
      StepParams q = p;
      q.n = tiled;
      constexpr int GMAX = KernelGroups<Rb>::G;
      const int n_tiles = tiled / TC::TE;
      const size_t table_bytes = sizeof(float) * (size_t)h->total_rows * RowLayout<Rb>::STRIDE;
      // shared memory: fixed block, (optionally) the clip table, a ring of nbuf tile buffers, per-group running sums.
      // Groups first (occupancy), then any extra ring slots, then the clip table if it still fits.
      const size_t acc_bytes = a->accumulate_stats ? TC::ACC_BYTES : 0;
      const size_t avail = TC::SMEM_BUDGET - TC::FIXED_BYTES;
      int G = (int)(avail / (TC::BUF_BYTES + acc_bytes));
      G = G > GMAX ? GMAX : G;
      if (h->max_groups_override > 0 && h->max_groups_override < G) G = h->max_groups_override;
      const int want = (n_tiles + h->num_sms - 1) / h->num_sms;  // spread small batches over all SMs first
      G = want < G ? (want < 1 ? 1 : want) : G;
      // Batches of a few tiles per SM cannot fill GMAX groups anyway: they run the instantiation compiled for two groups
      // per CTA, whose threads get the registers they want (161-168 humanoid3d, 233-246 unitree_g1, no spills) instead of
      // 128 / 168.  Measured (kernel time per step): humanoid3d 32768 envs 16.4 -> 11.4 us, 16384 envs 11.8 -> 11.2 us;
      // G1 combined 65536 envs 33.2 -> 29.1 us; from ~1000 tiles on the occupancy of GMAX groups wins again.
      const bool wide = !a->terms && h->wide_override != 0 && (want <= KernelGroups<Rb>::WIDE_WANT || h->wide_override == 2);
      if (wide && G > KernelGroups<Rb>::GWIDE) G = KernelGroups<Rb>::GWIDE;
      // ... and some robots have a middle shape (KernelGroups::GMID groups without a register cap plus a lookahead slot)
      using KG = KernelGroups<Rb>;
      static const int mid_hi_env = getenv("DM_MID_HI") ? atoi(getenv("DM_MID_HI")) : 0;
      const int mid_hi = mid_hi_env > 0 ? mid_hi_env : KG::MID_WANT_HI;
      const bool mid = !wide && !a->terms && KG::GMID > 0 && h->wide_override != 0 && want >= KG::MID_WANT_LO && want <= mid_hi &&
                       G >= KG::GMID && h->max_groups_override == 0 && h->nbuf_override == 0 &&
                       table_bytes + KG::MID_NBUF * TC::BUF_BYTES + KG::GMID * acc_bytes <= avail;  // only with the clip table in shared memory
      if (mid) G = KG::GMID;
      const int grid = (n_tiles + G - 1) / G < h->num_sms ? (n_tiles + G - 1) / G : h->num_sms;
      const int tiles_per_cta = (n_tiles + grid - 1) / grid;
      // ring slots: measured at 1 M / 4 M envs, slots beyond the groups (tiles loading while every group computes)
      // are 3-5 % SLOWER than one slot per group, so the default ring is exactly the groups; DM_NBUF overrides
      int nbuf_want = h->nbuf_override > 0 ? h->nbuf_override : (mid ? KG::MID_NBUF : G);
      nbuf_want = nbuf_want > TC::MAXBUF ? TC::MAXBUF : nbuf_want;
      nbuf_want = nbuf_want > tiles_per_cta ? tiles_per_cta : nbuf_want;  // no more slots than tiles
      nbuf_want = nbuf_want < G ? G : nbuf_want;
      auto fit = [&](bool table) {
        const size_t rest = avail - G * acc_bytes;
        if (table && table_bytes > rest) return 0;
        const int nb = (int)((rest - (table ? table_bytes : 0)) / TC::BUF_BYTES);
        return nb > nbuf_want ? nbuf_want : nb;
      };
      const int nb_table = fit(true), nb_plain = fit(false);
      bool table_smem = nb_table >= G && (nb_table >= nb_plain || nb_table >= G + 1);
      if (h->table_smem_override >= 0) table_smem = h->table_smem_override != 0 && nb_table >= G;
      const int nbuf = table_smem ? nb_table : nb_plain;
      size_t smem = TC::FIXED_BYTES + (table_smem ? table_bytes : 0) + (size_t)nbuf * TC::BUF_BYTES + G * acc_bytes;
      // a table that stays in global memory: in the wide shape there is room to stage every environment's own row a tile ahead
      const size_t row_stage_bytes = (size_t)G * TC::TE * RowLayout<Rb>::STRIDE * sizeof(float);
      const bool row_stage = wide && !table_smem && !h->skip_compute && h->row_stage_override != 0 && G <= 3 &&
                             smem + row_stage_bytes <= TC::SMEM_BUDGET;
      if (row_stage) smem += row_stage_bytes;
      TileExtra ex{a->accumulate_stats ? h->d_stats : nullptr, n_tiles, (int)(table_bytes / sizeof(float)), nbuf, h->skip_compute, grid, 0, 0};
      static const int prefetch_next = getenv("DM_PREFETCH_NEXT") ? atoi(getenv("DM_PREFETCH_NEXT")) : 1;
      ex.prefetch_next = prefetch_next;
      static const int no_prefetch = getenv("DM_NO_PREFETCH") ? atoi(getenv("DM_NO_PREFETCH")) : 0;
      ex.no_prefetch = no_prefetch;
      const bool fused = h->fuse_now && a->accumulate_stats;  // one more CTA: the statistics exchange (dm_peer.cuh)
      if (fused) {
        static const int dbg = getenv("DM_PEER_DEBUG") ? atoi(getenv("DM_PEER_DEBUG")) : 1;
        ex.peer = peer_params(h, dbg);
      }
      // the kernels themselves live in dm_tile_{h,g,c}.cu, one translation unit per (kind, robot)
      TileLaunch tl{grid + (fused ? 1 : 0), G, smem, s, h->pdl != 0, a->terms != nullptr, a->accumulate_stats != 0,
                    row_stage ? 2 : (table_smem ? 1 : 0), wide ? 1 : (mid ? 2 : 0)};
      cudaError_t e = launch_tile<KIND, Rb>(tl, q, ex);
      CU(h, e);
      h->launches++;
      if (fused) h->fused_done = true;
    }
    const int rest = a->n - tiled;
    if (rest) {
      StepParams q = p;
      q.n = rest;
      if (tiled && soa_tiled) {  // SoA: the tail starts `tiled` environments into every row
        q.qpos = off(p.qpos, tiled); q.qvel = off(p.qvel, tiled); q.action = off(p.action, tiled); q.contact = off(p.contact, tiled);
        q.idx = off(p.idx, tiled); q.ep_len = off(p.ep_len, tiled); q.clip = off(p.clip, tiled);
        q.obs = off(p.obs, tiled); q.reward = off(p.reward, tiled); q.done = off(p.done, tiled); q.reason = off(p.reason, tiled);
        q.torque = off(p.torque, tiled); q.terms = off(p.terms, tiled); q.ep_return = off(p.ep_return, tiled);
      } else if (tiled) {  // AoS
        q.qpos = off(p.qpos, (long long)tiled * Rb::NQ); q.qvel = off(p.qvel, (long long)tiled * Rb::NV);
        q.action = off(p.action, (long long)tiled * Rb::NA); q.contact = off(p.contact, tiled);
        q.idx = off(p.idx, tiled); q.ep_len = off(p.ep_len, tiled); q.clip = off(p.clip, tiled);
        q.obs = off(p.obs, (long long)tiled * OL::DIM); q.reward = off(p.reward, tiled); q.done = off(p.done, tiled);
        q.reason = off(p.reason, tiled); q.torque = off(p.torque, (long long)tiled * Rb::NU);
        q.terms = off(p.terms, (long long)tiled * 8); q.ep_return = off(p.ep_return, tiled);
      }
      const int block = 128, grid = (rest + block - 1) / block;
      if (a->terms) step_strided_kernel<KIND, Rb, true><<<grid, block, 0, s>>>(q);
      else step_strided_kernel<KIND, Rb, false><<<grid, block, 0, s>>>(q);
      h->launches++;
      CU(h, cudaGetLastError());
      if (a->accumulate_stats) {
        StatsParams sp{rest, h->dims.obs_dim, q.es, q.aos, q.obs, q.reward, q.done, q.ep_return, q.ep_len, h->d_stats};
        int sgrid = (rest + 2047) / 2048;
        sgrid = sgrid < 1 ? 1 : (sgrid > 1184 ? 1184 : sgrid);
        stats_kernel<<<sgrid, dim3(32, 8), 0, s>>>(sp);
        h->launches++;
        CU(h, cudaGetLastError());
      }
    }
    timing_end(h, s);
    return DM_OK;
  });
}

static int issue_exchange(dm_handle* h, cudaStream_t s);

extern "C" int dm_step(dm_handle* h, const dm_step_args* a) {
  int rc = check_step_args(h, a, "dm_step");
  if (rc) return rc;
  CU(h, cudaSetDevice(h->device));
  // auto mode of the collective: every auto_every-th step with statistics comes with a lagged exchange (dm_peer.cuh).
  // Launches that leave SMs free (at most one tile per SM) get it as a one-CTA kernel on the side stream, with no
  // dependency on this stream in either direction: a kernel that has written peer memory completes ~1.2 us late, and
  // the next step must not wait for that (2 x 4096 envs: 9.1 us per step against 9.9 with the exchange CTA, 8.6 alone).
  // Launches that fill the GPU carry it as one extra CTA of the tile kernel, where it is free (1 M envs: 169.2 vs 169.5 us)
  // and a second kernel would wait for an SM.  DM_PEER_SIDE = 0 / 1 forces one or the other.
  const bool due = h->auto_every > 0 && a->accumulate_stats && a->n > 0 && (h->auto_count + 1) % h->auto_every == 0;
  const char* side_env = due ? getenv("DM_PEER_SIDE") : nullptr;  // read per exchange: the tests switch routes between steps
  const bool side = side_env ? atoi(side_env) != 0 : (a->n + 127) / 128 <= h->num_sms;
  h->fuse_now = due && !side;
  h->fused_done = false;
  if (h->fuse_now && h->exch_where == 2) {  // exchanges share the sequence number: this one goes behind those on the side stream
    CU(h, cudaEventRecord(h->coll_event, h->coll_stream));
    CU(h, cudaStreamWaitEvent((cudaStream_t)a->stream, h->coll_event, 0));
  }
  rc = launch_step(h, a, (cudaStream_t)a->stream);
  h->fuse_now = false;
  if (rc) return rc;
  if (h->auto_every > 0 && a->accumulate_stats && a->n > 0) ++h->auto_count;
  if (due && h->fused_done) h->exch_where = 1;
  if (!due || h->fused_done) return DM_OK;
  // (also the launch shapes without an exchange CTA: SoA, tails, the team kernel)
  if (h->exch_where == 1) {  // ... and this one behind the exchange CTA of an earlier step
    CU(h, cudaEventRecord(h->coll_event, (cudaStream_t)a->stream));
    CU(h, cudaStreamWaitEvent(h->coll_stream, h->coll_event, 0));
  }
  h->exch_where = 2;
  stats_exchange_lagged_kernel<<<1, kPeerStride, 0, h->coll_stream>>>(peer_params(h, 1));
  CU(h, cudaGetLastError());
  h->launches++;
  return DM_OK;
}

static int launch_reset(dm_handle* h, const dm_reset_args* a, cudaStream_t s) {
  if (a->n == 0) return DM_OK;
  return dispatch(h, [&](auto rb, auto kind) -> int {
    using Rb = decltype(rb);
    constexpr int KIND = decltype(kind)::value;
    ResetParams p{};
    p.n = a->n;
    p.aos = a->layout == DM_LAYOUT_AOS;
    p.es = p.aos ? 1 : a->ld;
    p.mask = a->mask; p.frame = a->frame; p.coin = a->coin; p.contact = a->contact;
    p.seed = a->seed; p.counter = a->counter;
    p.qpos = a->qpos; p.qvel = a->qvel; p.idx = a->idx; p.ep_len = a->ep_len; p.clip = a->clip;
    p.obs = a->obs; p.ep_return = a->ep_return; p.terminal_obs = a->terminal_obs; p.state_rows = h->d_state_rows; p.cs = h->cs;
    p.cs.contact_from_fk = h->contact_mode;
    reset_kernel<KIND, Rb><<<(a->n + 127) / 128, 128, 0, s>>>(p);
    h->launches++;
    CU(h, cudaGetLastError());
    return DM_OK;
  });
}

static int check_reset_args(dm_handle* h, const dm_reset_args* a, const char* who) {
  if (!h || !a) return fail(h, DM_E_ARG, "%s: NULL argument", who);
  if (!h->d_rows) return fail(h, DM_E_STATE, "%s: no clips uploaded", who);
  if (a->n < 0 || !a->idx || !a->ep_len || !a->obs) return fail(h, DM_E_ARG, "%s: idx / ep_len / obs are required", who);
  if (a->layout == DM_LAYOUT_SOA && a->ld < a->n) return fail(h, DM_E_ARG, "%s: SoA ld < n", who);
  if (h->kind == DM_ENV_COMBINED && !a->clip) return fail(h, DM_E_ARG, "%s: DPCombinedEnv needs the clip buffer", who);
  return DM_OK;
}

extern "C" int dm_reset(dm_handle* h, const dm_reset_args* a) {
  int rc = check_reset_args(h, a, "dm_reset");
  if (rc) return rc;
  CU(h, cudaSetDevice(h->device));
  return launch_reset(h, a, (cudaStream_t)a->stream);
}

// ---- host-buffer variants ---------------------------------------------------------------------
static int ensure_stage(dm_handle* h, Stage& st, size_t bytes) {
  if (bytes <= st.bytes) return DM_OK;
  cudaFree(st.d_mem);
  if (st.h_pin) cudaFreeHost(st.h_pin);
  st.d_mem = st.h_pin = nullptr;
  st.bytes = 0;
  const size_t cap = bytes + 4096;
  CU(h, cudaMalloc(&st.d_mem, cap));
  CU(h, cudaMallocHost(&st.h_pin, cap));
  st.bytes = cap;
  return DM_OK;
}

struct Carver {  // lays buffers out back to back (256-byte aligned) in the pinned / device staging areas
  char* hbase;
  char* dbase;
  size_t off = 0;
  template <class T>
  void take(size_t count, T** hp, T** dp) {
    off = (off + 255) & ~size_t(255);
    if (hp) *hp = (T*)(hbase + off);
    if (dp) *dp = (T*)(dbase + off);
    off += count * sizeof(T);
  }
};

// Staging layout of one host-buffer step: [inputs | in-out state | outputs], so that the H2D copy
// (inputs + state) and the D2H copy (state + outputs) are ONE contiguous transfer each.
struct StepStage {
  float *hq, *dq, *hv, *dv, *ha, *da;            // qpos qvel action
  uint32_t *hc, *dc;                             // contact
  int32_t *hi, *di, *hl, *dl, *hk, *dk;          // idx ep_len clip   (in-out)
  float *her, *der;                              // ep_return         (in-out)
  float *ho, *dob, *hr, *dr, *ht, *dt, *hm, *dmt;  // obs reward torque terms
  uint8_t *hd, *dd, *hz, *dz;                    // done reason
  size_t state0, out0, end;
};

static size_t step_stage_bytes(const dm_dims_t& d, size_t n) {
  return n * (sizeof(float) * (d.nq + d.nv + d.na + d.obs_dim + 1 + d.nu + DM_NTERMS + 1) + 4 * 4 + 2) + 16 * 256;
}

static StepStage carve_step(const dm_handle* h, size_t n) {
  const dm_dims_t& d = h->dims;
  Carver c{(char*)h->step_stage.h_pin, (char*)h->step_stage.d_mem};
  StepStage s;
  c.take(n * d.nq, &s.hq, &s.dq); c.take(n * d.nv, &s.hv, &s.dv); c.take(n * d.na, &s.ha, &s.da); c.take(n, &s.hc, &s.dc);
  c.off = (c.off + 255) & ~size_t(255);
  s.state0 = c.off;
  c.take(n, &s.hi, &s.di); c.take(n, &s.hl, &s.dl); c.take(n, &s.hk, &s.dk); c.take(n, &s.her, &s.der);
  c.off = (c.off + 255) & ~size_t(255);
  s.out0 = c.off;
  c.take(n * d.obs_dim, &s.ho, &s.dob); c.take(n, &s.hr, &s.dr); c.take(n * d.nu, &s.ht, &s.dt);
  c.take(n * DM_NTERMS, &s.hm, &s.dmt); c.take(n, &s.hd, &s.dd); c.take(n, &s.hz, &s.dz);
  s.end = c.off;
  return s;
}

// Pinned host buffers for n environments, laid out exactly as dm_step_host stages them.  A caller that keeps its
// per-step arrays HERE (and passes these very pointers) gets zero staging copies: dm_step_host then issues one
// H2D transfer, the kernel, and one D2H transfer.  The pointers stay valid until dm_host_buffers / dm_step_host is
// called with a larger n, or the handle is destroyed.
extern "C" int dm_host_buffers(dm_handle* h, int32_t n, dm_step_args* out) {
  if (!h || !out || n <= 0) return fail(h, DM_E_ARG, "dm_host_buffers: bad argument");
  CU(h, cudaSetDevice(h->device));
  int rc = ensure_stage(h, h->step_stage, step_stage_bytes(h->dims, (size_t)n));
  if (rc) return rc;
  const StepStage s = carve_step(h, (size_t)n);
  memset(out, 0, sizeof *out);
  out->n = n; out->layout = DM_LAYOUT_AOS; out->ld = 0;
  out->qpos = s.hq; out->qvel = s.hv; out->action = s.ha; out->contact = s.hc;
  out->idx = s.hi; out->ep_len = s.hl; out->clip = s.hk; out->ep_return = s.her;
  out->obs = s.ho; out->reward = s.hr; out->torque = s.ht; out->terms = s.hm; out->done = s.hd; out->reason = s.hz;
  return DM_OK;
}

// The mapped form of the host step.  The pinned block of dm_host_buffers is device-accessible (unified addressing),
// so the step kernel works on it directly: its TMA bulk loads pull the qpos / qvel / action tiles across PCIe, its
// TMA bulk store and plain stores push observations, torques and per-env scalars back, both directions at once and
// with no copy-engine launches in between (measured on the B200 box, 4096 envs, everything in place: 74 us against
// 101 us for H2D + kernel + D2H; the results are bit-identical, it is the same kernel on different addresses).
// Caller arrays that are not the pinned slots are staged in CHUNKS of whole tiles: chunk c is copied into the block
// and its kernel launched (chunks alternate between two streams) while the host already copies chunk c + 1; the
// outputs of chunk c are copied out as soon as its event fires, while later chunks still run.
static int step_host_mapped(dm_handle* h, const dm_step_args* a, const StepStage& g, CopyPool* pool, int64_t* h2d_bytes,
                            int64_t* d2h_bytes) {
  const dm_dims_t& d = h->dims;
  const size_t n = a->n;
  if (!h->host_stream2) CU(h, cudaStreamCreateWithFlags(&h->host_stream2, cudaStreamNonBlocking));
  struct Arr { void* slot; const void* src; void* dst; size_t row; };  // src: caller input (or NULL), dst: caller output (or NULL)
  // A big caller array that is itself pinned host memory (cudaHostAlloc / cudaHostRegister: torch pin_memory(), DPVecEnv's
  // own output arrays) needs no staging either: the kernel works on it where it lies.
  float *bq = g.hq, *bv = g.hv, *ba = g.ha, *bo = g.ho, *bt = g.ht;
  auto direct = [&](const void* p, size_t bytes) -> float* {
    if (!p || bytes < 64 * 1024 || (reinterpret_cast<uintptr_t>(p) & 15u)) return nullptr;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
      cudaGetLastError();
      return nullptr;
    }
    return at.type == cudaMemoryTypeHost ? (float*)at.devicePointer : nullptr;
  };
  const void *cq = a->qpos, *cv = a->qvel, *ca = a->action;
  void *co = a->obs, *ct = a->torque;
  if (cq != g.hq) if (float* q = direct(cq, sizeof(float) * n * d.nq)) { bq = q; cq = q; }
  if (cv != g.hv) if (float* q = direct(cv, sizeof(float) * n * d.nv)) { bv = q; cv = q; }
  if (ca != g.ha) if (float* q = direct(ca, sizeof(float) * n * d.na)) { ba = q; ca = q; }
  if (co != g.ho) if (float* q = direct(co, sizeof(float) * n * d.obs_dim)) { bo = q; co = q; }
  if (ct && ct != g.ht) if (float* q = direct(ct, sizeof(float) * n * d.nu)) { bt = q; ct = q; }
  const Arr ins[] = {{bq, cq, nullptr, sizeof(float) * d.nq}, {bv, cv, nullptr, sizeof(float) * d.nv},
                     {ba, ca, nullptr, sizeof(float) * d.na}, {g.hc, a->contact, nullptr, 4},
                     {g.hi, a->idx, nullptr, 4}, {g.hl, a->ep_len, nullptr, 4}, {g.hk, a->clip, nullptr, 4},
                     {g.her, a->ep_return, nullptr, 4}};
  const Arr outs[] = {{bo, nullptr, co, sizeof(float) * d.obs_dim}, {g.hr, nullptr, a->reward, 4}, {g.hd, nullptr, a->done, 1},
                      {g.hz, nullptr, a->reason, 1}, {bt, nullptr, ct, sizeof(float) * d.nu},
                      {g.hm, nullptr, a->terms, sizeof(float) * DM_NTERMS}, {g.hi, nullptr, a->idx, 4}, {g.hl, nullptr, a->ep_len, 4},
                      {g.hk, nullptr, a->clip, 4}, {g.her, nullptr, a->ep_return, 4}};
  // An absent contact / clip array reads as zeros.  The slot is zero-filled when it may hold something else (it was
  // carved, or an earlier step put words there) and is then in place like any other array, so that a caller who has
  // everything else in the pinned block keeps the single-launch path.
  auto absent = [&](const void* src, void* slot, bool& dirty) {
    if (src) { dirty = true; return; }
    if (dirty) memset(slot, 0, 4 * n);
    dirty = false;
  };
  absent(a->contact, g.hc, h->contact_slot_dirty);
  absent(a->clip, g.hk, h->clip_slot_dirty);
  bool staged_any = false;
  for (const Arr& x : ins) staged_any |= x.src && x.src != x.slot;
  for (const Arr& x : outs) staged_any |= x.dst && x.dst != x.slot;
  // chunks: whole tiles, at most 16 per step, at least 2048 environments each; one chunk when nothing is staged
  constexpr size_t kMaxChunks = 16, kTile = 128;
  size_t chunk = n;
  if (staged_any) {
    chunk = h->host_chunk > 0 ? (size_t)h->host_chunk : (n / 4 < 2048 ? 2048 : n / 4);
    chunk = (chunk + kTile - 1) / kTile * kTile;
    while ((n + chunk - 1) / chunk > kMaxChunks) chunk += kTile;
  }
  const size_t n_chunks = (n + chunk - 1) / chunk;
  using clk = std::chrono::steady_clock;
  auto t_prev = clk::now();
  const auto t_start = t_prev;
  h->ht[4] += 1;
  const bool timeline = h->host_trace >= 2 && staged_any && h->ht[4] >= h->host_trace && h->ht[4] < h->host_trace + 3;
  auto lap = [&](int k) {
    if (!h->host_trace) return;
    const auto t = clk::now();
    h->ht[k] += std::chrono::duration<double, std::micro>(t - t_prev).count();
    if (timeline)
      fprintf(stderr, "[dmhot]   t=%7.1f us  %s done\n", std::chrono::duration<double, std::micro>(t - t_start).count(),
              k == 0 ? "copy-in" : k == 1 ? "launch" : k == 2 ? "wait" : "copy-out");
    t_prev = t;
  };
  size_t next_out = 0;  // chunks [0, next_out) have been copied out
  auto copy_out = [&](size_t c) {
    const size_t e0 = c * chunk, e1 = e0 + chunk < n ? e0 + chunk : n, m = e1 - e0;
    CopyPool::Seg segs[CopyPool::kMaxSegs];
    int ns = 0;
    for (const Arr& x : outs)
      if (x.dst && x.dst != x.slot) segs[ns++] = CopyPool::Seg{(char*)x.dst + e0 * x.row, (const char*)x.slot + e0 * x.row, m * x.row};
    pool->copy_many(segs, ns);
    lap(3);
  };
  for (size_t c = 0; c < n_chunks; ++c) {
    const size_t e0 = c * chunk, e1 = e0 + chunk < n ? e0 + chunk : n, m = e1 - e0;
    CopyPool::Seg segs[CopyPool::kMaxSegs];
    int ns = 0;
    for (const Arr& x : ins) {  // absent arrays: contact / clip slots are zero already, ep_return is not read at all
      if (x.src == x.slot || !x.src) continue;
      segs[ns++] = CopyPool::Seg{(char*)x.slot + e0 * x.row, (const char*)x.src + e0 * x.row, m * x.row};
    }
    pool->copy_many(segs, ns);
    lap(0);
    dm_step_args k = *a;
    k.n = (int32_t)m;
    k.qpos = bq + e0 * d.nq; k.qvel = bv + e0 * d.nv; k.action = ba + e0 * d.na; k.contact = g.hc + e0;
    k.idx = g.hi + e0; k.ep_len = g.hl + e0; k.clip = g.hk + e0;
    k.obs = bo + e0 * d.obs_dim; k.reward = g.hr + e0; k.done = g.hd + e0; k.reason = g.hz + e0;
    k.torque = a->torque ? bt + e0 * d.nu : nullptr;
    k.terms = a->terms ? g.hm + e0 * DM_NTERMS : nullptr;
    k.ep_return = a->ep_return ? g.her + e0 : nullptr;
    cudaStream_t s = (c & 1) ? h->host_stream2 : h->host_stream;
    int rc = launch_step(h, &k, s);
    if (rc) return rc;
    if (!h->chunk_event[c]) CU(h, cudaEventCreateWithFlags(&h->chunk_event[c], cudaEventDisableTiming));
    CU(h, cudaEventRecord(h->chunk_event[c], s));
    lap(1);
    // outputs of chunks that have finished meanwhile leave now, while the GPU works on the later ones
    while (c + 1 < n_chunks && next_out < c && cudaEventQuery(h->chunk_event[next_out]) == cudaSuccess) copy_out(next_out++);
  }
  for (; next_out < n_chunks; ++next_out) {
    CU(h, cudaEventSynchronize(h->chunk_event[next_out]));
    lap(2);
    copy_out(next_out);
  }
  if (h2d_bytes) *h2d_bytes = (int64_t)g.out0;            // bytes the kernel reads across PCIe: inputs + in-out state
  if (d2h_bytes) *d2h_bytes = (int64_t)(g.end - g.state0);  // bytes it writes back: in-out state + outputs
  return DM_OK;
}

extern "C" int dm_set_host_mode(dm_handle* h, int mode) {
  if (!h || (mode != DM_HOST_STAGED && mode != DM_HOST_MAPPED)) return fail(h, DM_E_ARG, "dm_set_host_mode: bad argument");
  h->host_mode = mode;
  return DM_OK;
}

extern "C" int dm_step_host(dm_handle* h, const dm_step_args* a, int64_t* h2d_bytes, int64_t* d2h_bytes) {
  int rc = check_step_args(h, a, "dm_step_host");
  if (rc) return rc;
  if (a->layout != DM_LAYOUT_AOS) return fail(h, DM_E_ARG, "dm_step_host: host buffers are row-major (AoS) only");
  CU(h, cudaSetDevice(h->device));
  const dm_dims_t& d = h->dims;
  const size_t n = a->n;
  if (n == 0) return DM_OK;
  rc = ensure_stage(h, h->step_stage, step_stage_bytes(d, n));
  if (rc) return rc;
  const StepStage g = carve_step(h, n);
  if (h->slots_n != n) {
    h->slots_n = n;
    h->contact_slot_dirty = h->clip_slot_dirty = true;
  }
  cudaStream_t s = h->host_stream;
  // a buffer that already IS its staging slot (dm_host_buffers) needs no copy
  const bool small = sizeof(float) * n * (d.nq + d.nv + d.na) < 128 * 1024;  // never reaches the pool's parallel threshold
  if (!h->pool || (h->pool_small && !small)) {
    int helpers = h->copy_threads;
    if (helpers < 0) {
      const unsigned hw = std::thread::hardware_concurrency();
      helpers = hw >= 8 ? 3 : hw >= 4 ? 1 : 0;  // measured: 7 helpers are slower than 3 (the per-job hand-shake grows with the threads)
    }
    if (small) helpers = 0;
    h->pool_small = small;
    h->pool.reset(new CopyPool(helpers > 8 ? 8 : helpers));
  }
  CopyPool* pool = h->pool.get();
  // batches of at least one tile run on the pinned block itself; smaller ones (the N = 1 shims) are latency-bound
  // and keep the copy engines
  if (h->host_mode == DM_HOST_MAPPED && n >= 128) return step_host_mapped(h, a, g, pool, h2d_bytes, d2h_bytes);
  auto in = [pool](void* slot, const void* src, size_t bytes) { if (src != slot) { if (src) pool->copy(slot, src, bytes); else memset(slot, 0, bytes); } };
  auto out = [pool](void* dst, const void* slot, size_t bytes) { if (dst && dst != slot) pool->copy(dst, slot, bytes); };
  // Inputs that are not already in their staging slot are copied there one array at a time, each followed at once
  // by its own H2D transfer, so the DMA of one array overlaps the host memcpy of the next; when every input is in
  // place (dm_host_buffers) the whole input block goes in ONE transfer.
  const size_t in_bytes = g.out0;  // inputs + in-out state
  const bool in_place = a->qpos == g.hq && a->qvel == g.hv && a->action == g.ha && (const void*)a->contact == (const void*)g.hc &&
                        a->idx == g.hi && a->ep_len == g.hl && (const void*)a->clip == (const void*)g.hk &&
                        (!a->ep_return || a->ep_return == g.her);
  if (in_place) {
    CU(h, cudaMemcpyAsync(h->step_stage.d_mem, h->step_stage.h_pin, in_bytes, cudaMemcpyHostToDevice, s));
  } else {
    char* hb = (char*)h->step_stage.h_pin;
    char* db = (char*)h->step_stage.d_mem;
    auto ship = [&](const void* first, const void* last_end) -> cudaError_t {  // [first, last_end) of the pinned block
      const size_t o0 = (const char*)first - hb, o1 = (const char*)last_end - hb;
      return cudaMemcpyAsync(db + o0, hb + o0, o1 - o0, cudaMemcpyHostToDevice, s);
    };
    in(g.hq, a->qpos, sizeof(float) * n * d.nq);
    CU(h, ship(g.hq, g.hq + n * d.nq));
    in(g.hv, a->qvel, sizeof(float) * n * d.nv);
    CU(h, ship(g.hv, g.hv + n * d.nv));
    in(g.ha, a->action, sizeof(float) * n * d.na);
    CU(h, ship(g.ha, g.ha + n * d.na));
    in(g.hc, a->contact, 4 * n);
    in(g.hi, a->idx, 4 * n);
    in(g.hl, a->ep_len, 4 * n);
    in(g.hk, a->clip, 4 * n);
    if (a->ep_return) in(g.her, a->ep_return, 4 * n);
    CU(h, ship(g.hc, hb + in_bytes));  // contact words + the in-out state block behind them
  }
  dm_step_args k = *a;
  k.qpos = g.dq; k.qvel = g.dv; k.action = g.da; k.contact = g.dc; k.idx = g.di; k.ep_len = g.dl; k.clip = g.dk;
  k.obs = g.dob; k.reward = g.dr; k.done = g.dd; k.reason = g.dz; k.torque = a->torque ? g.dt : nullptr;
  k.terms = a->terms ? g.dmt : nullptr; k.ep_return = a->ep_return ? g.der : nullptr;
  rc = launch_step(h, &k, s);
  if (rc) return rc;
  const size_t back_bytes = g.end - g.state0;  // in-out state + outputs
  char* hb2 = (char*)h->step_stage.h_pin;
  char* db2 = (char*)h->step_stage.d_mem;
  if (a->obs == g.ho) {  // outputs stay in the pinned block: one transfer
    CU(h, cudaMemcpyAsync(hb2 + g.state0, db2 + g.state0, back_bytes, cudaMemcpyDeviceToHost, s));
    CU(h, cudaStreamSynchronize(s));
  } else {  // two transfers: the caller's observation copy overlaps the DMA of the remaining outputs
    const size_t obs_end = (size_t)((char*)(g.ho + n * d.obs_dim) - hb2);
    if (!h->obs_event) CU(h, cudaEventCreateWithFlags(&h->obs_event, cudaEventDisableTiming));
    CU(h, cudaMemcpyAsync(hb2 + g.state0, db2 + g.state0, obs_end - g.state0, cudaMemcpyDeviceToHost, s));
    CU(h, cudaEventRecord(h->obs_event, s));
    CU(h, cudaMemcpyAsync(hb2 + obs_end, db2 + obs_end, g.end - obs_end, cudaMemcpyDeviceToHost, s));
    CU(h, cudaEventSynchronize(h->obs_event));
    pool->copy(a->obs, g.ho, sizeof(float) * n * d.obs_dim);
    CU(h, cudaStreamSynchronize(s));
  }
  out(a->reward, g.hr, 4 * n);
  out(a->done, g.hd, n);
  out(a->reason, g.hz, n);
  out(a->torque, g.ht, sizeof(float) * n * d.nu);
  out(a->terms, g.hm, sizeof(float) * n * DM_NTERMS);
  out(a->idx, g.hi, 4 * n);
  out(a->ep_len, g.hl, 4 * n);
  out(a->clip, g.hk, 4 * n);
  out(a->ep_return, g.her, 4 * n);
  if (h2d_bytes) *h2d_bytes = (int64_t)in_bytes;
  if (d2h_bytes) *d2h_bytes = (int64_t)back_bytes;
  return DM_OK;
}
// The mapped form of the host reset: the reset kernel runs on pinned host memory directly (the handle's block, caller
// arrays that are themselves pinned, or small pinned staging slots for the rest).  It reads the n mask bytes and touches
// only the rows of the environments being reset, so an auto-reset of a few finished episodes moves a few KB across PCIe
// instead of shipping every array of the batch to the device and back (4096 envs: 2 x 3.3 MB through the copy engines).
static int reset_host_mapped(dm_handle* h, const dm_reset_args* a) {
  const dm_dims_t& d = h->dims;
  const size_t n = a->n;
  int rc = ensure_stage(h, h->reset_stage, n * (sizeof(float) * (d.nq + d.nv + 2 * d.obs_dim + 1) + 4 * 6 + 1) + 16 * 256);
  if (rc) return rc;
  Carver c{(char*)h->reset_stage.h_pin, (char*)h->reset_stage.d_mem};
  const char* blk0 = (const char*)h->step_stage.h_pin;
  const char* blk1 = blk0 ? blk0 + h->step_stage.bytes : nullptr;
  auto pinned = [&](const void* p) -> void* {  // the device-visible address of a pinned host pointer, else NULL
    if (!p) return nullptr;
    if (blk0 && (const char*)p >= blk0 && (const char*)p < blk1) return const_cast<void*>(p);
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
      cudaGetLastError();
      return nullptr;
    }
    return at.type == cudaMemoryTypeHost ? at.devicePointer : nullptr;
  };
  struct Back { void* dst; const void* slot; size_t bytes; };
  Back back[8];
  int nback = 0;
  // in: what the kernel reads for this array; absent arrays read as `fill` (or stay NULL when optional)
  auto bind = [&](const void* caller, size_t bytes, bool out, bool optional, int fill) -> void* {
    char* slot;
    c.take(bytes, &slot, (char**)nullptr);
    if (!caller) {
      if (optional) return nullptr;
      memset(slot, fill, bytes);
      return slot;
    }
    if (void* p = pinned(caller)) return p;
    memcpy(slot, caller, bytes);
    if (out) back[nback++] = Back{const_cast<void*>(caller), slot, bytes};
    return slot;
  };
  dm_reset_args k = *a;
  k.mask = (const uint8_t*)bind(a->mask, n, false, false, 1);
  k.frame = (const int32_t*)bind(a->frame, 4 * n, false, true, 0);
  k.coin = (const int32_t*)bind(a->coin, 4 * n, false, true, 0);
  k.contact = (const uint32_t*)bind(a->contact, 4 * n, false, false, 0);
  k.idx = (int32_t*)bind(a->idx, 4 * n, true, false, 0);
  k.ep_len = (int32_t*)bind(a->ep_len, 4 * n, true, false, 0);
  k.clip = (int32_t*)bind(a->clip, 4 * n, true, false, 0);
  k.ep_return = (float*)bind(a->ep_return, 4 * n, true, false, 0);
  k.qpos = (float*)bind(a->qpos, sizeof(float) * n * d.nq, true, true, 0);
  k.qvel = (float*)bind(a->qvel, sizeof(float) * n * d.nv, true, true, 0);
  k.obs = (float*)bind(a->obs, sizeof(float) * n * d.obs_dim, true, false, 0);
  k.terminal_obs = (float*)bind(a->terminal_obs, sizeof(float) * n * d.obs_dim, true, true, 0);
  cudaStream_t s = h->host_stream;
  rc = launch_reset(h, &k, s);
  if (rc) return rc;
  CU(h, cudaStreamSynchronize(s));
  for (int i = 0; i < nback; ++i) memcpy(back[i].dst, back[i].slot, back[i].bytes);
  return DM_OK;
}

extern "C" int dm_reset_host(dm_handle* h, const dm_reset_args* a) {
  int rc = check_reset_args(h, a, "dm_reset_host");
  if (rc) return rc;
  if (a->layout != DM_LAYOUT_AOS) return fail(h, DM_E_ARG, "dm_reset_host: AoS only");
  CU(h, cudaSetDevice(h->device));
  const dm_dims_t& d = h->dims;
  const size_t n = a->n;
  if (n == 0) return DM_OK;
  if (h->host_mode == DM_HOST_MAPPED && n >= 128) return reset_host_mapped(h, a);
  rc = ensure_stage(h, h->reset_stage, n * (sizeof(float) * (d.nq + d.nv + 2 * d.obs_dim + 1) + 4 * 6 + 1) + 16 * 256);
  if (rc) return rc;
  Carver c{(char*)h->reset_stage.h_pin, (char*)h->reset_stage.d_mem};
  cudaStream_t s = h->host_stream;
  uint8_t *hm, *dmk;
  int32_t *hf, *df, *hco, *dco, *hi, *di, *hl, *dl, *hk, *dk;
  uint32_t *hc, *dc;
  float *hq, *dq, *hv, *dv, *ho, *dob, *her, *der, *hto, *dto;
  c.take(n, &hm, &dmk); c.take(n, &hf, &df); c.take(n, &hco, &dco); c.take(n, &hc, &dc);
  c.take(n, &hi, &di); c.take(n, &hl, &dl); c.take(n, &hk, &dk); c.take(n, &her, &der);
  c.take(n * d.nq, &hq, &dq); c.take(n * d.nv, &hv, &dv); c.take(n * d.obs_dim, &ho, &dob);
  c.take(n * d.obs_dim, &hto, &dto);
  if (a->terminal_obs) memcpy(hto, a->terminal_obs, sizeof(float) * n * d.obs_dim);
  if (a->mask) memcpy(hm, a->mask, n); else memset(hm, 1, n);
  if (a->frame) memcpy(hf, a->frame, 4 * n);
  if (a->coin) memcpy(hco, a->coin, 4 * n);
  if (a->contact) memcpy(hc, a->contact, 4 * n); else memset(hc, 0, 4 * n);
  memcpy(hi, a->idx, 4 * n);
  memcpy(hl, a->ep_len, 4 * n);
  if (a->clip) memcpy(hk, a->clip, 4 * n); else memset(hk, 0, 4 * n);
  if (a->ep_return) memcpy(her, a->ep_return, 4 * n); else memset(her, 0, 4 * n);
  if (a->qpos) memcpy(hq, a->qpos, sizeof(float) * n * d.nq);
  if (a->qvel) memcpy(hv, a->qvel, sizeof(float) * n * d.nv);
  memcpy(ho, a->obs, sizeof(float) * n * d.obs_dim);
  CU(h, cudaMemcpyAsync(h->reset_stage.d_mem, h->reset_stage.h_pin, c.off, cudaMemcpyHostToDevice, s));
  dm_reset_args k = *a;
  k.mask = dmk; k.frame = a->frame ? df : nullptr; k.coin = a->coin ? dco : nullptr; k.contact = dc;
  k.idx = di; k.ep_len = dl; k.clip = dk; k.ep_return = der; k.qpos = dq; k.qvel = dv; k.obs = dob; k.terminal_obs = a->terminal_obs ? dto : nullptr;
  rc = launch_reset(h, &k, s);
  if (rc) return rc;
  CU(h, cudaMemcpyAsync(h->reset_stage.h_pin, h->reset_stage.d_mem, c.off, cudaMemcpyDeviceToHost, s));
  CU(h, cudaStreamSynchronize(s));
  memcpy(a->idx, hi, 4 * n);
  memcpy(a->ep_len, hl, 4 * n);
  if (a->clip) memcpy(a->clip, hk, 4 * n);
  if (a->ep_return) memcpy(a->ep_return, her, 4 * n);
  if (a->qpos) memcpy(a->qpos, hq, sizeof(float) * n * d.nq);
  if (a->qvel) memcpy(a->qvel, hv, sizeof(float) * n * d.nv);
  memcpy(a->obs, ho, sizeof(float) * n * d.obs_dim);
  if (a->terminal_obs) memcpy(a->terminal_obs, hto, sizeof(float) * n * d.obs_dim);
  return DM_OK;
}

// ---- statistics, bookkeeping ------------------------------------------------------------------
extern "C" int dm_stats_zero(dm_handle* h, void* stream) {
  if (!h) return fail(nullptr, DM_E_ARG, "dm_stats_zero: NULL handle");
  CU(h, cudaSetDevice(h->device));
  CU(h, cudaMemsetAsync(h->d_stats, 0, sizeof(double) * h->dims.n_stats, (cudaStream_t)stream));
  if (h->coll_stream) CU(h, cudaStreamSynchronize(h->coll_stream));
  if (h->d_total) {  // the collective's snapshot and global totals restart with the accumulators (not its sequence number)
    CU(h, cudaMemsetAsync(h->d_snap, 0, sizeof(double) * kPeerStride, (cudaStream_t)stream));
    CU(h, cudaMemsetAsync(h->d_total, 0, sizeof(double) * kPeerStride, (cudaStream_t)stream));
  }
  return DM_OK;
}
extern "C" int dm_stats_device_ptr(dm_handle* h, double** dev_ptr, int32_t* n) {
  if (!h || !dev_ptr || !n) return fail(h, DM_E_ARG, "dm_stats_device_ptr: NULL argument");
  *dev_ptr = h->d_stats;
  *n = h->dims.n_stats;
  return DM_OK;
}
extern "C" int dm_stats_read(dm_handle* h, double* host_out, void* stream) {
  if (!h || !host_out) return fail(h, DM_E_ARG, "dm_stats_read: NULL argument");
  CU(h, cudaSetDevice(h->device));
  CU(h, cudaMemcpyAsync(host_out, h->d_stats, sizeof(double) * h->dims.n_stats, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CU(h, cudaStreamSynchronize((cudaStream_t)stream));
  return DM_OK;
}
extern "C" int dm_launch_count(const dm_handle* h, int64_t* kernels) {
  if (!h || !kernels) return fail(nullptr, DM_E_ARG, "dm_launch_count: NULL argument");
  *kernels = h->launches;
  return DM_OK;
}
extern "C" int dm_set_timing(dm_handle* h, int enabled) {
  if (!h) return fail(nullptr, DM_E_ARG, "dm_set_timing: NULL handle");
  h->timing = enabled != 0;
  return DM_OK;
}
extern "C" int dm_step_kernel_time(dm_handle* h, int reset, double* avg_ms, int64_t* launches) {
  if (!h) return fail(nullptr, DM_E_ARG, "dm_step_kernel_time: NULL handle");
  CU(h, cudaSetDevice(h->device));
  for (size_t i = 0; i < h->ev_used; ++i) {
    CU(h, cudaEventSynchronize(h->ev[i].second));
    float ms = 0;
    CU(h, cudaEventElapsedTime(&ms, h->ev[i].first, h->ev[i].second));
    h->t_ms += ms;
    h->t_n++;
  }
  h->ev_used = 0;
  if (avg_ms) *avg_ms = h->t_n ? h->t_ms / h->t_n : 0.0;
  if (launches) *launches = h->t_n;
  if (reset) { h->t_ms = 0; h->t_n = 0; }
  return DM_OK;
}
extern "C" int dm_set_contact_mode(dm_handle* h, int mode) {
  if (!h || (mode != DM_CONTACT_CALLER && mode != DM_CONTACT_FROM_FK)) return fail(h, DM_E_ARG, "dm_set_contact_mode: mode must be 0 or 1");
  h->contact_mode = mode;
  return DM_OK;
}
extern "C" int dm_set_kernel_variant(dm_handle* h, int variant) {
  if (!h || variant < -1 || variant > 2) return fail(h, DM_E_ARG, "dm_set_kernel_variant: variant must be -1, 0, 1 or 2");
  h->variant = variant;
  return DM_OK;
}

// ---- policy (K6) -------------------------------------------------------------------------------------------------------
extern "C" int dm_policy_upload(dm_handle* h, const dm_policy_t* pol) {
  if (!h || !pol) return fail(h, DM_E_ARG, "dm_policy_upload: NULL argument");
  if (!pol->W0 || !pol->b0 || !pol->W1 || !pol->b1 || !pol->W2 || !pol->b2) return fail(h, DM_E_ARG, "dm_policy_upload: a weight pointer is NULL");
  if (pol->in_dim < 1 || pol->in_dim > PMAX_IN || pol->h0 < 1 || pol->h0 > PMAX_H0 || pol->h1 < 1 || pol->h1 > PMAX_H1 ||
      pol->out_dim < 1 || pol->out_dim > PMAX_OUT)
    return fail(h, DM_E_ARG, "dm_policy_upload: supported sizes are in <= %d, hidden <= [%d, %d], out <= %d (got %d -> %d -> %d -> %d)",
                PMAX_IN, PMAX_H0, PMAX_H1, PMAX_OUT, pol->in_dim, pol->h0, pol->h1, pol->out_dim);
  CU(h, cudaSetDevice(h->device));
  const int in = pol->in_dim, in_pad = (in + PK - 1) / PK * PK, h0 = pol->h0, h0p = (h0 + 127) / 128 * 128, h1 = pol->h1, h1p = 128,
            out = pol->out_dim;
  const int big = (h0p / 128) * (in_pad / PK) + h0p / PK, small = h1p / PK;
  std::vector<float> packed((size_t)big * PCHUNK + (size_t)small * PK * PHEADC, 0.0f), bias(h0p + h1p + PHEADC, 0.0f);
  float* w = packed.data();
  for (int n0 = 0; n0 < h0p; n0 += 128)        // layer 0, one 128-column pass at a time
    for (int k0 = 0; k0 < in_pad; k0 += PK, w += PCHUNK)
      for (int kk = 0; kk < PK; ++kk)
        for (int c = 0; c < 128; ++c)
          if (k0 + kk < in && n0 + c < h0) w[kk * 128 + c] = pol->W0[(size_t)(k0 + kk) * h0 + n0 + c];
  for (int k0 = 0; k0 < h0p; k0 += PK, w += PCHUNK)  // layer 1
    for (int kk = 0; kk < PK; ++kk)
      for (int c = 0; c < 128; ++c)
        if (k0 + kk < h0 && c < h1) w[kk * 128 + c] = pol->W1[(size_t)(k0 + kk) * h1 + c];
  for (int k0 = 0; k0 < h1p; k0 += PK, w += PK * PHEADC)  // head
    for (int kk = 0; kk < PK; ++kk)
      for (int c = 0; c < PHEADC; ++c)
        if (k0 + kk < h1 && c < out) w[kk * PHEADC + c] = pol->W2[(size_t)(k0 + kk) * out + c];
  // the weights for the tensor-core kernel: layer 0 (one 128-column half after the other) | layer 1 | head; per k-step of
  // 16 [B_hi | B_lo], each [2 k-chunks][N][8] fp16 (K-major core matrices, see dm_policy_tc.cuh), N = 128 (head: PHEADC)
  std::vector<__half> p16((size_t)((h0p / 128) * (in_pad / 16) + h0p / 16) * 4096 + (size_t)8 * PHEADC * 32, __float2half_rn(0.0f));
  {
    __half* o = p16.data();
    auto emit = [&](int n_cols, auto&& weight /* (k, n) -> float, 0 outside */, int k0, int n0) {
      for (int part = 0; part < 2; ++part)  // hi, lo
        for (int kc = 0; kc < 2; ++kc)
          for (int n = 0; n < n_cols; ++n)
            for (int k8 = 0; k8 < 8; ++k8) {
              const float x = weight(k0 + kc * 8 + k8, n0 + n);
              const __half hi = __float2half_rn(x);
              *o++ = part == 0 ? hi : __float2half_rn(x - __half2float(hi));
            }
    };
    auto w0 = [&](int k, int n) { return k < in && n < h0 ? pol->W0[(size_t)k * h0 + n] : 0.0f; };
    auto w1 = [&](int k, int n) { return k < h0 && n < h1 ? pol->W1[(size_t)k * h1 + n] : 0.0f; };
    auto w2 = [&](int k, int n) { return k < h1 && n < out ? pol->W2[(size_t)k * out + n] : 0.0f; };
    for (int hf = 0; hf < h0p / 128; ++hf)
      for (int k0 = 0; k0 < in_pad; k0 += 16) emit(128, w0, k0, 128 * hf);
    for (int k0 = 0; k0 < h0p; k0 += 16) emit(128, w1, k0, 0);
    for (int k0 = 0; k0 < 128; k0 += 16) emit(PHEADC, w2, k0, 0);
  }
  for (int i = 0; i < h0; ++i) bias[i] = pol->b0[i];
  for (int i = 0; i < h1; ++i) bias[h0p + i] = pol->b1[i];
  for (int i = 0; i < out; ++i) bias[h0p + h1p + i] = pol->b2[i];
  cudaFree(h->d_pol_packed);
  cudaFree(h->d_pol_bias);
  h->d_pol_packed = h->d_pol_bias = nullptr;
  cudaFree(h->d_pol_packed16);
  h->d_pol_packed16 = nullptr;
  CU(h, cudaMalloc(&h->d_pol_packed16, p16.size() * sizeof(__half)));
  CU(h, cudaMemcpy(h->d_pol_packed16, p16.data(), p16.size() * sizeof(__half), cudaMemcpyHostToDevice));
  CU(h, cudaMalloc(&h->d_pol_packed, packed.size() * sizeof(float)));
  CU(h, cudaMalloc(&h->d_pol_bias, bias.size() * sizeof(float)));
  CU(h, cudaMemcpy(h->d_pol_packed, packed.data(), packed.size() * sizeof(float), cudaMemcpyHostToDevice));
  CU(h, cudaMemcpy(h->d_pol_bias, bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice));
  if (!h->d_pol_norm) {
    std::vector<float> id(2 * PMAX_IN + 1, 0.0f);
    for (int i = PMAX_IN; i <= 2 * PMAX_IN; ++i) id[i] = 1.0f;
    CU(h, cudaMalloc(&h->d_pol_norm, id.size() * sizeof(float)));
    CU(h, cudaMemcpy(h->d_pol_norm, id.data(), id.size() * sizeof(float), cudaMemcpyHostToDevice));
  }
  PolicyParams& q = h->pol;
  q = PolicyParams{};
  q.in = in; q.in_pad = in_pad; q.h0 = h0p; q.h1 = h1p; q.out = out;
  q.big_chunks = big; q.chunks_per_tile = big + small;
  return DM_OK;
}

extern "C" int dm_policy_set_normalizer(dm_handle* h, int32_t dim, const float* mean, const float* var, float eps, float clip_obs,
                                        float ret_var, float clip_reward) {
  if (!h) return fail(h, DM_E_ARG, "dm_policy_set_normalizer: NULL handle");
  if (!h->d_pol_norm) return fail(h, DM_E_STATE, "dm_policy_set_normalizer: no policy uploaded (call dm_policy_upload first)");
  if (dim < 0 || dim > PMAX_IN || (mean && !var)) return fail(h, DM_E_ARG, "dm_policy_set_normalizer: dim %d, mean / var inconsistent", dim);
  CU(h, cudaSetDevice(h->device));
  std::vector<float> v(2 * PMAX_IN + 1, 0.0f);
  for (int i = 0; i < PMAX_IN; ++i) {
    v[i] = mean && i < dim ? mean[i] : 0.0f;
    v[PMAX_IN + i] = mean && i < dim ? (float)(1.0 / sqrt((double)var[i] + (double)eps)) : 1.0f;
  }
  v[2 * PMAX_IN] = (float)(1.0 / sqrt((double)ret_var + (double)eps));
  CU(h, cudaMemcpy(h->d_pol_norm, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice));
  h->pol_clip_obs = clip_obs;
  h->pol_clip_r = clip_reward;
  return DM_OK;
}

extern "C" int dm_policy_normalizer_from_stats(dm_handle* h, float eps, float clip_obs, float clip_reward, void* stream) {
  if (!h) return fail(h, DM_E_ARG, "dm_policy_normalizer_from_stats: NULL handle");
  if (!h->d_pol_norm) return fail(h, DM_E_STATE, "dm_policy_normalizer_from_stats: no policy uploaded (call dm_policy_upload first)");
  if (h->dims.obs_dim > PMAX_IN) return fail(h, DM_E_ARG, "dm_policy_normalizer_from_stats: obs_dim %d > %d", h->dims.obs_dim, PMAX_IN);
  CU(h, cudaSetDevice(h->device));
  policy_norm_from_stats_kernel<<<1, PMAX_IN, 0, (cudaStream_t)stream>>>(h->d_total ? h->d_total : h->d_stats, h->dims.obs_dim, eps, h->d_pol_norm,
                                                                        h->d_pol_norm + 2 * PMAX_IN);
  h->launches++;
  CU(h, cudaGetLastError());
  h->pol_clip_obs = clip_obs;
  h->pol_clip_r = clip_reward;
  return DM_OK;
}

// debug hook for tools/policy_trace.py (not part of the public header): per-tile event clocks of CTA 0
extern "C" int dm_policy_set_trace(dm_handle* h, long long* device_buffer, int cta) {
  if (!h) return DM_E_ARG;
  h->pol_trace = device_buffer;
  h->pol_trace_cta = cta;
  return DM_OK;
}

extern "C" int dm_set_policy_variant(dm_handle* h, int variant) {
  if (!h || variant < 0 || variant > 1) return fail(h, DM_E_ARG, "dm_set_policy_variant: variant must be 0 (fp32 CUDA cores) or 1 (tensor cores)");
  h->pol_variant = variant;
  return DM_OK;
}

extern "C" int dm_policy_act(dm_handle* h, const dm_policy_args* a) {
  if (!h || !a) return fail(h, DM_E_ARG, "dm_policy_act: NULL argument");
  if (!h->d_pol_packed) return fail(h, DM_E_STATE, "dm_policy_act: no policy uploaded (call dm_policy_upload first)");
  if (a->n < 0) return fail(h, DM_E_ARG, "dm_policy_act: n = %d", a->n);
  if (a->n == 0) return DM_OK;
  if (!a->obs || !a->action) return fail(h, DM_E_ARG, "dm_policy_act: obs / action is NULL");
  if ((a->reward_norm != nullptr) != (a->reward != nullptr)) return fail(h, DM_E_ARG, "dm_policy_act: reward and reward_norm go together");
  CU(h, cudaSetDevice(h->device));
  if (!h->pol_smem_set) {
    CU(h, cudaFuncSetAttribute(policy_mlp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPolicySmemBytes));
    h->pol_smem_set = true;
  }
  PolicyParams p = h->pol;
  p.n = a->n;
  p.n_tiles = (a->n + PM - 1) / PM;
  p.obs = a->obs; p.action = a->action; p.obs_norm = a->obs_norm; p.reward = a->reward; p.reward_norm = a->reward_norm;
  p.packed = h->d_pol_packed; p.bias = h->d_pol_bias; p.norm = h->d_pol_norm;
  p.clip_obs = h->pol_clip_obs; p.clip_r = h->pol_clip_r; p.act_lo = a->act_lo; p.act_hi = a->act_hi;
  p.normalize = a->normalize != 0;
  p.use_tma = (reinterpret_cast<uintptr_t>(a->obs) & 15u) == 0;
  const int grid = p.n_tiles < h->num_sms ? p.n_tiles : h->num_sms;
  if (h->pol_variant == 1) {
    if (!h->pol_tc_smem_set) {
      CU(h, cudaFuncSetAttribute(policy_mlp_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
      h->pol_tc_smem_set = true;
    }
    PolicyTcParams q{p, h->d_pol_packed16, policy_tc_stages(p.in), policy_tc_raw_bytes(p.in), h->pol_trace, h->pol_trace_cta};
    policy_mlp_tc_kernel<<<grid, TC_THREADS, policy_tc_smem_bytes(p.in), (cudaStream_t)a->stream>>>(q);
    h->launches++;
    CU(h, cudaGetLastError());
    return DM_OK;
  }
  policy_mlp_kernel<<<grid, 256, kPolicySmemBytes, (cudaStream_t)a->stream>>>(p);
  h->launches++;
  CU(h, cudaGetLastError());
  return DM_OK;
}

// ---- dm_batch_stats_noalloc: the statistics reduction (K4) without any allocation ----------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The existing route to these sums --
// dm_step(accumulate_stats = 1) + dm_stats_read -- keeps its accumulators in memory cudaMalloc'd per handle; this variant
// of stats_kernel keeps its scratch (per-block partial sums) in a file-scope device array instead and calls no cudaMalloc,
// cudaFree or allocating library routine.
//   * Size: kNoAllocBlocks x kNoAllocRow doubles = 256 KiB, held by every GPU that loads the library until exit.
//   * Limits: obs_dim <= DM_NOALLOC_MAX_OBS_DIM (6 + 2 obs_dim <= kNoAllocRow), n <= DM_NOALLOC_MAX_ENVS.
//   * Every call shares the array and it keeps the last call's contents.  Nothing in it has to start at zero: stage 1
//     WRITES (never adds to) the rows of the blocks it launches, stage 2 reads exactly those rows, so stale rows of an
//     earlier, larger call are never read.  Calls that overlap on the GPU (two streams, two host threads) RACE on the
//     array: serialise them (one stream, or an event between them).
//   * In host code `g_noalloc_scratch` is a host-side stand-in, not a device address; the address the kernels get comes
//     from cudaGetSymbolAddress, once per device with that device current.
constexpr int kNoAllocRow = 256;     // doubles per block: [6 scalars | sum obs[D] | sum obs^2[D]], D <= 125
constexpr int kNoAllocBlocks = 128;  // stage-1 blocks at most
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocBlocks * kNoAllocRow];
static double* g_noalloc_ptr[64] = {};  // per device: device address of g_noalloc_scratch
static std::mutex g_noalloc_mutex;

// stage 1: block b reduces its slab of environments into scratch row b (blockDim = (32, 8); fixed summation order)
static __global__ void __launch_bounds__(256) stats_partial_kernel(int n, int D, const float* __restrict__ obs,
                                                                   const float* __restrict__ reward, const uint8_t* __restrict__ done,
                                                                   const float* __restrict__ ep_return, const int* __restrict__ ep_len,
                                                                   double* __restrict__ scratch) {
  __shared__ double sh[8][64];
  const int lane = threadIdx.x, wy = threadIdx.y;
  const int per_block = (n + gridDim.x - 1) / gridDim.x;
  const int e0 = blockIdx.x * per_block, e1 = min(n, e0 + per_block);
  double* row = scratch + (size_t)blockIdx.x * kNoAllocRow;
  if (wy == 0) {  // the six scalars: lanes walk environments
    double c = 0, sr = 0, srr = 0, nd = 0, ser = 0, sel = 0;
    for (int e = e0 + lane; e < e1; e += 32) {
      const double r = reward[e];
      c += 1.0; sr += r; srr += r * r;
      if (done[e]) {
        nd += 1.0;
        if (ep_return) ser += ep_return[e];
        if (ep_len) sel += ep_len[e];
      }
    }
    c = warp_sum(c); sr = warp_sum(sr); srr = warp_sum(srr); nd = warp_sum(nd); ser = warp_sum(ser); sel = warp_sum(sel);
    if (lane == 0) { row[0] = c; row[1] = sr; row[2] = srr; row[3] = nd; row[4] = ser; row[5] = sel; }
  }
  for (int j0 = 0; j0 < D; j0 += 32) {  // observation columns: lanes walk columns of a row (coalesced), warps walk rows
    const int j = j0 + lane;
    double s = 0, ss = 0;
    if (j < D)
      for (int e = e0 + wy; e < e1; e += 8) {
        const double v = obs[(long long)e * D + j];
        s += v; ss += v * v;
      }
    sh[wy][lane] = s;
    sh[wy][32 + lane] = ss;
    __syncthreads();
    if (wy == 0 && j < D) {
      double a = 0, b = 0;
#pragma unroll
      for (int w = 0; w < 8; ++w) { a += sh[w][lane]; b += sh[w][32 + lane]; }
      row[6 + j] = a;
      row[6 + D + j] = b;
    }
    __syncthreads();
  }
}
// stage 2: out[j] = sum over the blocks of stage 1, in block order
static __global__ void __launch_bounds__(kNoAllocRow) stats_final_kernel(int blocks, int n_stats, const double* __restrict__ scratch,
                                                                         double* __restrict__ out) {
  const int j = threadIdx.x;
  if (j >= n_stats) return;
  double v = 0;
  for (int b = 0; b < blocks; ++b) v += scratch[(size_t)b * kNoAllocRow + j];
  out[j] = v;
}

extern "C" int dm_batch_stats_noalloc(int device, int32_t n, int32_t obs_dim, const float* obs, const float* reward,
                                      const uint8_t* done, const float* ep_return, const int32_t* ep_len, double* out_stats,
                                      void* stream) {
  if (n < 0 || n > DM_NOALLOC_MAX_ENVS)
    return fail(nullptr, DM_E_ARG, "dm_batch_stats_noalloc: n = %d is outside [0, %d] (DM_NOALLOC_MAX_ENVS)", n, DM_NOALLOC_MAX_ENVS);
  if (obs_dim < 1 || obs_dim > DM_NOALLOC_MAX_OBS_DIM)
    return fail(nullptr, DM_E_ARG, "dm_batch_stats_noalloc: obs_dim = %d is outside [1, %d] (DM_NOALLOC_MAX_OBS_DIM: the static scratch "
                "holds %d doubles per block)", obs_dim, DM_NOALLOC_MAX_OBS_DIM, kNoAllocRow);
  static_assert(6 + 2 * DM_NOALLOC_MAX_OBS_DIM <= kNoAllocRow, "scratch row too short");
  if (!out_stats || (n > 0 && (!obs || !reward || !done)))
    return fail(nullptr, DM_E_ARG, "dm_batch_stats_noalloc: a required buffer is NULL (obs reward done out_stats)");
  if (device < 0 || device >= 64) return fail(nullptr, DM_E_ARG, "dm_batch_stats_noalloc: device %d out of range", device);
  CU(nullptr, cudaSetDevice(device));
  double* scratch;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {  // once per device, with that device current
      void* p = nullptr;
      CU(nullptr, cudaGetSymbolAddress(&p, g_noalloc_scratch));
      g_noalloc_ptr[device] = (double*)p;
    }
    scratch = g_noalloc_ptr[device];
  }
  cudaStream_t s = (cudaStream_t)stream;
  int blocks = (n + 2047) / 2048;
  blocks = blocks < 1 ? 1 : (blocks > kNoAllocBlocks ? kNoAllocBlocks : blocks);
  stats_partial_kernel<<<blocks, dim3(32, 8), 0, s>>>(n, obs_dim, obs, reward, done, ep_return, ep_len, scratch);
  CU(nullptr, cudaGetLastError());
  stats_final_kernel<<<1, kNoAllocRow, 0, s>>>(blocks, 6 + 2 * obs_dim, scratch, out_stats);
  CU(nullptr, cudaGetLastError());
  return DM_OK;
}

// ---- get_abspos_obs as an opt-in post-pass (K8, dm_kernels.cuh) -----------------------------------------------------
extern "C" int dm_obs_abspos(dm_handle* h, int32_t n, const float* qpos, const float* obs_in, float* obs_out, float* reward,
                             uint8_t* done, uint8_t* reason, const uint8_t* mask, void* stream) {
  if (!h) return fail(nullptr, DM_E_ARG, "dm_obs_abspos: NULL handle");
  if (n < 0 || (n > 0 && (!qpos || !obs_in || !obs_out))) return fail(h, DM_E_ARG, "dm_obs_abspos: a required buffer is NULL (qpos obs_in obs_out)");
  if (reward && !done) return fail(h, DM_E_ARG, "dm_obs_abspos: step mode (reward given) needs the done flags");
  if (obs_in == obs_out) return fail(h, DM_E_ARG, "dm_obs_abspos: the extended rows are wider than the input rows: not in place");
  if (n == 0) return DM_OK;
  CU(h, cudaSetDevice(h->device));
  AbsPosParams p{n, qpos, obs_in, obs_out, reward, done, reason, mask};
  int rc = dispatch(h, [&](auto rb, auto kind) -> int {
    using Rb = decltype(rb);
    constexpr int KIND = decltype(kind)::value;
    obs_abspos_kernel<KIND, Rb><<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(p);
    return DM_OK;
  });
  if (rc) return rc;
  CU(h, cudaGetLastError());
  h->launches++;
  return DM_OK;
}

// ---- the statistics collective (K7, dm_peer.cuh) ------------------------------------------------------------------------
static int peer_ensure(dm_handle* h) {  // device buffers of the collective, allocated on first use
  if (h->d_total) return DM_OK;
  CU(h, cudaSetDevice(h->device));
  CU(h, cudaMalloc(&h->d_snap, sizeof(double) * kPeerStride));
  CU(h, cudaMalloc(&h->d_total, sizeof(double) * kPeerStride));
  CU(h, cudaMalloc(&h->d_peer_state, sizeof(PeerState)));
  CU(h, cudaMemset(h->d_snap, 0, sizeof(double) * kPeerStride));
  CU(h, cudaMemset(h->d_total, 0, sizeof(double) * kPeerStride));
  CU(h, cudaMemset(h->d_peer_state, 0, sizeof(PeerState)));
  return DM_OK;
}

extern "C" int dm_stats_peer_export(dm_handle* h, int32_t rank, int32_t world, void* handle_out) {
  if (!h || !handle_out) return fail(h, DM_E_ARG, "dm_stats_peer_export: NULL argument");
  if (world < 1 || world > kPeerMaxWorld || rank < 0 || rank >= world)
    return fail(h, DM_E_ARG, "dm_stats_peer_export: rank %d / world %d (at most %d ranks: the GPUs of one box)", rank, world, kPeerMaxWorld);
  if (h->dims.n_stats > kPeerStride) return fail(h, DM_E_STATE, "dm_stats_peer_export: %d statistics exceed the slot size", h->dims.n_stats);
  if (h->peer_world > 1 || h->peer_comm[h->peer_rank]) return fail(h, DM_E_STATE, "dm_stats_peer_export: this handle is already part of a group");
  int rc = peer_ensure(h);
  if (rc) return rc;
  static_assert(sizeof(cudaIpcMemHandle_t) == DM_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t size");
  PeerComm* c = nullptr;
  CU(h, cudaMalloc(&c, sizeof(PeerComm)));
  CU(h, cudaMemset(c, 0, sizeof(PeerComm)));
  CU(h, cudaDeviceSynchronize());
  cudaIpcMemHandle_t ih;
  cudaError_t e = cudaIpcGetMemHandle(&ih, c);
  if (e != cudaSuccess) {
    cudaFree(c);
    return fail(h, DM_E_CUDA, "dm_stats_peer_export: cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  }
  h->peer_rank = rank;
  h->peer_comm[rank] = c;
  memcpy(handle_out, &ih, sizeof ih);
  return DM_OK;
}

extern "C" int dm_stats_peer_import(dm_handle* h, int32_t world, const void* handles) {
  if (!h || !handles) return fail(h, DM_E_ARG, "dm_stats_peer_import: NULL argument");
  if (!h->peer_comm[h->peer_rank]) return fail(h, DM_E_STATE, "dm_stats_peer_import: call dm_stats_peer_export first");
  if (world < 1 || world > kPeerMaxWorld || h->peer_rank >= world) return fail(h, DM_E_ARG, "dm_stats_peer_import: world %d", world);
  CU(h, cudaSetDevice(h->device));
  for (int r = 0; r < world; ++r) {
    if (r == h->peer_rank) continue;
    cudaIpcMemHandle_t ih;
    memcpy(&ih, (const char*)handles + (size_t)r * DM_IPC_HANDLE_BYTES, sizeof ih);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, ih, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      for (int q = 0; q < r; ++q)
        if (q != h->peer_rank && h->peer_comm[q]) { cudaIpcCloseMemHandle(h->peer_comm[q]); h->peer_comm[q] = nullptr; }
      return fail(h, DM_E_CUDA, "dm_stats_peer_import: cudaIpcOpenMemHandle of rank %d: %s (use the NCCL route: dm_stats_snapshot / "
                  "dm_stats_commit)", r, cudaGetErrorString(e));
    }
    h->peer_comm[r] = (PeerComm*)p;
  }
  h->peer_world = world;
  return DM_OK;
}

extern "C" int dm_stats_allreduce(dm_handle* h, void* stream) {
  if (!h) return fail(nullptr, DM_E_ARG, "dm_stats_allreduce: NULL handle");
  int rc = peer_ensure(h);
  if (rc) return rc;
  CU(h, cudaSetDevice(h->device));
  if (h->coll_stream) {  // exchanges are numbered: one issued by the caller goes behind those of the auto mode
    CU(h, cudaEventRecord(h->coll_event, h->coll_stream));
    CU(h, cudaStreamWaitEvent((cudaStream_t)stream, h->coll_event, 0));
  }
  rc = issue_exchange(h, (cudaStream_t)stream);
  if (rc == DM_OK && h->coll_stream) {  // ... and those the auto mode issues next go behind this one (they share the sequence number)
    CU(h, cudaEventRecord(h->coll_event, (cudaStream_t)stream));
    CU(h, cudaStreamWaitEvent(h->coll_stream, h->coll_event, 0));
    h->exch_where = 0;
  }
  return rc;
}

static PeerParams peer_params(const dm_handle* h, int enable) {
  PeerParams p{};
  p.acc = h->d_stats; p.snap = h->d_snap; p.total = h->d_total; p.st = h->d_peer_state;
  for (int r = 0; r < kPeerMaxWorld; ++r) p.comm[r] = h->peer_comm[r];
  p.rank = h->peer_rank; p.world = h->peer_world; p.n = h->dims.n_stats; p.enable = enable;
  return p;
}

static int issue_exchange(dm_handle* h, cudaStream_t stream) {
  stats_exchange_kernel<<<1, kPeerStride, 0, stream>>>(peer_params(h, 0));
  CU(h, cudaGetLastError());
  h->launches++;
  return DM_OK;
}

extern "C" int dm_stats_set_auto(dm_handle* h, int32_t every) {
  if (!h || every < 0) return fail(h, DM_E_ARG, "dm_stats_set_auto: every must be >= 0");
  int rc = peer_ensure(h);
  if (rc) return rc;
  CU(h, cudaSetDevice(h->device));
  if (every > 0 && !h->coll_stream) {
    CU(h, cudaStreamCreateWithFlags(&h->coll_stream, cudaStreamNonBlocking));
    CU(h, cudaEventCreateWithFlags(&h->coll_event, cudaEventDisableTiming));
  }
  h->auto_every = every;
  h->auto_count = 0;
  return DM_OK;
}

// `stream` waits for every exchange issued so far (the auto mode's side stream): call it before reading the totals on
// `stream`, before timing marks, before dm_stats_zero
extern "C" int dm_stats_join(dm_handle* h, void* stream) {
  if (!h) return fail(nullptr, DM_E_ARG, "dm_stats_join: NULL handle");
  if (!h->coll_stream) return DM_OK;
  CU(h, cudaSetDevice(h->device));
  CU(h, cudaEventRecord(h->coll_event, h->coll_stream));
  CU(h, cudaStreamWaitEvent((cudaStream_t)stream, h->coll_event, 0));
  return DM_OK;
}

extern "C" int dm_stats_snapshot(dm_handle* h, double** snapshot_dev_ptr, void* stream) {
  if (!h) return fail(nullptr, DM_E_ARG, "dm_stats_snapshot: NULL handle");
  int rc = peer_ensure(h);
  if (rc) return rc;
  CU(h, cudaSetDevice(h->device));
  stats_snapshot_kernel<<<1, kPeerStride, 0, (cudaStream_t)stream>>>(h->d_stats, h->d_snap, h->dims.n_stats);
  CU(h, cudaGetLastError());
  h->launches++;
  if (snapshot_dev_ptr) *snapshot_dev_ptr = h->d_snap;
  return DM_OK;
}

extern "C" int dm_stats_commit(dm_handle* h, void* stream) {
  if (!h || !h->d_total) return fail(h, DM_E_STATE, "dm_stats_commit: no snapshot (call dm_stats_snapshot first)");
  CU(h, cudaSetDevice(h->device));
  stats_commit_kernel<<<1, kPeerStride, 0, (cudaStream_t)stream>>>(h->d_snap, h->d_total, h->dims.n_stats);
  CU(h, cudaGetLastError());
  h->launches++;
  return DM_OK;
}

extern "C" int dm_stats_read_total(dm_handle* h, double* host_out, void* stream) {
  if (!h || !host_out) return fail(h, DM_E_ARG, "dm_stats_read_total: NULL argument");
  int rc = peer_ensure(h);
  if (rc) return rc;
  CU(h, cudaSetDevice(h->device));
  unsigned int err = 0;
  if (h->coll_stream) CU(h, cudaStreamSynchronize(h->coll_stream));
  CU(h, cudaMemcpyAsync(host_out, h->d_total, sizeof(double) * h->dims.n_stats, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CU(h, cudaMemcpyAsync(&err, &h->d_peer_state->err, sizeof err, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CU(h, cudaStreamSynchronize((cudaStream_t)stream));
  if (err) return fail(h, DM_E_STATE, "dm_stats_read_total: a rank did not reach dm_stats_allreduce within %.1f s; the totals lack its share",
                       kPeerTimeoutNs * 1e-9);
  return DM_OK;
}
