/* dmhot -- C ABI of the B200 batched DeepMimic environment hot path (libdmhot.so).
 *
 * The reference (danieldugas/DeepMimic_mujoco) is pure Python and has no FFI layer; the boundary
 * this library sits behind is its gym / VecEnv API.  Each entry point below names the reference
 * code it replaces (paths relative to the reference root).  INTEGRATION.md shows the ctypes stub a
 * maintainer would add to src/deepmimic_env.py / src/combined_env.py.
 *
 * Conventions: every function returns 0 on success and a negative DM_E_* code on failure, never
 * throws, never allocates caller-visible memory.  dm_last_error() returns a message for the last
 * failure on that handle (or the last global failure for handle == NULL).  All `const float*` /
 * `float*` step buffers are DEVICE pointers owned by the caller (PyTorch); calls are asynchronous
 * on the given CUDA stream (a cudaStream_t passed as void*).  The *_host variants take HOST
 * pointers, stage through pinned memory owned by the handle and return after the results are on
 * the host.  A handle belongs to one device and is not thread-safe.
 */
#ifndef DMHOT_H_
#define DMHOT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DM_ABI_VERSION 1

enum { DM_ROBOT_HUMANOID3D = 0, DM_ROBOT_UNITREE_G1 = 1 }; /* src/config.py:3-22 RobotConfig */
enum { DM_ENV_SINGLE = 0,     /* DPEnv          src/deepmimic_env.py:273-510 */
       DM_ENV_COMBINED = 1 }; /* DPCombinedEnv  src/combined_env.py:102-533 (unitree_g1 only) */
enum { DM_LAYOUT_AOS = 0,     /* row-major [n, dim]: what gym / SB3 hand over */
       DM_LAYOUT_SOA = 1 };   /* [dim, ld] with ld >= n: one coalesced column per quantity */

enum { DM_CLIP_FLOOR = 1,   /* motion in MotionConfig.floor_motions      src/config.py:37 */
       DM_CLIP_ACYCLIC = 2, /* motion in MotionConfig.acyclical_motions  src/config.py:36 */
       DM_CLIP_G1_RUN = 4 };/* "run" on unitree_g1: roll/pitch limit     src/deepmimic_env.py:426-433 */

/* done_reason codes (info["done_reason"], src/deepmimic_env.py:424-442, src/combined_env.py:436,445);
 * 8 = observation-bounds guard (returns obs*0, 0, True, {} -- src/deepmimic_env.py:465-476) */
enum { DM_REASON_NONE = 0, DM_REASON_LOW_Z, DM_REASON_HIGH_Z, DM_REASON_RUN_ROLL, DM_REASON_RUN_PITCH,
       DM_REASON_MAX_EP_LEN, DM_REASON_ACYCLICAL_END, DM_REASON_FALLEN_NO_AMNESTY, DM_REASON_OBS_BOUNDS };

enum { DM_OK = 0, DM_E_ARG = -1, DM_E_CUDA = -2, DM_E_STATE = -3, DM_E_NOMEM = -4 };

#define DM_MAX_CLIPS 8
#define DM_NTERMS 8 /* reward_config, reward_qvel, reward_end_eff, reward_com, reward_joint_limit,
                       imitation_reward, task_reward, z_com */
/* contact word: bit 0 rfoot&floor, bit 1 lfoot&floor (get_foot_contact_obs, deepmimic_env.py:78-105),
 * bits 2..9 the eight extra-contact geoms & floor (get_extra_contact_obs, :107-121) */

typedef struct dm_handle dm_handle;

typedef struct {
  int32_t nq, nv, nu, na;      /* na = exposed action dims (nu - 14 hand motors on G1) */
  int32_t obs_dim;             /* 67 / 85 / 98 */
  int32_t nbody, ngeom, nsel;  /* nsel = joints in the reward (deepmimic_env.py:206-207) */
  int32_t row_stride;          /* floats per packed clip row */
  int32_t n_stats;             /* doubles in the statistics vector: 6 + 2 * obs_dim */
  int32_t robot, env_kind;
} dm_dims_t;

/* One mocap clip = the tables MocapDM.read_raw_data builds (src/mujoco/mocap_v2.py:33-336), float64,
 * row-major, HOST pointers.  fixed_row >= 0 makes it a MotionTransition (src/combined_env.py:67-99:
 * every lookup returns that frame, `len` is its virtual length). */
typedef struct {
  int32_t len;            /* get_length() */
  int32_t n_rows;         /* frames actually present in the arrays below (== len unless fixed_row >= 0) */
  int32_t flags;          /* DM_CLIP_* */
  int32_t fixed_row;      /* -1, or the constant frame (1 for MTToGetup) */
  const double* qpos;      /* [n_rows, nq]         data_config    */
  const double* qvel;      /* [n_rows, nv]         data_vel       */
  const double* body_xpos; /* [n_rows, nbody, 3]   data_body_xpos */
  const double* geom_xpos; /* [n_rows, ngeom, 3]   data_geom_xpos */
} dm_clip_t;

/* Per-step buffers.  For DM_ENV_SINGLE `idx` is idx_curr and `clip` selects the clip of each env
 * (NULL = clip 0 for all); for DM_ENV_COMBINED `idx` is current_motion_n_steps and `clip` the motion
 * state (0 walk, 1 run, 2 getup, 3 to_getup).  idx / ep_len / clip are updated in place. */
typedef struct {
  int32_t n;
  int32_t layout;           /* DM_LAYOUT_* for every float buffer */
  int64_t ld;               /* SoA leading dimension (ignored for AoS) */
  const float* qpos;        /* [n, nq]  state after the physics step (force_state semantics) */
  const float* qvel;        /* [n, nv] */
  const float* action;      /* [n, na] */
  const uint32_t* contact;  /* [n] contact word, NULL = no contacts */
  int32_t* idx;             /* [n] in/out */
  int32_t* ep_len;          /* [n] in/out episode_length */
  int32_t* clip;            /* [n] in (single) / in-out (combined); may be NULL for single */
  float* obs;               /* [n, obs_dim] */
  float* reward;            /* [n] */
  uint8_t* done;            /* [n] */
  uint8_t* reason;          /* [n] DM_REASON_*; may be NULL */
  float* torque;            /* [n, nu] gear * clip(scale * action, ctrlrange); may be NULL */
  float* terms;             /* [n, DM_NTERMS]; may be NULL */
  float* ep_return;         /* [n] in/out episode_reward accumulator; may be NULL */
  int32_t accumulate_stats; /* != 0: add this batch to the handle's statistics accumulators */
  void* stream;             /* cudaStream_t */
} dm_step_args;

/* Reset (DPEnv.reset / reset_model / reference_state_init, src/deepmimic_env.py:312-316,496-510;
 * DPCombinedEnv.reset, src/combined_env.py:208-244) of the envs whose mask byte is non-zero.
 * frame: injected random.randint draw (NULL = counter-based device RNG from seed/counter).
 * coin : combined only -- 0 walk (+150+10 amnesty offset), 1 getup, 2 = rsi=False (getup, n_steps 0);
 *        NULL = device RNG coin flip.  qpos/qvel receive the reference frame (set_state). */
typedef struct {
  int32_t n;
  int32_t layout;
  int64_t ld;
  const uint8_t* mask;     /* [n] or NULL = all */
  const int32_t* frame;    /* [n] or NULL */
  const int32_t* coin;     /* [n] or NULL */
  const uint32_t* contact; /* [n] contacts at the reset pose, NULL = none */
  uint64_t seed, counter;
  float* qpos;             /* [n, nq] out */
  float* qvel;             /* [n, nv] out */
  int32_t* idx;            /* [n] out */
  int32_t* ep_len;         /* [n] out (0) */
  int32_t* clip;           /* [n] in (single, may be NULL) / out (combined) */
  float* obs;              /* [n, obs_dim] out (rows of reset envs only) */
  float* ep_return;        /* [n] out (0), may be NULL */
  float* terminal_obs;     /* [n, obs_dim] or NULL: rows of reset envs receive the obs they had before the reset
                              (SB3 VecEnv infos[i]["terminal_observation"]) */
  void* stream;
} dm_reset_args;

/* lifetime */
int dm_abi_version(void);
int dm_create(int device, int robot, int env_kind, dm_handle** out);
int dm_destroy(dm_handle* h);
const char* dm_last_error(const dm_handle* h);
int dm_get_dims(const dm_handle* h, dm_dims_t* out);

/* baked model constants (what MuJoCo's compiled model exposes: model.body_mass, model.jnt_range,
 * actuator gear / ctrlrange) -- lets the host check them against its own MJCF compilation */
int dm_get_model(const dm_handle* h, double* body_mass /*[nbody]*/, float* lim_lo /*[nsel]*/, float* lim_hi /*[nsel]*/,
                 float* gear /*[nu]*/, float* ctrl_lo /*[nu]*/, float* ctrl_hi /*[nu]*/, int32_t* ee_geom /*[4]*/);

/* FK of n key frames in float64 on the device: the load-time stage of MocapDM.read_raw_data
 * (src/mujoco/mocap_v2.py:292-307: set_state per frame, harvest body_xpos / geom_xpos).  HOST pointers. */
int dm_fk_tables(dm_handle* h, int32_t n, const double* qpos, const double* qvel, double* body_xpos, double* geom_xpos);

/* pack + upload the clip tables (replaces the per-process MocapDM objects a step indexes:
 * mocap.get_qpos/get_qvel/get_geom_xpos/get_body_xpos, src/mujoco/mocap_v2.py:338-348) */
int dm_upload_clips(dm_handle* h, int32_t n_clips, const dm_clip_t* clips);
/* Raw mocap clip = the parsed "Frames" array of a DeepMimic motion file (src/mujoco/motions/*.txt), float64, HOST.
 * same_as >= 0 reuses the compiled tables of that earlier clip of the same call (the MotionTransition pose source,
 * src/combined_env.py:67-99) and ignores frames. */
typedef struct {
  int32_t n_frames, n_cols;  /* 44 columns for DeepMimic rows (humanoid3d), nq + 1 for "Format": "direct_qpos" */
  int32_t direct_qpos;       /* "Format" == "direct_qpos"            src/mujoco/mocap_v2.py:49-51,271-272 */
  int32_t getup;             /* "getup" in the file path: VMX = 5    src/mujoco/mocap_v2.py:156-158 */
  int32_t flags;             /* DM_CLIP_* */
  int32_t fixed_row;         /* as dm_clip_t */
  int32_t virtual_len;       /* get_length() of a fixed_row clip; ignored otherwise */
  int32_t same_as;           /* -1, or an earlier clip index */
  const double* frames;      /* [n_frames, n_cols]; frames[0] of each row is the frame duration */
} dm_raw_clip_t;
/* The clip compiler on the device: MocapDM.read_raw_data (src/mujoco/mocap_v2.py:33-336: axis alignment, Euler
 * conversion, the shoulder / hip continuity fix, finite-difference velocities, FK of every key frame, x k resampling
 * to the simulator step) in float64 CUDA kernels, followed by the fp32 packing of dm_upload_clips -- the tables never
 * visit the host.  Replaces dm_upload_clips(MocapDM tables); fails with the reference's message for clips whose frame
 * duration is not a multiple of the simulator step (backflip, spin). */
int dm_compile_clips(dm_handle* h, int32_t n_clips, const dm_raw_clip_t* raw);
/* read back one compiled clip (float64, HOST, any pointer may be NULL): what MocapDM exposes as dt, data_config,
 * data_vel, data_body_xpos, data_geom_xpos.  Only valid after dm_compile_clips. */
int dm_get_compiled_clip(const dm_handle* h, int32_t clip, int32_t* len, double* dt, double* qpos, double* qvel,
                         double* body_xpos, double* geom_xpos);
/* read back the packed fp32 rows (host, [total_rows, row_stride]); rows == NULL just returns the count */
int dm_get_packed_rows(const dm_handle* h, float* rows, int32_t* total_rows);

/* the hot path: get_obs + calc_imitation_reward + termination / motion state machine + counters +
 * obs-bounds guard + action -> torque for n environments (DPEnv.step src/deepmimic_env.py:335-484,
 * DPCombinedEnv.step src/combined_env.py:246-493, with force_state semantics :355-357) */
int dm_step(dm_handle* h, const dm_step_args* a);
int dm_reset(dm_handle* h, const dm_reset_args* a);

/* same, HOST pointers, AoS only; returns when the outputs are in the caller's arrays.  Batches of at least one
 * tile run in the mode of dm_set_host_mode (default: the kernel works on pinned host memory directly -- the handle's
 * block, or the caller's own arrays where those are pinned); smaller ones by H2D copies, kernel, D2H copies.
 * h2d/d2h (may be NULL) receive the bytes that cross PCIe. */
int dm_step_host(dm_handle* h, const dm_step_args* a, int64_t* h2d_bytes, int64_t* d2h_bytes);
int dm_reset_host(dm_handle* h, const dm_reset_args* a);
/* Pinned host buffers for n environments, laid out exactly as dm_step_host stages them: every pointer field of
 * *out (qpos ... reason, incl. torque / terms / ep_return) receives a HOST pointer owned by the handle.  A caller
 * that keeps its per-step arrays there and passes those very pointers to dm_step_host pays no staging copies --
 * one kernel launch on the mapped block (DM_HOST_MAPPED), or one H2D transfer, the kernel, one D2H transfer
 * (DM_HOST_STAGED).  (The reference's SubprocVecEnv pickles every observation
 * through a pipe instead, src/sb3_ppo.py:275-278.)  Valid until a call with a larger n or dm_destroy. */
int dm_host_buffers(dm_handle* h, int32_t n, dm_step_args* out);
/* How dm_step_host moves a batch of at least one tile (128 environments) between host and device:
 *   DM_HOST_MAPPED  the step kernel reads and writes the pinned block of dm_host_buffers directly (its TMA bulk copies
 *                   and stores cross PCIe, both directions at once); caller arrays that are not the pinned slots are
 *                   staged in chunks of whole tiles, each chunk's kernel launched while the next is copied   [default]
 *   DM_HOST_STAGED  copy engines: H2D into a device block, the kernel on device memory, D2H
 * Same kernel and same results either way; batches smaller than a tile always take the staged form.
 * dm_reset_host follows the same mode: mapped, the reset kernel works on the pinned memory and only the rows of the
 * environments being reset cross PCIe. */
enum { DM_HOST_STAGED = 0, DM_HOST_MAPPED = 1 };
int dm_set_host_mode(dm_handle* h, int mode);

/* reward / observation-normaliser statistics (new functionality, no reference counterpart):
 * [count, sum r, sum r^2, n_done, sum ep_return(done), sum ep_len(done), sum obs[D], sum obs^2[D]] */
int dm_stats_zero(dm_handle* h, void* stream);
int dm_stats_device_ptr(dm_handle* h, double** dev_ptr, int32_t* n); /* for an in-place NCCL all-reduce */
int dm_stats_read(dm_handle* h, double* host_out, void* stream);

/* The collective of the path (the reference has none: SubprocVecEnv pipes only, src/sb3_ppo.py:275-278).
 * One process per GPU, every rank owns a contiguous block of environments and its own handle.  The step kernels add
 * into the rank's accumulators; dm_stats_allreduce SNAPSHOTS-AND-ZEROES them, exchanges the snapshot with every other
 * rank and adds the sum (taken in rank order: the same bits on every rank) to the handle's GLOBAL TOTALS, which
 * dm_stats_read_total returns and dm_policy_normalizer_from_stats uses from then on.  Only snapshots are ever reduced,
 * never running sums.  The exchange is ONE single-CTA kernel on `stream` that writes into the other GPUs' memory over
 * NVLink (cudaIpc peer mappings; every value travels with its sequence number, so there are no flags and no fences) and
 * waits for the peers' values there: no host round trip, no NCCL launch, capturable in a CUDA graph.
 * Every rank must call it the same number of times; a rank that waits longer than 2 s for a peer gives up and
 * dm_stats_read_total then fails with DM_E_STATE.  With one rank (no dm_stats_peer_import) it is the local
 * snapshot -> totals move.  dm_stats_zero also clears the snapshot and the totals.
 *   set-up, once:  dm_stats_peer_export(h, rank, world, handle[DM_IPC_HANDLE_BYTES])  -> all-gather the handles by any
 *                  means (the launcher's process group, MPI, a file) -> dm_stats_peer_import(h, world, handles[world][64]) on every rank.
 *   fallback where cudaIpc is not available (dm_stats_peer_import fails): dm_stats_snapshot(h, &dev_ptr, stream),
 *                  ncclAllReduce(dev_ptr, n_stats doubles, sum), dm_stats_commit(h, stream). */
#define DM_IPC_HANDLE_BYTES 64
int dm_stats_peer_export(dm_handle* h, int32_t rank, int32_t world, void* handle_out);
int dm_stats_peer_import(dm_handle* h, int32_t world, const void* handles);
int dm_stats_allreduce(dm_handle* h, void* stream);
int dm_stats_snapshot(dm_handle* h, double** snapshot_dev_ptr, void* stream);
int dm_stats_commit(dm_handle* h, void* stream);
int dm_stats_read_total(dm_handle* h, double* host_out, void* stream);
/* Auto mode: every `every`-th dm_step with accumulate_stats = 1 comes with a LAGGED exchange (0 = off): it takes in what
 * every rank posted one exchange earlier -- those values have had a whole step to cross NVLink -- and posts the sums
 * accumulated since; nothing waits for the exchange it posts.  Steps of at most one tile per SM (<= 128 x SM-count
 * environments) get it as a one-CTA kernel on a side stream owned by the handle, with no dependency on the step's
 * stream in either direction; larger steps carry it as ONE EXTRA CTA of the step kernel (no extra launch).  The global
 * totals therefore trail the steps by up to two exchanges while the loop runs; one dm_stats_allreduce after the last
 * step takes in what is pending and exchanges the remainder, after which dm_stats_read_total is exact.  Snapshots are
 * taken atomically per value, so no contribution is ever lost or counted twice.
 * dm_stats_join makes `stream` wait for every exchange issued on the side stream so far (dm_stats_allreduce does so itself). */
int dm_stats_set_auto(dm_handle* h, int32_t every);
int dm_stats_join(dm_handle* h, void* stream);

/* dm_batch_stats_noalloc -- NO COUNTERPART IN THE REFERENCE.
 * The same statistics vector as above for ONE batch, computed without any allocation: no cudaMalloc, cudaFree or
 * allocating library routine is called, so it can be used where those are forbidden (CUDA graph capture, tight serving
 * loops).  Its scratch (per-block partial sums) is a file-scope static __device__ array of 256 KiB that every GPU which
 * loads the library holds until exit.  No handle is needed.
 *   device      CUDA device ordinal (made current)
 *   n, obs_dim  environments in the batch, observation width; n <= DM_NOALLOC_MAX_ENVS, 1 <= obs_dim <=
 *               DM_NOALLOC_MAX_OBS_DIM, anything beyond is rejected with DM_E_ARG and a message in dm_last_error(NULL)
 *   obs [n, obs_dim] row-major, reward [n], done [n]    device pointers, required
 *   ep_return [n], ep_len [n]                            device pointers or NULL (their sums are then 0)
 *   out_stats   device double[6 + 2 obs_dim], fully overwritten:
 *               [count, sum r, sum r^2, n_done, sum ep_return(done), sum ep_len(done), sum obs[D], sum obs^2[D]]
 *   stream      cudaStream_t (NULL = default stream); asynchronous
 * Every call shares the one scratch array: calls that overlap on the GPU (different streams or host threads without
 * an ordering between them) race.  Results are deterministic (fixed summation order, no atomics). */
#define DM_NOALLOC_MAX_OBS_DIM 125
#define DM_NOALLOC_MAX_ENVS (1 << 22)
int dm_batch_stats_noalloc(int device, int32_t n, int32_t obs_dim, const float* obs, const float* reward, const uint8_t* done,
                           const float* ep_return, const int32_t* ep_len, double* out_stats, void* stream);

/* dm_obs_abspos -- get_abspos_obs (src/deepmimic_env.py:133-137: `mjdata.geom_xpos.flatten()` appended to the observation
 * when ENV_CFG.ADD_ABSPOS_OBS is set; False in every configuration the reference ships, src/deepmimic_env.py:267,
 * src/combined_env.py:31).  An opt-in POST-PASS over device buffers, after dm_step / dm_reset on the same stream:
 *   obs_out[e] = [ obs_in[e][0 : phase] | geom_xpos(qpos[e]) as 3 * ngeom floats | obs_in[e][phase : obs_dim] ]
 * (the reference's concatenation order, src/deepmimic_env.py:45; the world body's floor geom is at the origin).
 *   qpos [n, nq], obs_in [n, obs_dim], obs_out [n, obs_dim + 3 * ngeom]   row-major device pointers; not in place
 *   reward / done / reason   step mode (reward != NULL): the observation-bounds guard (src/deepmimic_env.py:465-476) is
 *                            applied to the new block: |coordinate| > 100 -> (obs * 0, 0, True, {}), reason = obs-bounds; a
 *                            row the step already zeroed gets its new block zeroed too.  NULL: rows are only extended
 *                            (after a reset).
 *   mask     NULL or [n]: only rows with mask != 0 are written (the environments a masked reset touched)
 * get_joint_force_obs (src/deepmimic_env.py:123-131) needs MuJoCo's constraint solver and is not provided. */
int dm_obs_abspos(dm_handle* h, int32_t n, const float* qpos, const float* obs_in, float* obs_out, float* reward, uint8_t* done,
                  uint8_t* reason, const uint8_t* mask, void* stream);

/* launch bookkeeping for bench.py ("gpu_launches") */
int dm_launch_count(const dm_handle* h, int64_t* kernels);
/* name + average device time (CUDA events on the launch stream) of the step kernel since the last call
 * with reset != 0; used for the roofline line */
int dm_step_kernel_time(dm_handle* h, int reset, double* avg_ms, int64_t* launches);
int dm_set_timing(dm_handle* h, int enabled);
/* Where the floor-contact bits of the observation come from (get_foot_contact_obs / get_extra_contact_obs,
 * src/deepmimic_env.py:78-121, read mjdata.contact):
 *   DM_CONTACT_CALLER  the `contact` words of dm_step_args / dm_reset_args (NULL = no contacts)      [default]
 *   DM_CONTACT_FROM_FK derived inside the kernels from forward kinematics: MuJoCo's plane-sphere / plane-box test of
 *                      the foot / heel / toe geoms against the floor plane, contact iff dist <= max(margins); the
 *                      caller's words are ignored.  Restated from MuJoCo 2.1 (parity unpinned: no MuJoCo here). */
enum { DM_CONTACT_CALLER = 0, DM_CONTACT_FROM_FK = 1 };
int dm_set_contact_mode(dm_handle* h, int mode);
/* 0 = generic strided kernel (K1a), 1 = shared-memory tile kernel with TMA bulk copies (K1b, AoS only), 2 = team kernel
 * (K1c, AoS only: one environment split over four warps, rotation-matrix FK; results agree with 0 / 1 to fp32 round-off,
 * not bit for bit); default -1 = auto = 1 */
int dm_set_kernel_variant(dm_handle* h, int variant);

/* ---- policy inference on the device -------------------------------------------------------------
 * The reference's ExtractedPolicy.act (src/extracted_policy.py:471-478): two tanh layers and a linear head, the
 * architecture sb3_ppo.py trains (net_arch [256, 128], src/sb3_ppo.py:265), preceded by the observation normaliser
 * clip((obs - mean) / sqrt(var + eps), +-clip_obs) fed from this library's running statistics.  One fused kernel;
 * the observations the step wrote never leave the GPU between environment and policy. */
typedef struct {
  int32_t in_dim, h0, h1, out_dim;     /* in_dim <= 128, h0 <= 256, h1 <= 128, out_dim <= 32 */
  const float *W0, *b0, *W1, *b1, *W2, *b2; /* HOST; W row-major [fan_in, fan_out]: np.matmul(x, W) + b */
} dm_policy_t;
typedef struct {
  int32_t n;
  int32_t normalize;      /* != 0: apply the observation normaliser before the first layer */
  const float* obs;       /* DEVICE [n, in_dim] */
  float* action;          /* DEVICE [n, out_dim]: the action mean (deterministic policy) */
  float* obs_norm;        /* DEVICE [n, in_dim] or NULL: the normalised observations (for a learner) */
  const float* reward;    /* DEVICE [n] or NULL */
  float* reward_norm;     /* DEVICE [n] or NULL: clip(reward * reward_scale, +-clip_reward) */
  float act_lo, act_hi;   /* clip of the action mean (SB3 predict() clips to the Box); -INF / +INF for none */
  void* stream;
} dm_policy_args;
int dm_policy_upload(dm_handle* h, const dm_policy_t* policy);
/* mean / var HOST [dim]; NULL mean = identity.  reward scale = 1 / sqrt(ret_var + eps). */
int dm_policy_set_normalizer(dm_handle* h, int32_t dim, const float* mean, const float* var, float eps, float clip_obs,
                             float ret_var, float clip_reward);
/* same, computed on the device from the statistics accumulators (dm_stats_*; population variance) */
int dm_policy_normalizer_from_stats(dm_handle* h, float eps, float clip_obs, float clip_reward, void* stream);
int dm_policy_act(dm_handle* h, const dm_policy_args* a);
/* 0 = fp32 on the CUDA cores, 1 = tcgen05 tensor cores with every fp32 operand split into two fp16 halves and three
 * MMAs per product (fp32-level accuracy); default 1 */
int dm_set_policy_variant(dm_handle* h, int variant);

#ifdef __cplusplus
}
#endif
#endif /* DMHOT_H_ */
