/*
 * bcp.h — C ABI of the B200-native batched BulletCartpole reset/step path.
 *
 * The reference (rebcabin/cartpoleplusplus) has no FFI layer of its own: its hot
 * path is `bullet_cartpole.py` calling the third-party `pybullet` C extension
 * in-process.  This header is the boundary a maintainer would bind instead of
 * those pybullet calls; every entry point names the reference interface it
 * replaces (file:line into the reference tree).
 *
 * Conventions
 *   - extern "C", POD only, no torch types.  All tensor arguments are raw DEVICE
 *     pointers owned by the caller (PyTorch allocates them); nothing is retained
 *     past the call.  `stream` is a cudaStream_t passed as void*.
 *   - every function returns 0 on success, <0 on error; the message is available
 *     through bcp_last_error().  Nothing throws across the ABI.
 *   - work is enqueued asynchronously on `stream`; no host synchronisation inside
 *     reset/step/rollout.
 */
#ifndef BCP_H
#define BCP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BCP_ABI_VERSION 8

/* ---- dtype of the physics arithmetic ------------------------------------ */
#define BCP_DTYPE_F32 0 /* fast mode                                          */
#define BCP_DTYPE_F64 1 /* parity mode (pybullet's btScalar is double)        */

/* ---- done reasons (info['done_reason'], bullet_cartpole.py:246,251,256) -- */
#define BCP_REASON_NONE 0
#define BCP_REASON_POSITION 1    /* 'out of position bounds'    :245-248 */
#define BCP_REASON_ORIENTATION 2 /* 'out of orientation bounds' :249-253 */
#define BCP_REASON_LENGTH 3      /* 'episode length'            :255-257 */
/* pipelined auto-reset only (no reference counterpart; see bcp_config.auto_reset): */
#define BCP_REASON_RESETTING 4   /* env is inside its 130-substep reset: obs untouched,
                                    reward 0, done 1                                  */
#define BCP_REASON_FRESH 5       /* reset finished in this call: obs = reset observation
                                    (:338-339), reward 0, done 0                      */

/* ---- what bcp_step does with an env that is done ------------------------- */
#define BCP_AUTO_RESET_OFF 0       /* the reference: 'calling step after done????' -
                                      old obs, reward 0, done 1 (:179-181); the caller
                                      resets with bcp_reset(mask)                     */
#define BCP_AUTO_RESET_PIPELINED 1 /* the env runs its reset (:313-346) R*S substeps per
                                      bcp_step call, exactly as many substeps as a live
                                      env, and reports RESETTING / FRESH meanwhile     */

/* ---- applyExternalForce frame (bullet_cartpole.py:203 vs the remnant :352) */
#define BCP_FRAME_LINK 0
#define BCP_FRAME_WORLD 1

/* ---- initial-push direction draw (bullet_cartpole.py:329-332,354-359) ---- */
#define BCP_THETA_PER_SUBSTEP 0 /* this fork: a new theta on every push substep */
#define BCP_THETA_PER_RESET 1   /* upstream: one theta per reset                */

/* One box body as the host URDF parser extracts it (models/{ground,cart,pole}.urdf). */
typedef struct bcp_body {
  double mass;            /* 0 => static (ground.urdf:6)                          */
  double inertia[3];      /* effective local inertia diagonal, inertia_scaling
                             already applied                                       */
  double half_extents[3]; /* collision box half sizes                              */
  double friction;        /* <lateral_friction>; Bullet default 0.5 when absent    */
  double restitution;     /* <restitution>; Bullet default 0                       */
  double init_pos[3];     /* loadURDF / reset position, bullet_cartpole.py:154-156 */
} bcp_body;

typedef struct bcp_config {
  int32_t abi_version; /* must be BCP_ABI_VERSION */
  int32_t num_envs;    /* envs owned by this handle (this rank's shard) */
  int64_t env_id_offset; /* global id of local env 0 (RNG key => shard-invariant) */
  uint64_t seed;
  int32_t dtype;          /* BCP_DTYPE_* */
  int32_t device_ordinal; /* CUDA device */

  /* --- env options: add_opts, bullet_cartpole.py:15-40; ctor :50-122 --- */
  int32_t discrete_actions;    /* ctor arg :50,:88 */
  int32_t action_repeats;      /* --action-repeats :23 */
  int32_t steps_per_repeat;    /* --steps-per-repeat :25 */
  int32_t max_episode_len;     /* --max-episode-len :31 */
  int32_t initial_force_steps; /* 30, :76 */
  int32_t settle_steps;        /* 100, :326 */
  int32_t random_theta;        /* !--no-random-theta :81 */
  int32_t theta_mode;          /* BCP_THETA_* */
  int32_t force_frame;         /* BCP_FRAME_* */
  int32_t termination_enabled; /* the rule of :243-253 (commented out in the fork) */
  int32_t carry_over_reset;    /* 1: reset keeps the pending force and the contact
                                  caches exactly as pybullet does (:319-327)      */
  int32_t auto_reset;          /* BCP_AUTO_RESET_* */
  double action_force;    /* --action-force :18 */
  double initial_force;   /* --initial-force :20 */
  double pos_threshold;   /* 3.0, :58 */
  double angle_threshold; /* 0.35, :62 */
  double sin_angle_threshold; /* host-precomputed sin/tan of the above: the done */
  double tan_angle_threshold; /* test is evaluated trig-free     */

  /* --- the three boxes --- */
  bcp_body ground, cart, pole;

  /* --- physics engine parameters (pybullet defaults; [BULLET-RECALL]) --- */
  double dt;         /* 1/240 */
  double gravity[3]; /* (0,0,-9.81) bullet_cartpole.py:152 */
  int32_t solver_iterations; /* 50 */
  int32_t friction_dirs;     /* 2 (btMultiBodyDynamicsWorld) */
  int32_t cone_friction;     /* 1: implicit cone over the two directions */
  int32_t manifold_order;    /* 0: rows in reverse pair-creation order
                                   (ground-pole, cart-pole, ground-cart);
                                1: creation order */
  double residual_threshold; /* leastSquaresResidualThreshold 1e-7 */
  double erp2;               /* 0.08 */
  double linear_slop;        /* 1e-5 */
  double global_cfm;         /* 0 */
  double friction_erp;       /* 0.2 (unused: no friction anchors) */
  double friction_cfm;       /* 0 */
  double restitution_velocity_threshold; /* 0.2 */
  double linear_damping;     /* 0.04f */
  double angular_damping;    /* 0.04f */
  double max_coordinate_velocity; /* 100 */
  double contact_breaking_factor; /* gContactBreakingThreshold 0.02 */
  double max_friction;       /* 10 (combined friction clamp) */
  double warmstart_factor;   /* cached normal impulse x factor seeds the next solve;
                                0.85 (btContactSolverInfo default): a calibrated
                                choice, not pinned; 0 = off */
  int32_t gyro_term;         /* 1 */
  int32_t reset_settle_cache; /* 1 (only with carry_over_reset = 0): the 100 settle substeps of reset() (:326-327), which then
                                 start from the same state in every env, are computed once per handle and their result is
                                 replayed; a reset costs its 30 push substeps.  0: every reset runs them (the reference)    */
  double cart_ground_friction; /* >= 0: lateral friction of the cart AGAINST THE GROUND, when it differs from cart.friction
                                  (what the pole touches): the fork's compound cart, models/double_cart_1.urdf:15-80 - upper half
                                  1.0, lower half 0.0.  < 0: cart.friction serves both pairs (cart.urdf)                       */
} bcp_config;

/* ---- canonical per-env state record (doubles), used by get/set_state ------
 * and by the parity tests to start oracle and kernel from identical states.   */
#define BCP_S_CART 0      /* pos3, quat4 (x,y,z,w), linvel3, angvel3 */
#define BCP_S_POLE 13
#define BCP_S_PEND_F 26   /* pending external force on the cart (world)  */
#define BCP_S_PEND_T 29   /* pending external torque on the cart (world) */
#define BCP_S_MANI 32     /* 3 manifolds x 4 points x 11 scalars         */
#define BCP_MANI_POINT 11 /* localA3, localB3, normalWorldOnB3, distance,
                             applied normal impulse (warm-start cache)   */
#define BCP_MANI_STRIDE 44
#define BCP_S_MANI_CNT 164 /* 3 point counts */
#define BCP_S_STEPS 167
#define BCP_S_DONE 168
#define BCP_S_EPISODE 169  /* number of resets so far (RNG counter word) */
#define BCP_S_RESET_LEFT 170 /* pipelined auto-reset: reset substeps still to run */
#define BCP_STATE_DOUBLES 176

/* manifold (collision pair) indices, in Bullet's pair-creation order */
#define BCP_PAIR_GROUND_CART 0
#define BCP_PAIR_CART_POLE 1
#define BCP_PAIR_GROUND_POLE 2

/* stats record filled by bcp_stats */
#define BCP_STATS_HEADER 4 /* episodes, env_steps, substeps, resets */

typedef struct bcp_env bcp_env; /* opaque handle: one per CUDA device / shard */

/* version / sizes (lets a binding check its struct mirror) */
int bcp_abi_version(void);
size_t bcp_config_size(void);
const char* bcp_last_error(void);

/* Replaces BulletCartpole.__init__'s world construction:
 * p.connect/setGravity/loadURDF x3, bullet_cartpole.py:151-156.            */
int bcp_create(const bcp_config* cfg, bcp_env** out);
int bcp_destroy(bcp_env* env);

/* Replaces BulletCartpole.reset, bullet_cartpole.py:313-346: teleport cart and
 * pole (:319-320), 100 settle substeps (:326-327), 30 push substeps (:329-332),
 * observation bootstrap (:338-339).
 *   mask    : device u8[num_envs] or NULL (NULL = reset every env)
 *   obs_out : device f32[num_envs][R][2][7] or NULL                          */
int bcp_reset(bcp_env* env, const uint8_t* mask, float* obs_out, void* stream);

/* Replaces BulletCartpole.step, bullet_cartpole.py:178-275: R x S times
 * [p.stepSimulation(); p.applyExternalForce(cart, action*action_force)]
 * (:197-203), pose capture per repeat (:237), steps/done/reward (:239-260).
 *   actions : discrete  -> device i32[num_envs] in 0..4
 *             continuous-> device f32[num_envs][2] in [-1,1]
 *   obs_out f32[num_envs][R][2][7], reward_out f32[num_envs],
 *   done_out u8[num_envs], reason_out u8[num_envs] (may be NULL)
 * An env that is already done returns its old observation, reward 0, done 1
 * (:179-181).                                                                */
int bcp_step(bcp_env* env, const void* actions, float* obs_out, float* reward_out,
             uint8_t* done_out, uint8_t* reason_out, void* stream);

/* Random-agent protocol of README.md:76-93 run entirely on the device, state in
 * registers across the whole launch: n_calls rounds of R*S substeps per env with
 * actions drawn in-kernel (Philox4x32-10 keyed by global env id, episode, step);
 * an env that finishes records its episode length in the histogram and runs its
 * reset (:313-346) in the following rounds, R*S substeps per round, so every lane
 * of a warp executes one physics substep per iteration whatever its phase.
 *   action_set_mask : discrete only, bit a set => action a may be drawn
 *                     (--actions="0,1,2,3,4" == 0x1f)
 *   max_episodes    : > 0: an env idles once it has finished that many episodes
 *                     (the protocol's --num-eval split over the envs; makes the
 *                     histogram independent of n_calls and of the sharding);
 *                     0: no limit
 *   obs_out may be NULL; else the envs' latest observation
 * fp32 handles of >= 4 096 envs run the n_calls rounds through the split pipeline
 * (the action draw as a kernel of its own before each round's kernels); smaller
 * and fp64 handles in one launch.  Same machine and action stream either way.  */
int bcp_rollout_random(bcp_env* env, int32_t n_calls, uint32_t action_set_mask,
                       int32_t max_episodes, float* obs_out, void* stream);

/* Episode statistics accumulated on the device since the last clear:
 *   hist : host i64[max_episode_len+1], hist[L] = episodes that ended at length L
 *   header: host i64[BCP_STATS_HEADER] = episodes, env_steps, substeps, resets
 *   reward_sum: host f64[1]                                                   */
int bcp_stats(bcp_env* env, int64_t* hist, int64_t* header, double* reward_sum,
              void* stream);
int bcp_stats_clear(bcp_env* env, void* stream);

/* Canonical state export/import (host buffers, BCP_STATE_DOUBLES doubles per
 * env); the FP64 parity tests start oracle and kernel from identical states.   */
int bcp_get_state(bcp_env* env, double* host_buf, size_t n_doubles, void* stream);
int bcp_set_state(bcp_env* env, const double* host_buf, size_t n_doubles, void* stream);

/* bcp_step for a HOST-side agent loop: actions in from, and the four outputs out to, host buffers
 * (pinned memory for the copies to overlap), all enqueued on `stream`.  The observation leaves
 * repeat by repeat - the slab of repeat r is copied on a second stream as soon as the kernel that
 * completes it is done, beside the remaining substeps - so it arrives REPEAT-MAJOR:
 *   h_obs_rm f32[R][num_envs][2][7]   (h_obs_rm[r][e] == obs_out[e][r] of bcp_step)
 *   h_actions i32[num_envs] | f32[num_envs][2];  h_reward f32[num_envs], h_done / h_reason u8[num_envs]
 * bcp_step_host_finish synchronises `stream` and completes the buffers on the host (the envs whose
 * pipelined reset ended in the call receive their reset observation in every repeat slot only when
 * the call closes, :338-339); call it before reading.                                            */
int bcp_step_host(bcp_env* env, const void* h_actions, float* h_obs_rm, float* h_reward, uint8_t* h_done,
                  uint8_t* h_reason, void* stream);
int bcp_step_host_finish(bcp_env* env, float* h_obs_rm, const uint8_t* h_reason, void* stream);

/* One bare p.stepSimulation() for every env (no force application, no
 * bookkeeping) - bullet_cartpole.py:199 alone; used by single-substep parity. */
int bcp_substep(bcp_env* env, void* stream);

/* Measurement helper (bench.py only, no reference counterpart): dependent-free FFMA /
 * DFMA chains on every SM for ~`iters` rounds; writes the achieved TFLOP/s
 * (2 flops per FMA).  Gives the FP32 / FP64 CUDA-core pipe peak the roofline of this
 * compute-bound path is quoted against (MEASURED_PEAKS.json has no such entry). */
int bcp_bench_fma_peak(int device_ordinal, int use_f64, int iters, double* tflops_out);

/* Which pipeline bcp_step runs on this handle (reporting only; results agree to fp32 rounding):
 * the one-launch kernel (fp64, small batches), the split pipeline with the row-space solver, or
 * the split pipeline with the impulse-space solve phase.              */
#define BCP_SOLVER_ONE_LAUNCH 0
#define BCP_SOLVER_SPLIT_ROW 1
#define BCP_SOLVER_SPLIT_IMPULSE 2
int bcp_solver_kind(const bcp_env* env);

/* number of kernel launches this handle has enqueued (bench's gpu_launches) */
int64_t bcp_launch_count(const bcp_env* env);

/* Measurement helper (bench.py only): between begin and end every kernel bcp_step launches
 * is bracketed by CUDA events on the caller's stream.  end() synchronises and writes
 *   out4 = { ms in k_prep (or the one-launch k_step), ms in k_solve, launches of the
 *            former, launches of the latter }                                          */
int bcp_profile_begin(bcp_env* env);
int bcp_profile_end(bcp_env* env, double* out4);

/* Measurement helper (bench.py only): what the constraint solve of the LAST substep did, per
 * env, read back from the split pipeline's solve records:
 *   host_out[4 * e + 0..3] = { ground-cart, cart-pole, ground-pole contact points, PGS iterations }
 * (stale for an env that did not run that substep).  n_ints = 4 * num_envs.  Synchronises.
 * Error unless the handle runs the split pipeline (fp32, >= 4096 envs).                  */
int bcp_last_solve_info(bcp_env* env, int32_t* host_out, size_t n_ints, void* stream);

/* Test instrumentation (no reference counterpart): with tracing enabled, every bcp_step on the split pipeline also
 * keeps, for each of its R*S substeps, what bcp_last_solve_info reports for the last one; the parity tests classify
 * an env whose fp32 pose leaves the fp64 oracle's by more than 1e-4 by the discrete decision that differs
 * (tests/test_gpu_pipeline_parity.py).  One device-to-device copy per substep while enabled, nothing otherwise.
 *   host_out[(k * num_envs + e) * 4 + 0..3] = substep k of the last call, env e: the three contact counts, iterations
 *   n_ints = 4 * R * S * num_envs.  read synchronises.                                                     */
int bcp_solve_trace_enable(bcp_env* env, int on);
int bcp_solve_trace_read(bcp_env* env, int32_t* host_out, size_t n_ints, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BCP_H */
