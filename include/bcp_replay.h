/*
 * bcp_replay.h — C ABI of the device-side replay memory fed by the batched env step
 * (same library, libbcp.so, same conventions as bcp.h).
 *
 * Replaces the reference's learner-side `ReplayMemory` (replay_memory.py:11-165): a ring of
 * (state_1_idx, action, reward, terminal_mask, state_2_idx) rows over a shared float16 state
 * pool with a FIFO list of free pool slots.  Everything lives in HBM; after any sequence of
 * calls the ring, the pool and the free list are bit-identical to what the reference class
 * holds after `add_episode` of the same episodes in the same order (tests/test_gpu_replay.py).
 *
 * Batched feeding (what ddpg_cartpole.py:308-326 / naf_cartpole.py:340-361 do one env at a
 * time — cache (action, reward, state_2) per step, add_episode when done): bcp_replay_begin
 * after a reset, bcp_replay_observe after every env.step.  Transitions are staged per env in
 * HBM (float16 states) and an episode enters the ring in the call in which it finishes;
 * episodes finishing in the same call enter in env-id order.
 *
 * Conventions: extern "C", POD only; tensor arguments are raw DEVICE pointers owned by the
 * caller unless stated; `stream` is a cudaStream_t as void*; 0 on success, <0 on error with
 * the message in bcp_last_error(); asynchronous unless stated.
 */
#ifndef BCP_REPLAY_H
#define BCP_REPLAY_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BCP_REPLAY_ABI_VERSION 1

/* error bits reported by bcp_replay_info (sticky) */
#define BCP_REPLAY_ERR_UNDERFLOW 1 /* the free-slot list ran dry: the reference raises IndexError
                                      ("pop from empty list", replay_memory.py:66,105)            */
#define BCP_REPLAY_ERR_STAGING 2   /* an env produced more transitions than max_episode_len        */
#define BCP_REPLAY_ERR_INDEX 4     /* bcp_replay_batch was given an index outside [0, size)        */
#define BCP_REPLAY_ERR_ROUNDS 8    /* internal: a commit needed more passes than were launched     */

typedef struct bcp_replay_config {
  int32_t abi_version;       /* must be BCP_REPLAY_ABI_VERSION */
  int32_t device_ordinal;
  int64_t buffer_size;       /* rows of the ring, replay_memory.py:12,14            */
  int64_t state_buffer_size; /* pool slots = int(buffer_size * load_factor), :31    */
  int32_t state_words;       /* prod(state_shape), :32; (R,2,7) => 14 R             */
  int32_t action_dim;        /* :21                                                 */
  int32_t num_envs;          /* envs feeding through begin/observe (0: add_episode only) */
  int32_t max_episode_len;   /* staging capacity per env (--max-episode-len)        */
} bcp_replay_config;

#define BCP_REPLAY_INFO_WORDS 10
/* info[0..9] = insert, full, size(), free slots, stats['>add_episode'], stats['>add'],
 *              stats['cache_evicted_s2'], error bits, episodes / transitions of the last commit */

typedef struct bcp_replay bcp_replay; /* opaque handle */

size_t bcp_replay_config_size(void);

/* ReplayMemory.__init__, replay_memory.py:12-41 */
int bcp_replay_create(const bcp_replay_config* cfg, bcp_replay** out);
int bcp_replay_destroy(bcp_replay* rp);

/* env.reset() of a batched loop: obs[e] becomes the initial state of a new episode of env e
 * (ddpg_cartpole.py:308-311).
 *   obs  : device f32[num_envs][state_words]
 *   mask : device u8[num_envs] or NULL (NULL = every env)                                 */
int bcp_replay_begin(bcp_replay* rp, const float* obs, const uint8_t* mask, void* stream);

/* after env.step(actions) -> (obs, reward, done): one transition per env whose episode is
 * open (ddpg_cartpole.py:316-322), then add_episode (:326, replay_memory.py:64-72) of every
 * episode that finished, in env-id order.
 *   actions : device f32[num_envs][action_dim], or i32[num_envs] when actions_are_int32
 *             (stored as float, action_dim must be 1)
 *   obs f32[num_envs][state_words], reward f32[num_envs], done u8[num_envs]
 *   reason  : bcp_step's reason_out or NULL.  BCP_REASON_RESETTING (4): no transition;
 *             BCP_REASON_FRESH (5): obs[e] is a reset observation => begins an episode.
 * An env whose episode already ended and was not begun again is skipped (step after done,
 * bullet_cartpole.py:179-181).                                                             */
int bcp_replay_observe(bcp_replay* rp, const void* actions, int32_t actions_are_int32, const float* obs,
                       const float* reward, const uint8_t* done, const uint8_t* reason, void* stream);

/* ReplayMemory.add_episode(initial_state, action_reward_state_sequence), replay_memory.py:64-72,
 * for callers that hold a whole episode.  Pointers may be device or host memory
 * (cudaMemcpyDefault); host memory must stay valid until the stream has run the call.
 *   initial_state f32[state_words], actions f32[n][action_dim], rewards f32[n],
 *   states2 f32[n][state_words], n >= 1                                                    */
int bcp_replay_add_episode(bcp_replay* rp, const float* initial_state, const float* actions, const float* rewards,
                           const float* states2, int32_t n, void* stream);

/* ReplayMemory.batch, replay_memory.py:131-138: gather of n rows.
 *   idxs : device i64[n] row indexes, or NULL: drawn in-kernel, uniform over [0, size())
 *          (random_indexes :123-129), Philox4x32-10 keyed by (seed, counter, i)
 *   out_state_1 / out_state_2 : device f16[n][state_words]; out_action f32[n][action_dim];
 *   out_reward f32[n]; out_mask f32[n]; out_idx i64[n] or NULL (the indexes used)
 * On an empty memory nothing is written (the reference returns empty arrays, :126-127).   */
int bcp_replay_batch(bcp_replay* rp, const int64_t* idxs, int32_t n, uint64_t seed, uint64_t counter, void* out_state_1,
                     float* out_action, float* out_reward, float* out_mask, void* out_state_2, int64_t* out_idx,
                     void* stream);

/* size / insert / full / stats (replay_memory.py:120-121,162-165) and error bits; host i64
 * [BCP_REPLAY_INFO_WORDS]; synchronises the stream.                                        */
int bcp_replay_info(bcp_replay* rp, int64_t* host_info, void* stream);

/* the internals, to HOST buffers (any may be NULL); synchronises.  Parity tests compare these
 * with the reference's arrays.
 *   state_1_idx i32[N], action f32[N][A], reward f32[N], terminal_mask f32[N], state_2_idx i32[N],
 *   state f16[M][W], free_slots i32[M] in list order (first info[3] entries are valid)        */
int bcp_replay_export(bcp_replay* rp, int32_t* state_1_idx, float* action, float* reward, float* terminal_mask,
                      int32_t* state_2_idx, void* state_f16, int32_t* free_slots, void* stream);

/* kernels enqueued so far (bench.py's gpu_launches) */
int64_t bcp_replay_launch_count(const bcp_replay* rp);

#ifdef __cplusplus
}
#endif
#endif /* BCP_REPLAY_H */
