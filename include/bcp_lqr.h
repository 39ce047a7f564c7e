/* bcp_lqr.h — the fork's 8-state observation and LQR feedback for a batch of envs.
 *
 * Replaces, for a whole batch resident in HBM, what the fork's own env class does per cart-pole
 * with four pybullet queries and a numpy dot product:
 *   random_action_agent.py:121-135  _observe_state   pole pose / Euler angles / velocity -> 8 floats
 *   random_action_agent.py:108-119  _check_done      3.0 m / pi/4 thresholds on that state
 *   random_action_agent.py:92-95    lqr_control_forces   -(K . (state - zero_point))
 *   random_action_agent.py:888-903  the loop's  disturbance + control  action
 * One kernel reads the pole's body out of every env record (13 scalars) and writes state, action
 * and done flag; nothing here is on the path of bcp_step.
 *
 * State layout (random_action_agent.py:55-62):
 *   0 X  pole x - x0    1 X_DOT  pole vx    2 Y  pole y    3 Y_DOT  pole vy
 *   4 ROLL              5 ROLL_DOT  wx      6 PITCH        7 PITCH_DOT  wy
 * Roll and pitch are pybullet's getEulerFromQuaternion of the pole's orientation (pybullet is a
 * third-party dependency that is absent from the reference tree; its published formula is
 * restated in DESIGN.md section 11, parity against pybullet itself is unpinned).
 */
#ifndef BCP_LQR_H
#define BCP_LQR_H

#include <stddef.h>
#include <stdint.h>

#include "bcp.h"

#ifdef __cplusplus
extern "C" {
#endif

#define BCP_LQR_STATE 8
#define BCP_LQR_ACTIONS 2

/* done codes written to done_out (random_action_agent.py:108-119) */
#define BCP_LQR_RUNNING 0
#define BCP_LQR_POSITION 1 /* 'cart position bounds exceeded'  |X| or |Y| > position_threshold  */
#define BCP_LQR_ANGLE 2    /* 'pole angle bounds exceeded'     |ROLL| or |PITCH| > angle_threshold */

typedef struct bcp_lqr_args {
  /* inputs, DEVICE pointers, any of them may be NULL */
  const float* gains;       /* f32[G][2][8] feedback matrices, G = 1 (shared) or num_envs; NULL: no control */
  int32_t gains_per_env;    /* 0: one matrix for every env, 1: one per env */
  const float* zero_point;  /* f32[8] the state the controller holds (NULL: zeros)                 */
  const float* x0;          /* f32[num_envs] initial pole x of each env (NULL: 0)                  */
  const float* disturbance; /* f32[D][2] force added to the control, D = 1 or num_envs (NULL: 0)  */
  int32_t disturbance_per_env;
  float position_threshold; /* the fork's game uses 3.0      (random_action_agent.py:694) */
  float angle_threshold;    /* ... and pi/4                  (random_action_agent.py:695) */
  /* outputs, DEVICE pointers, any of them may be NULL */
  float* state_out;   /* f32[num_envs][8]                                                    */
  float* action_out;  /* f32[num_envs][2] disturbance - K.(state - zero_point): the next bcp_step's
                         continuous action when the env was created with action_force = 1      */
  uint8_t* done_out;  /* u8[num_envs] BCP_LQR_* */
} bcp_lqr_args;

size_t bcp_lqr_args_size(void);

/* One launch over every env of the handle, enqueued on `stream`. */
int bcp_lqr_control(bcp_env* env, const bcp_lqr_args* args, void* stream);

#ifdef __cplusplus
}
#endif
#endif
