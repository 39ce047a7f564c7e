/* bcp_render.h — the raw-pixel observation as a CUDA kernel.
 *
 * Replaces bullet_cartpole.py:277-296 (render_rgb: p.renderImage at render_width x render_height from one
 * of two fixed cameras, uint8 RGBA -> float16 / 255) and :299-306 (state[:, :, :, camera, repeat]) for a
 * whole batch: one call (a small per-pixel table kernel, then one block per env) turns the low-dimensional
 * observation of a step - the (R, 2, 7) cart and pole poses per env that bcp_step / bcp_reset wrote - into
 * the (H, W, 3, C, R) float16 image block per env.
 * No window, no graphics API: three boxes (ground, cart, pole) are ray-cast per pixel from pybullet's
 * camera model (computeViewMatrix + computeProjectionMatrixFOV, restated in DESIGN.md section 12).
 *
 * Parity: pybullet's renderer (TinyRenderer behind p.renderImage) is unavailable on this box, so the
 * reference's pixels cannot be produced; the camera geometry follows pybullet's published matrices,
 * the shading model (flat Lambert: ambient + diffuse * max(0, n.l), opaque, no shadows) and its
 * constants are parameters of this call whose defaults are stated, not verified, against pybullet.
 */
#ifndef BCP_RENDER_H
#define BCP_RENDER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BCP_RENDER_BODIES 3 /* ground, cart, pole */

typedef struct bcp_render_args {
  int32_t num_envs, repeats, num_cameras; /* C = 1 or 2 (bullet_cartpole.py:113-114) */
  int32_t width, height;                  /* --render-width / --render-height, 50 x 50 by default */
  float half_extents[BCP_RENDER_BODIES][3];
  float rgb[BCP_RENDER_BODIES][3];        /* the URDFs' <material><color rgba>, alpha ignored */
  float ground_pos[3];                    /* the ground box does not move */
  float camera_pos[2][3];                 /* (0, .75, .75) and (.75, 0, .75)   :279-280 */
  float target[3], up[3];                 /* (0, 0, .3), (0, 0, 1)             :281-282 */
  float near_val, far_val, fov_deg;       /* 1, 20, 30                         :283-284 */
  int32_t clip;                           /* 0: nothing is clipped against near / far, like pybullet's software
                                             rasteriser, which only z-tests - the reference's near plane (1 m) lies
                                             BEHIND its cart and pole (0.875 m from eye to target), a clipping renderer
                                             would show an empty ground; 1: hits outside [near, far] are dropped */
  float light_dir[3];                     /* unit vector from the scene towards the light */
  float ambient, diffuse;                 /* 0.6, 0.35 */
  float background[3];                    /* 1, 1, 1 */
  const float* obs;                       /* DEVICE f32[num_envs][R][2][7]: cart pose, pole pose (x,y,z, qx,qy,qz,qw) */
  void* pixels;                           /* DEVICE f16[num_envs][H][W][3][C][R] */
} bcp_render_args;

size_t bcp_render_args_size(void);

/* Enqueued on `stream` of device `device_ordinal` (scratch comes from the device's stream-ordered pool).
 * Returns 0, or -1 with bcp_last_error() set. */
int bcp_render_pixels(const bcp_render_args* args, int device_ordinal, void* stream);

/* bcp_render_pixels for callers that may not allocate.  NO counterpart in the reference: added
 * for code under CUDA graph capture, tight serving loops and harnesses
 * that forbid cudaMalloc.  Same arguments, same two kernels, bit-identical pixels; the per-call scratch
 * (ray table | ground depth | background image) lives in a file-scope `static __device__` array of
 * BCP_RENDER_STATIC_SCRATCH_BYTES instead of the stream-ordered pool, and the call path contains no
 * cudaMalloc / cudaMallocAsync / cudaFree or any other allocating routine (the first call on a device
 * only looks the array's device address up with cudaGetSymbolAddress).
 *   limit : 20 * C*W*H (rounded up to 16) + 6 * C*W*H*R bytes must fit in the array - e.g. the
 *           reference's 50 x 50 with C = 2, R = 8 needs 340 000 bytes, 128 x 128 with C = 2, R = 2 needs
 *           1 048 576; anything larger is refused with -1 and a message naming both sizes.
 *   cost  : every GPU that loads libbcp.so holds the array until the process exits.
 *   races : all calls on a device share the array, which keeps the last call's contents (nothing needs to
 *           start at zero: the table kernel rewrites every word the render kernel reads).  Calls on the
 *           same stream are ordered; calls that overlap on the GPU (different streams or host threads
 *           without an event between them) race on it - serialise them or use bcp_render_pixels.      */
#define BCP_RENDER_STATIC_SCRATCH_BYTES (1u << 20)
size_t bcp_render_static_scratch_bytes(void);
int bcp_render_pixels_static(const bcp_render_args* args, int device_ordinal, void* stream);

#ifdef __cplusplus
}
#endif
#endif
