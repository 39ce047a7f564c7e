// bcp_render.cu — the raw-pixel observation (include/bcp_render.h): ground, cart and pole boxes ray-cast
// per pixel from pybullet's camera model, every (camera, repeat) image of an env in one pass.
// HBM-bound on the image it writes: 3*C*R halves per pixel, contiguous per thread, coalesced per warp
// ((H, W, 3, C, R) is the reference's own layout, bullet_cartpole.py:299-306).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>

#include "../../include/bcp_render.h"

namespace bcp {
int abi_fail(const char* what, cudaError_t e);
}

namespace {

constexpr int kMaxR = 8;       // action repeats held in shared memory
constexpr int kThreads = 256;

struct Camera {  // pybullet's computeViewMatrix + computeProjectionMatrixFOV, as a ray generator
  float eye[3], s[3], u[3], f[3];  // right, up, forward
  float tx, ty;                    // tan(fov / 2) * aspect, tan(fov / 2)
};
struct Scene {
  int B, R, C, W, H;
  float h[BCP_RENDER_BODIES][3], rgb[BCP_RENDER_BODIES][3], ground_pos[3];
  float near_val, far_val, light[3], ambient, diffuse, bg[3];
  Camera cam[2];
};
struct Box {
  float c[3], ax[9];  // centre, the three axes (columns of the rotation) one after the other
};

__device__ __forceinline__ void quat_axes(const float* q, float* ax) {  // q = x, y, z, w
  const float x = q[0], y = q[1], z = q[2], w = q[3];
  ax[0] = 1.f - 2.f * (y * y + z * z); ax[1] = 2.f * (x * y + w * z); ax[2] = 2.f * (x * z - w * y);
  ax[3] = 2.f * (x * y - w * z); ax[4] = 1.f - 2.f * (x * x + z * z); ax[5] = 2.f * (y * z + w * x);
  ax[6] = 2.f * (x * z + w * y); ax[7] = 2.f * (y * z - w * x); ax[8] = 1.f - 2.f * (x * x + y * y);
}

// slab test in the box frame; on a hit nearer than `best` (and inside [near, far]) updates best / normal
__device__ __forceinline__ bool ray_box(const float* o, const float* d, const Box& b, const float* h, float tlo, float thi,
                                        float& best, float* n) {
  float rel[3] = {o[0] - b.c[0], o[1] - b.c[1], o[2] - b.c[2]};
  float t0 = tlo, t1 = thi;
  int axis = -1;
  float sign = 0.f;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const float* e = b.ax + 3 * a;
    const float lo = rel[0] * e[0] + rel[1] * e[1] + rel[2] * e[2];
    const float ld = d[0] * e[0] + d[1] * e[1] + d[2] * e[2];
    if (fabsf(ld) < 1e-12f) {
      if (fabsf(lo) > h[a]) return false;
      continue;
    }
    const float inv = 1.f / ld;
    float ta = (-h[a] - lo) * inv, tb = (h[a] - lo) * inv, sg = -1.f;  // entering through the -a face ...
    if (ta > tb) { const float t = ta; ta = tb; tb = t; sg = 1.f; }    // ... or through the +a face
    if (ta > t0) { t0 = ta; axis = a; sign = sg; }
    if (tb < t1) t1 = tb;
    if (t0 > t1) return false;
  }
  if (axis < 0 || t0 >= best) return false;  // the eye is never inside a box
  best = t0;
  n[0] = sign * b.ax[3 * axis]; n[1] = sign * b.ax[3 * axis + 1]; n[2] = sign * b.ax[3 * axis + 2];
  return true;
}

__global__ void __launch_bounds__(kThreads) k_render(const Scene sc, const float* __restrict__ obs, __half* __restrict__ out) {
  __shared__ Box box[kMaxR][2];
  __shared__ Box ground;
  const int env = blockIdx.x;
  if (threadIdx.x < sc.R * 2) {
    const int r = threadIdx.x >> 1, body = threadIdx.x & 1;
    const float* p = obs + (((size_t)env * sc.R + r) * 2 + body) * 7;
    Box& b = box[r][body];
    b.c[0] = p[0]; b.c[1] = p[1]; b.c[2] = p[2];
    quat_axes(p + 3, b.ax);
  }
  if (threadIdx.x == 32) {
    for (int i = 0; i < 3; ++i) ground.c[i] = sc.ground_pos[i];
    for (int i = 0; i < 9; ++i) ground.ax[i] = (i % 4 == 0) ? 1.f : 0.f;
  }
  __syncthreads();
  const int pix = blockIdx.y * kThreads + threadIdx.x;
  if (pix >= sc.W * sc.H) return;
  const int py = pix / sc.W, px = pix - py * sc.W;
  const float nx = (2.f * ((float)px + 0.5f)) / (float)sc.W - 1.f, ny = 1.f - (2.f * ((float)py + 0.5f)) / (float)sc.H;
  const int CR = sc.C * sc.R;
  __half* dst = out + ((size_t)env * sc.W * sc.H + pix) * 3 * CR;
  for (int c = 0; c < sc.C; ++c) {
    const Camera& cam = sc.cam[c];
    const float dx = nx * cam.tx, dy = ny * cam.ty;
    float d[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) d[i] = cam.s[i] * dx + cam.u[i] * dy + cam.f[i];  // depth along f is the ray parameter
    float gbest = sc.far_val, gn[3] = {0.f, 0.f, 0.f};
    const bool ghit = ray_box(cam.eye, d, ground, sc.h[0], sc.near_val, sc.far_val, gbest, gn);
    for (int r = 0; r < sc.R; ++r) {
      float best = gbest, n[3] = {gn[0], gn[1], gn[2]};
      int body = ghit ? 0 : -1;
      if (ray_box(cam.eye, d, box[r][0], sc.h[1], sc.near_val, sc.far_val, best, n)) body = 1;
      if (ray_box(cam.eye, d, box[r][1], sc.h[2], sc.near_val, sc.far_val, best, n)) body = 2;
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) {
        float v;
        if (body < 0) {
          v = sc.bg[ch];
        } else {
          const float ndl = fmaxf(0.f, n[0] * sc.light[0] + n[1] * sc.light[1] + n[2] * sc.light[2]);
          v = sc.rgb[body][ch] * (sc.ambient + sc.diffuse * ndl);
        }
        const int u8 = min(255, max(0, (int)(v * 255.f)));  // the renderer's uint8 image ...
        dst[ch * CR + c * sc.R + r] = __float2half_rn((float)u8 / 255.f);  // ... as float16 / 255 (:289-294)
      }
    }
  }
}

void norm3(float* v) {
  const float n = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  for (int i = 0; i < 3; ++i) v[i] /= n;
}
void cross3(const float* a, const float* b, float* o) {
  o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}

}  // namespace

extern "C" {

size_t bcp_render_args_size(void) { return sizeof(bcp_render_args); }

int bcp_render_pixels(const bcp_render_args* a, int device_ordinal, void* stream) {
  if (!a || !a->obs || !a->pixels) return bcp::abi_fail("bcp_render_pixels: null argument", cudaSuccess);
  if (a->num_envs <= 0 || a->width <= 0 || a->height <= 0) return bcp::abi_fail("bcp_render_pixels: empty batch or image", cudaSuccess);
  if (a->num_cameras != 1 && a->num_cameras != 2) return bcp::abi_fail("bcp_render_pixels: num_cameras must be 1 or 2", cudaSuccess);
  if (a->repeats < 1 || a->repeats > kMaxR) return bcp::abi_fail("bcp_render_pixels: repeats must be 1..8", cudaSuccess);
  if (!(a->near_val > 0.f) || !(a->far_val > a->near_val) || !(a->fov_deg > 0.f && a->fov_deg < 180.f))
    return bcp::abi_fail("bcp_render_pixels: bad near / far / fov", cudaSuccess);
  Scene sc;
  sc.B = a->num_envs; sc.R = a->repeats; sc.C = a->num_cameras; sc.W = a->width; sc.H = a->height;
  for (int b = 0; b < BCP_RENDER_BODIES; ++b)
    for (int i = 0; i < 3; ++i) { sc.h[b][i] = a->half_extents[b][i]; sc.rgb[b][i] = a->rgb[b][i]; }
  for (int i = 0; i < 3; ++i) { sc.ground_pos[i] = a->ground_pos[i]; sc.light[i] = a->light_dir[i]; sc.bg[i] = a->background[i]; }
  norm3(sc.light);
  sc.near_val = a->clip ? a->near_val : 1e-6f; sc.far_val = a->clip ? a->far_val : 1e30f; sc.ambient = a->ambient; sc.diffuse = a->diffuse;
  const float ty = tanf(0.5f * a->fov_deg * 3.14159265358979323846f / 180.f);
  const float aspect = (float)a->width / (float)a->height;  // p.renderImage's legacy form: aspect of the image
  for (int c = 0; c < sc.C; ++c) {
    Camera& cam = sc.cam[c];
    float up[3] = {a->up[0], a->up[1], a->up[2]};
    for (int i = 0; i < 3; ++i) { cam.eye[i] = a->camera_pos[c][i]; cam.f[i] = a->target[i] - a->camera_pos[c][i]; }
    norm3(cam.f);
    norm3(up);
    cross3(cam.f, up, cam.s);
    norm3(cam.s);
    cross3(cam.s, cam.f, cam.u);
    cam.tx = ty * aspect;
    cam.ty = ty;
  }
  int prev = -1;
  cudaGetDevice(&prev);
  if (prev != device_ordinal) cudaSetDevice(device_ordinal);
  const dim3 grid((unsigned)a->num_envs, (unsigned)((a->width * a->height + kThreads - 1) / kThreads));
  k_render<<<grid, kThreads, 0, (cudaStream_t)stream>>>(sc, a->obs, reinterpret_cast<__half*>(a->pixels));
  const cudaError_t e = cudaGetLastError();
  if (prev != device_ordinal) cudaSetDevice(prev);
  if (e != cudaSuccess) return bcp::abi_fail("bcp_render_pixels: launch", e);
  return 0;
}

}  // extern "C"
