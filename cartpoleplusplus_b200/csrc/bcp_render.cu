// bcp_render.cu — the raw-pixel observation (include/bcp_render.h): ground, cart and pole boxes ray-cast
// per pixel from pybullet's camera model, every (camera, repeat) image of an env in one pass.
// HBM-bound on the image it writes: 3*C*R halves per pixel, contiguous per thread, coalesced per warp
// ((H, W, 3, C, R) is the reference's own layout, bullet_cartpole.py:299-306).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

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

// slab test in the box frame; on a hit nearer than `best` (and inside [tlo, thi]) updates best / normal
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
// the renderer's uint8 channel as float16 / 255 (bullet_cartpole.py:289-294)
__device__ __forceinline__ __half to_pixel(float v) {
  const int u8 = min(255, max(0, (int)(v * 255.f)));
  return __float2half_rn((float)u8 / 255.f);
}
// What a pixel sees when neither cart nor pole is in the way does not depend on the env: its ray, the
// ground hit and the finished colour are worked out once per (camera, pixel) ...
__global__ void k_render_table(const Scene sc, float4* __restrict__ rays, float* __restrict__ depth, __half* __restrict__ base) {
  const int npix = sc.W * sc.H, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= sc.C * npix) return;
  const int c = i / npix, pix = i - c * npix;
  const int py = pix / sc.W, px = pix - py * sc.W;
  const float nx = (2.f * ((float)px + 0.5f)) / (float)sc.W - 1.f, ny = 1.f - (2.f * ((float)py + 0.5f)) / (float)sc.H;
  const Camera& cam = sc.cam[c];
  const float dx = nx * cam.tx, dy = ny * cam.ty;
  float d[3];
  for (int k = 0; k < 3; ++k) d[k] = cam.s[k] * dx + cam.u[k] * dy + cam.f[k];  // depth along f is the ray parameter
  Box ground;
  for (int k = 0; k < 3; ++k) ground.c[k] = sc.ground_pos[k];
  for (int k = 0; k < 9; ++k) ground.ax[k] = (k % 4 == 0) ? 1.f : 0.f;
  float best = sc.far_val, n[3] = {0.f, 0.f, 0.f};
  const bool hit = ray_box(cam.eye, d, ground, sc.h[0], sc.near_val, sc.far_val, best, n);
  rays[i] = make_float4(d[0], d[1], d[2], 1.f / (d[0] * d[0] + d[1] * d[1] + d[2] * d[2]));
  depth[i] = best;
  const float shade = hit ? sc.ambient + sc.diffuse * fmaxf(0.f, n[0] * sc.light[0] + n[1] * sc.light[1] + n[2] * sc.light[2]) : 1.f;
  for (int ch = 0; ch < 3; ++ch) base[(size_t)i * 3 + ch] = to_pixel(hit ? sc.rgb[0][ch] * shade : sc.bg[ch]);
}

// ... and the per-env kernel only tests the two moving boxes, behind a screen-space bounding rectangle per
// (camera, repeat, box): the corners are projected once per env, most pixels fail four integer compares and
// take the table's colour.  One block per env; after the set-up its warps are independent: a warp renders 32
// consecutive pixels at a time and writes their 32 * 3*C*R halves, contiguous in the reference's layout,
// with the widest stores the alignment allows.
constexpr int kWarps = kThreads / 32;
struct Rect { int x0, x1, y0, y1; };
__global__ void __launch_bounds__(kThreads) k_render(const Scene sc, const float* __restrict__ obs, const float4* __restrict__ rays,
                                                     const float* __restrict__ depth, const __half* __restrict__ base,
                                                     __half* __restrict__ out) {
  __shared__ Box box[kMaxR][2];
  __shared__ Rect rect[2][kMaxR][2];  // [camera][repeat][cart, pole]
  __shared__ __align__(16) __half stage_all[kWarps][32 * 3 * 2 * kMaxR];
  const int env = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int npix = sc.W * sc.H, CR = sc.C * sc.R;
  if (threadIdx.x < sc.R * 2) {
    const int r = threadIdx.x >> 1, body = threadIdx.x & 1;
    const float* p = obs + (((size_t)env * sc.R + r) * 2 + body) * 7;
    Box& b = box[r][body];
    b.c[0] = p[0]; b.c[1] = p[1]; b.c[2] = p[2];
    quat_axes(p + 3, b.ax);
  }
  __syncthreads();
  if (threadIdx.x < sc.C * sc.R * 2) {  // screen rectangle of one box in one image: its eight corners projected
    const int c = threadIdx.x / (sc.R * 2), rb = threadIdx.x - c * sc.R * 2, r = rb >> 1, body = rb & 1;
    const Camera& cam = sc.cam[c];
    const Box& b = box[r][body];
    const float* h = sc.h[1 + body];
    float xmin = 1e30f, xmax = -1e30f, ymin = 1e30f, ymax = -1e30f;
    bool all = false;
    for (int k = 0; k < 8; ++k) {
      const float sx = (k & 1) ? h[0] : -h[0], sy = (k & 2) ? h[1] : -h[1], sz = (k & 4) ? h[2] : -h[2];
      float p[3];
      for (int i = 0; i < 3; ++i) p[i] = b.c[i] + sx * b.ax[i] + sy * b.ax[3 + i] + sz * b.ax[6 + i] - cam.eye[i];
      const float z = p[0] * cam.f[0] + p[1] * cam.f[1] + p[2] * cam.f[2];
      if (!(z > 1e-3f)) { all = true; break; }  // a corner at or behind the eye: no useful bound
      const float x = (p[0] * cam.s[0] + p[1] * cam.s[1] + p[2] * cam.s[2]) / (z * cam.tx);
      const float y = (p[0] * cam.u[0] + p[1] * cam.u[1] + p[2] * cam.u[2]) / (z * cam.ty);
      const float fx = (x + 1.f) * 0.5f * (float)sc.W - 0.5f, fy = (1.f - y) * 0.5f * (float)sc.H - 0.5f;  // pixel-centre coordinates
      xmin = fminf(xmin, fx); xmax = fmaxf(xmax, fx); ymin = fminf(ymin, fy); ymax = fmaxf(ymax, fy);
    }
    Rect q;
    if (all || !(xmax - xmin < 1e6f)) { q.x0 = 0; q.x1 = sc.W - 1; q.y0 = 0; q.y1 = sc.H - 1; }
    else {  // one pixel of slack against rounding
      q.x0 = (int)floorf(fmaxf(xmin, -2.f)) - 1; q.x1 = (int)ceilf(fminf(xmax, (float)sc.W + 2.f)) + 1;
      q.y0 = (int)floorf(fmaxf(ymin, -2.f)) - 1; q.y1 = (int)ceilf(fminf(ymax, (float)sc.H + 2.f)) + 1;
    }
    rect[c][r][body] = q;
  }
  __syncthreads();
  __half* stage = stage_all[warp];
  for (int pix0 = warp * 32; pix0 < npix; pix0 += kThreads) {
    const int pix = pix0 + lane;
    if (pix < npix) {
      const int py = pix / sc.W, px = pix - py * sc.W;
      __half* dst = stage + (size_t)lane * 3 * CR;
      for (int c = 0; c < sc.C; ++c) {
        const Camera& cam = sc.cam[c];
        const __half* bc = base + (size_t)(c * npix + pix) * 3;
        const __half b0 = bc[0], b1 = bc[1], b2 = bc[2];
        for (int r = 0; r < sc.R; ++r) {
          __half v0 = b0, v1 = b1, v2 = b2;
          const Rect q0 = rect[c][r][0], q1 = rect[c][r][1];
          const bool in0 = px >= q0.x0 && px <= q0.x1 && py >= q0.y0 && py <= q0.y1;
          const bool in1 = px >= q1.x0 && px <= q1.x1 && py >= q1.y0 && py <= q1.y1;
          if (in0 || in1) {
            const float4 ray = rays[c * npix + pix];
            const float d[3] = {ray.x, ray.y, ray.z};
            float best = depth[c * npix + pix], n[3] = {0.f, 0.f, 0.f};
            int body = 0;
            if (in0 && ray_box(cam.eye, d, box[r][0], sc.h[1], sc.near_val, sc.far_val, best, n)) body = 1;
            if (in1 && ray_box(cam.eye, d, box[r][1], sc.h[2], sc.near_val, sc.far_val, best, n)) body = 2;
            if (body) {
              const float shade = sc.ambient + sc.diffuse * fmaxf(0.f, n[0] * sc.light[0] + n[1] * sc.light[1] + n[2] * sc.light[2]);
              v0 = to_pixel(sc.rgb[body][0] * shade); v1 = to_pixel(sc.rgb[body][1] * shade); v2 = to_pixel(sc.rgb[body][2] * shade);
            }
          }
          dst[c * sc.R + r] = v0; dst[CR + c * sc.R + r] = v1; dst[2 * CR + c * sc.R + r] = v2;
        }
      }
    }
    __syncwarp();
    const int live = min(32, npix - pix0);
    const unsigned nbytes = (unsigned)(live * 3 * CR * sizeof(__half));
    char* g = reinterpret_cast<char*>(out + ((size_t)env * npix + pix0) * 3 * CR);
    const char* sm = reinterpret_cast<const char*>(stage);
    const size_t al = reinterpret_cast<size_t>(g) | nbytes;
    if ((al & 15) == 0) {
      for (unsigned i = lane; i < nbytes / 16; i += 32) reinterpret_cast<uint4*>(g)[i] = reinterpret_cast<const uint4*>(sm)[i];
    } else if ((al & 3) == 0) {
      for (unsigned i = lane; i < nbytes / 4; i += 32) reinterpret_cast<uint32_t*>(g)[i] = reinterpret_cast<const uint32_t*>(sm)[i];
    } else {
      for (unsigned i = lane; i < nbytes / 2; i += 32) reinterpret_cast<uint16_t*>(g)[i] = reinterpret_cast<const uint16_t*>(sm)[i];
    }
    __syncwarp();  // the slice is reused by the warp's next 32 pixels
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
  cudaStream_t st = (cudaStream_t)stream;
  const size_t nt = (size_t)sc.C * a->width * a->height;
  // per-call scratch from the stream-ordered pool: rays f32x4 | ground depth f32 | base colour f16x3
  char* scratch = nullptr;
  {  // keep freed scratch in the device's default pool across synchronisations (the default gives it back to the driver)
    static bool pool_set[64] = {false};
    if (device_ordinal >= 0 && device_ordinal < 64 && !pool_set[device_ordinal]) {
      cudaMemPool_t pool;
      if (cudaDeviceGetDefaultMemPool(&pool, device_ordinal) == cudaSuccess) {
        unsigned long long keep = 64ull << 20;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
      }
      pool_set[device_ordinal] = true;
    }
  }
  cudaError_t e0 = cudaMallocAsync((void**)&scratch, nt * (16 + 4 + 8), st);
  if (e0 != cudaSuccess) {
    if (prev != device_ordinal) cudaSetDevice(prev);
    return bcp::abi_fail("bcp_render_pixels: scratch allocation", e0);
  }
  float4* rays = reinterpret_cast<float4*>(scratch);
  float* depth = reinterpret_cast<float*>(scratch + nt * 16);
  __half* base = reinterpret_cast<__half*>(scratch + nt * 20);
  k_render_table<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(sc, rays, depth, base);
  const dim3 grid((unsigned)a->num_envs, 1);
  k_render<<<grid, kThreads, 0, st>>>(sc, a->obs, rays, depth, base, reinterpret_cast<__half*>(a->pixels));
  cudaFreeAsync(scratch, st);
  const cudaError_t e = cudaGetLastError();
  if (prev != device_ordinal) cudaSetDevice(prev);
  if (e != cudaSuccess) return bcp::abi_fail("bcp_render_pixels: launch", e);
  return 0;
}

}  // extern "C"
