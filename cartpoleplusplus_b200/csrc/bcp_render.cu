// bcp_render.cu — the raw-pixel observation (include/bcp_render.h): ground, cart and pole boxes ray-cast
// per pixel from pybullet's camera model, every (camera, repeat) image of an env in one pass.
// HBM-bound on the image it writes: 3*C*R halves per pixel, contiguous per thread, coalesced per warp
// ((H, W, 3, C, R) is the reference's own layout, bullet_cartpole.py:299-306).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

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
// ground hit and the finished colour are worked out once per (camera, pixel); `img` is that background
// already in the reference's layout, [pixel][3][C][R] ...
__global__ void k_render_table(const Scene sc, float4* __restrict__ rays, float* __restrict__ depth, __half* __restrict__ img) {
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
  rays[i] = make_float4(d[0], d[1], d[2], 0.f);
  depth[i] = best;
  const float shade = hit ? sc.ambient + sc.diffuse * fmaxf(0.f, n[0] * sc.light[0] + n[1] * sc.light[1] + n[2] * sc.light[2]) : 1.f;
  const int CR = sc.C * sc.R;
  for (int ch = 0; ch < 3; ++ch) {
    const __half v = to_pixel(hit ? sc.rgb[0][ch] * shade : sc.bg[ch]);
    for (int r = 0; r < sc.R; ++r) img[((size_t)pix * 3 + ch) * CR + c * sc.R + r] = v;
  }
}

// ... and the per-env kernel, one block per env, (1) streams that background into the env's block with the
// widest stores the alignment allows and (2) re-renders only the pixels inside the screen rectangles of the
// cart and the pole of each (camera, repeat) - the eight corners of a box projected once - as a flat list of
// (rectangle, pixel) tasks spread over the block, so that the few hundred rays that can hit something do not
// drag whole warps of background pixels through the intersection code.
struct Rect { int x0, y0, w, h, c, r; float inv_w; };  // inv_w: row of a task without an integer division
constexpr int kMaxRects = 2 * kMaxR * 2;
__global__ void __launch_bounds__(kThreads) k_render(const Scene sc, const float* __restrict__ obs, const float4* __restrict__ rays,
                                                     const float* __restrict__ depth, const __half* __restrict__ img,
                                                     __half* __restrict__ out) {
  __shared__ Box box[kMaxR][2];
  __shared__ Rect rect[kMaxRects];      // index (c * R + r) * 2 + body
  __shared__ int first[kMaxRects + 1];  // prefix of the rectangles' pixel counts
  const int env = blockIdx.x;
  const int npix = sc.W * sc.H, CR = sc.C * sc.R, nrect = CR * 2;
  if (threadIdx.x < sc.R * 2) {
    const int r = threadIdx.x >> 1, body = threadIdx.x & 1;
    const float* p = obs + (((size_t)env * sc.R + r) * 2 + body) * 7;
    Box& b = box[r][body];
    b.c[0] = p[0]; b.c[1] = p[1]; b.c[2] = p[2];
    quat_axes(p + 3, b.ax);
  }
  // (1) the background, while the poses arrive
  {
    const size_t nbytes = (size_t)npix * 3 * CR * sizeof(__half);
    char* g = reinterpret_cast<char*>(out + (size_t)env * npix * 3 * CR);
    const char* src = reinterpret_cast<const char*>(img);
    const size_t al = reinterpret_cast<size_t>(g) | reinterpret_cast<size_t>(src) | nbytes;
    if ((al & 15) == 0) {
      for (size_t i = threadIdx.x; i < nbytes / 16; i += kThreads) reinterpret_cast<uint4*>(g)[i] = reinterpret_cast<const uint4*>(src)[i];
    } else if ((al & 3) == 0) {
      for (size_t i = threadIdx.x; i < nbytes / 4; i += kThreads) reinterpret_cast<uint32_t*>(g)[i] = reinterpret_cast<const uint32_t*>(src)[i];
    } else {
      for (size_t i = threadIdx.x; i < nbytes / 2; i += kThreads) reinterpret_cast<uint16_t*>(g)[i] = reinterpret_cast<const uint16_t*>(src)[i];
    }
  }
  __syncthreads();
  // screen rectangle of one box in one image: its eight corners projected, one corner per lane, four
  // rectangles per warp at a time
  {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, corner = lane & 7;
    for (int k0 = warp * 4; k0 < nrect; k0 += kThreads / 32 * 4) {
      const int k = k0 + (lane >> 3), kk = min(k, nrect - 1);
      const int cr = kk >> 1, body = kk & 1, c = cr / sc.R, r = cr - c * sc.R;
      const Camera& cam = sc.cam[c];
      const Box& b = box[r][body];
      const float* h = sc.h[1 + body];
      const float sx = (corner & 1) ? h[0] : -h[0], sy = (corner & 2) ? h[1] : -h[1], sz = (corner & 4) ? h[2] : -h[2];
      float p[3];
      for (int i = 0; i < 3; ++i) p[i] = b.c[i] + sx * b.ax[i] + sy * b.ax[3 + i] + sz * b.ax[6 + i] - cam.eye[i];
      const float z = p[0] * cam.f[0] + p[1] * cam.f[1] + p[2] * cam.f[2];
      const bool bad = !(z > 1e-3f);  // a corner at or behind the eye: no useful bound
      const float zz = bad ? 1.f : z;
      const float x = (p[0] * cam.s[0] + p[1] * cam.s[1] + p[2] * cam.s[2]) / (zz * cam.tx);
      const float y = (p[0] * cam.u[0] + p[1] * cam.u[1] + p[2] * cam.u[2]) / (zz * cam.ty);
      const float fx = (x + 1.f) * 0.5f * (float)sc.W - 0.5f, fy = (1.f - y) * 0.5f * (float)sc.H - 0.5f;  // pixel-centre coordinates
      float xmin = fx, xmax = fx, ymin = fy, ymax = fy;
      int all = bad ? 1 : 0;
#pragma unroll
      for (int o = 1; o < 8; o <<= 1) {
        xmin = fminf(xmin, __shfl_xor_sync(0xffffffffu, xmin, o)); xmax = fmaxf(xmax, __shfl_xor_sync(0xffffffffu, xmax, o));
        ymin = fminf(ymin, __shfl_xor_sync(0xffffffffu, ymin, o)); ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
        all |= __shfl_xor_sync(0xffffffffu, all, o);
      }
      if (corner == 0 && k < nrect) {
        int x0 = 0, x1 = sc.W - 1, y0 = 0, y1 = sc.H - 1;
        if (!all && xmax - xmin < 1e6f && ymax - ymin < 1e6f) {
          // pixel centres inside the projected hull's bounds (a convex box projects inside the bounds of its corners);
          // 0.05 pixel of slack against rounding, clamped to the image
          x0 = max(0, (int)ceilf(fmaxf(xmin, -2.f) - 0.05f)); x1 = min(sc.W - 1, (int)floorf(fminf(xmax, (float)sc.W + 2.f) + 0.05f));
          y0 = max(0, (int)ceilf(fmaxf(ymin, -2.f) - 0.05f)); y1 = min(sc.H - 1, (int)floorf(fminf(ymax, (float)sc.H + 2.f) + 0.05f));
        }
        Rect q;
        q.x0 = x0; q.y0 = y0; q.w = max(0, x1 - x0 + 1); q.h = max(0, y1 - y0 + 1);
        q.c = c; q.r = r; q.inv_w = 1.f / (float)max(1, q.w);
        rect[k] = q;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < 32) {  // exclusive prefix of the rectangles' pixel counts (nrect <= 32): one warp scan
    const int lane = threadIdx.x;
    const int area = lane < nrect ? rect[lane].w * rect[lane].h : 0;
    int inc = area;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += v;
    }
    if (lane < nrect) first[lane] = inc - area;
    if (lane == nrect - 1) first[nrect] = inc;
  }
  __syncthreads();
  // (2) the pixels a box can cover
  const int total = first[nrect];
  for (int t = threadIdx.x; t < total; t += kThreads) {
    int k = 0;
    while (t >= first[k + 1]) ++k;
    const Rect q = rect[k];
    const int local = t - first[k];
    int ly = (int)(((float)local + 0.5f) * q.inv_w);  // local < 2^15: exact up to the rounding of inv_w, fixed below
    ly -= (ly * q.w > local);
    ly += ((ly + 1) * q.w <= local);
    const int px = q.x0 + (local - ly * q.w), py = q.y0 + ly;
    const int c = q.c, r = q.r, mine = k & 1;
    const Rect o = rect[k ^ 1];  // the other box of the same image
    const bool in_other = px >= o.x0 && px < o.x0 + o.w && py >= o.y0 && py < o.y0 + o.h;
    if (in_other && mine == 1) continue;  // the cart's task renders the shared pixel
    const Camera& cam = sc.cam[c];
    const int pix = py * sc.W + px;
    const float4 ray = rays[c * npix + pix];
    const float d[3] = {ray.x, ray.y, ray.z};
    float best = depth[c * npix + pix], n[3] = {0.f, 0.f, 0.f};
    int body = 0;
    if ((mine == 0 || in_other) && ray_box(cam.eye, d, box[r][0], sc.h[1], sc.near_val, sc.far_val, best, n)) body = 1;
    if ((mine == 1 || in_other) && ray_box(cam.eye, d, box[r][1], sc.h[2], sc.near_val, sc.far_val, best, n)) body = 2;
    if (!body) continue;  // the background is already there
    const float shade = sc.ambient + sc.diffuse * fmaxf(0.f, n[0] * sc.light[0] + n[1] * sc.light[1] + n[2] * sc.light[2]);
    __half* dst = out + ((size_t)env * npix + pix) * 3 * CR + c * sc.R + r;
    dst[0] = to_pixel(sc.rgb[body][0] * shade); dst[CR] = to_pixel(sc.rgb[body][1] * shade); dst[2 * CR] = to_pixel(sc.rgb[body][2] * shade);
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
  // per-call scratch from the stream-ordered pool: rays f32x4 | ground depth f32 | background image f16
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
  cudaError_t e0 = cudaMallocAsync((void**)&scratch, nt * (16 + 4) + 16 + nt * 3 * (size_t)sc.R * sizeof(__half), st);
  if (e0 != cudaSuccess) {
    if (prev != device_ordinal) cudaSetDevice(prev);
    return bcp::abi_fail("bcp_render_pixels: scratch allocation", e0);
  }
  float4* rays = reinterpret_cast<float4*>(scratch);
  float* depth = reinterpret_cast<float*>(scratch + nt * 16);
  __half* base = reinterpret_cast<__half*>(scratch + (nt * 20 + 15) / 16 * 16);  // [pixel][3][C][R], 16-byte aligned
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

// ---- bcp_render_pixels_static: the same two kernels with NO allocation anywhere on the call path ----
// (include/bcp_render.h; no counterpart in the reference).  The per-call scratch of bcp_render_pixels -
// rays | ground depth | background image - lives in a file-scope device array instead of the stream-ordered
// pool, for callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid it).
// Every GPU that loads libbcp.so holds BCP_RENDER_STATIC_SCRATCH_BYTES of it until the process exits.
namespace {

__device__ __align__(256) unsigned char g_render_scratch[BCP_RENDER_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
// In host code `g_render_scratch` is only the host-side stand-in of the symbol, not a device address:
// each device's copy is looked up once, with that device current, and the kernels get that pointer.
unsigned char* g_scratch_addr[kMaxDevices] = {nullptr};

int static_scratch(int device_ordinal, unsigned char** out) {  // the caller has made device_ordinal current
  if (device_ordinal < 0 || device_ordinal >= kMaxDevices) return bcp::abi_fail("bcp_render_pixels_static: device_ordinal out of range", cudaSuccess);
  if (!g_scratch_addr[device_ordinal]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_render_scratch);
    if (e != cudaSuccess || !p) return bcp::abi_fail("bcp_render_pixels_static: cudaGetSymbolAddress(g_render_scratch)", e);
    g_scratch_addr[device_ordinal] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_addr[device_ordinal];
  return 0;
}

}  // namespace

extern "C" {

size_t bcp_render_static_scratch_bytes(void) { return BCP_RENDER_STATIC_SCRATCH_BYTES; }

int bcp_render_pixels_static(const bcp_render_args* a, int device_ordinal, void* stream) {
  if (!a || !a->obs || !a->pixels) return bcp::abi_fail("bcp_render_pixels_static: null argument", cudaSuccess);
  if (a->num_envs <= 0 || a->width <= 0 || a->height <= 0) return bcp::abi_fail("bcp_render_pixels_static: empty batch or image", cudaSuccess);
  if (a->num_cameras != 1 && a->num_cameras != 2) return bcp::abi_fail("bcp_render_pixels_static: num_cameras must be 1 or 2", cudaSuccess);
  if (a->repeats < 1 || a->repeats > kMaxR) return bcp::abi_fail("bcp_render_pixels_static: repeats must be 1..8", cudaSuccess);
  if (!(a->near_val > 0.f) || !(a->far_val > a->near_val) || !(a->fov_deg > 0.f && a->fov_deg < 180.f))
    return bcp::abi_fail("bcp_render_pixels_static: bad near / far / fov", cudaSuccess);
  // the same scratch layout as bcp_render_pixels: rays f32x4 | ground depth f32 | (16-byte aligned) background f16
  const size_t nt = (size_t)a->num_cameras * (size_t)a->width * (size_t)a->height;
  const size_t img_at = (nt * 20 + 15) / 16 * 16;
  const size_t need = img_at + nt * 3 * (size_t)a->repeats * sizeof(__half);
  if (need > (size_t)BCP_RENDER_STATIC_SCRATCH_BYTES) {
    static thread_local char msg[200];
    snprintf(msg, sizeof msg,
             "bcp_render_pixels_static: num_cameras x width x height x repeats needs %zu bytes of scratch, the static array holds %zu "
             "(use bcp_render_pixels)", need, (size_t)BCP_RENDER_STATIC_SCRATCH_BYTES);
    return bcp::abi_fail(msg, cudaSuccess);
  }
  Scene sc;
  sc.B = a->num_envs; sc.R = a->repeats; sc.C = a->num_cameras; sc.W = a->width; sc.H = a->height;
  for (int b = 0; b < BCP_RENDER_BODIES; ++b)
    for (int i = 0; i < 3; ++i) { sc.h[b][i] = a->half_extents[b][i]; sc.rgb[b][i] = a->rgb[b][i]; }
  for (int i = 0; i < 3; ++i) { sc.ground_pos[i] = a->ground_pos[i]; sc.light[i] = a->light_dir[i]; sc.bg[i] = a->background[i]; }
  norm3(sc.light);
  sc.near_val = a->clip ? a->near_val : 1e-6f; sc.far_val = a->clip ? a->far_val : 1e30f; sc.ambient = a->ambient; sc.diffuse = a->diffuse;
  const float ty = tanf(0.5f * a->fov_deg * 3.14159265358979323846f / 180.f);
  const float aspect = (float)a->width / (float)a->height;
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
  if (prev != device_ordinal) {
    const cudaError_t es = cudaSetDevice(device_ordinal);
    if (es != cudaSuccess) {
      cudaGetLastError();  // not sticky: leave nothing for the next call's launch check
      return bcp::abi_fail("bcp_render_pixels_static: cudaSetDevice", es);
    }
  }
  unsigned char* scratch = nullptr;
  if (static_scratch(device_ordinal, &scratch)) {
    if (prev != device_ordinal) cudaSetDevice(prev);
    return -1;
  }
  cudaStream_t st = (cudaStream_t)stream;
  float4* rays = reinterpret_cast<float4*>(scratch);
  float* depth = reinterpret_cast<float*>(scratch + nt * 16);
  __half* base = reinterpret_cast<__half*>(scratch + img_at);
  // Nothing in the array has to start at zero: k_render_table writes every ray, depth and background word that
  // k_render reads (all `need` bytes in use) before k_render runs on the same stream; what the last call left
  // beyond them is never read.  Calls that overlap on the GPU (different streams / threads) race on the array.
  k_render_table<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(sc, rays, depth, base);
  const dim3 grid((unsigned)a->num_envs, 1);
  k_render<<<grid, kThreads, 0, st>>>(sc, a->obs, rays, depth, base, reinterpret_cast<__half*>(a->pixels));
  const cudaError_t e = cudaGetLastError();
  if (prev != device_ordinal) cudaSetDevice(prev);
  if (e != cudaSuccess) return bcp::abi_fail("bcp_render_pixels_static: launch", e);
  return 0;
}

}  // extern "C"
