// Headless rasteriser for `render('rgb_array')`, sm_100a.
//
// The reference draws every env with pyglet / fixed-function OpenGL (environment.py:263-412, rendering.py:44-112,
// 223-250): a 700 x 700 window cleared to white, camera bounds +-cam_range around the origin (shared viewer), and per
// entity (agents first, then landmarks: world.entities, core.py:142-144) a FilledPolygon circle of `res` = 30 vertices
// and radius entity.size, blended SRC_ALPHA / ONE_MINUS_SRC_ALPHA (agents alpha 0.5), followed by its outline
// (GL_LINE_LOOP, colour and alpha halved, width 2); after all entities come the communication circles of the
// non-silent agents (dim_c circles of radius size / dim_c along the agent's diameter, colour 1 - state.c[ci]).
// No graphics stack exists on this box, so this is a restatement of that fixed-function raster as a CUDA kernel that
// writes the RGB image straight into a device buffer:
//   * pixel-centre sampling: a pixel belongs to a polygon when its centre lies inside (all edge functions >= 0);
//   * the outline covers a pixel when its centre is within line_width / 2 of an edge segment (hard coverage: GL's
//     smooth-line antialiasing is implementation defined), blended once per primitive;
//   * the colour buffer is 8 bits per channel: every blend result is rounded to 1/255 like an RGBA8 framebuffer.
// All arithmetic is float64 without FMA contraction (this file is compiled with -fmad=false) in the operation order of
// oracle/render.py, so images match the oracle bit for bit.
//
// HBM-bound: 3 bytes written per pixel, 48 + 16 bytes of state read per primitive and image.  One thread shades four
// horizontally adjacent pixels and stores them as three 32-bit words; a CTA owns a band of 16 image rows, loads the
// draw list once, and each of its warps walks 32 x 4 pixel patches, culling the list against its patch with a ballot,
// so background pixels cost one shared-memory test.
#include "common.cuh"

namespace mappo {
namespace {

constexpr int kMaxPrim = 96;       // entities + communication circles per image
constexpr int kMaxRes = 64;        // vertices per circle
constexpr int kBandH = 16, kWarps = 8, kPix = 4;

struct RenderArgs {
  mappo_env_state st;
  int N, M, L, W, H, E, res;
  double cam, half_width;
  const int32_t* env_ids;
  const float* size;               // [E]
  const float* rgba;               // [images][E][4]
  const int32_t* comm_dim;         // [E] or null
  const double* unit;              // [res][2]
  uint8_t* out;
};

struct Prim {
  double cx, cy, r;                // world coordinates
  double c[4];                     // r g b alpha
};

__device__ __forceinline__ double quant(double v) { return floor(v * 255.0 + 0.5) / 255.0; }

__global__ void __launch_bounds__(256) render_kernel(const RenderArgs a) {
  __shared__ Prim prims[kMaxPrim];
  __shared__ double ux[kMaxRes], uy[kMaxRes];
  __shared__ int active_s[kWarps][kMaxPrim];
  __shared__ int cdim_s[kMaxPrim], said_s[kMaxPrim];
  __shared__ int n_prim_s;
  const int img = blockIdx.y;
  const int n = a.env_ids ? a.env_ids[img] : img;
  const int tid = threadIdx.y * 32 + threadIdx.x;
  const double sx = (double)a.W / (a.cam + a.cam), sy = (double)a.H / (a.cam + a.cam);
  if (tid < a.res) { ux[tid] = a.unit[2 * tid]; uy[tid] = a.unit[2 * tid + 1]; }
  // primitive list in draw order: entity circles, then the communication circles (viewer.geoms, environment.py:367-374).
  // One thread per entity fetches its state (independent loads in flight together); thread 0 then appends the
  // communication circles from shared memory.
  if (tid < a.E) {
    const int e = tid;
    double wx, wy;
    if (e < a.M) { wx = a.st.p_pos[(size_t)(0 * a.M + e) * a.N + n]; wy = a.st.p_pos[(size_t)(1 * a.M + e) * a.N + n]; }
    else { const int l = e - a.M; wx = a.st.lm_pos[(size_t)(0 * a.L + l) * a.N + n]; wy = a.st.lm_pos[(size_t)(1 * a.L + l) * a.N + n]; }
    Prim& p = prims[e];
    p.cx = wx; p.cy = wy; p.r = (double)a.size[e];
    const float4 c = *reinterpret_cast<const float4*>(a.rgba + ((size_t)img * a.E + e) * 4);
    p.c[0] = (double)c.x; p.c[1] = (double)c.y; p.c[2] = (double)c.z; p.c[3] = (double)c.w;
    cdim_s[e] = a.comm_dim ? a.comm_dim[e] : 0;
    said_s[e] = (e < a.M) ? a.st.comm[(size_t)e * a.N + n] : -1;
  }
  __syncthreads();
  if (tid == 0) {
    int np_ = a.E;
    for (int e = 0; e < a.E; ++e) {
      const int dc = cdim_s[e];
      if (dc <= 0 || np_ + dc > kMaxPrim) continue;
      const double es = prims[e].r;
      const double cs = es / (double)dc;                                   // comm_size = entity.size / dim_c
      const int said = said_s[e];
      for (int ci = 0; ci < dc; ++ci) {
        Prim& p = prims[np_++];
        p.cx = prims[e].cx + (((double)ci * cs) * 2.0 - es + cs);          // offset ci * comm_size * 2 - size + comm_size
        p.cy = prims[e].cy;
        p.r = cs;
        const double col = (said == ci) ? 0.0 : 1.0;                       // 1 - state.c[ci]
        p.c[0] = col; p.c[1] = col; p.c[2] = col; p.c[3] = 1.0;
      }
    }
    n_prim_s = np_;
  }
  __syncthreads();
  const int np_ = n_prim_s;
  // A CTA owns a band of kBandH image rows; each warp walks 32 x 4 pixel patches of the band (8 threads x 4 pixels
  // wide, 4 rows: compact, so few warps straddle a circle) and culls the draw list against its own patch with one
  // ballot per 32 primitives -- no block-wide synchronisation after the prologue.
  const int wrp = threadIdx.y, lane = threadIdx.x;
  int* active = active_s[wrp];
  const int patches_x = (a.W + 31) / 32;
  const int n_patches = patches_x * (kBandH / 4);
  const double hw2 = a.half_width * a.half_width;
  const double smax = sx > sy ? sx : sy;
  for (int pt = wrp; pt < n_patches; pt += kWarps) {
    const int x0 = (pt % patches_x) * 32, y0 = blockIdx.x * kBandH + (pt / patches_x) * 4;   // patch origin (row 0 = top)
    if (y0 >= a.H) break;
    // cull (bounds in GL window coordinates, one pixel of slack), compacted in draw order
    const double tx0 = (double)x0 - 1.0, tx1 = (double)(x0 + 32) + 1.0;
    const double ty1 = (double)(a.H - y0) + 1.0, ty0 = (double)(a.H - y0 - 4) - 1.0;
    int na = 0;
    __syncwarp();
    for (int base = 0; base < np_; base += 32) {
      const int p = base + lane;
      bool hit = false;
      if (p < np_) {
        const double pcx = (prims[p].cx + a.cam) * sx, pcy = (prims[p].cy + a.cam) * sy;
        const double rr = prims[p].r * smax + a.half_width;
        hit = pcx + rr >= tx0 && pcx - rr <= tx1 && pcy + rr >= ty0 && pcy - rr <= ty1;
      }
      const unsigned mask = __ballot_sync(0xffffffffu, hit);
      if (hit) active[na + __popc(mask & ((1u << lane) - 1u))] = p;
      na += __popc(mask);
    }
    __syncwarp();
    const int px0 = x0 + (lane & 7) * kPix, row = y0 + (lane >> 3);
    if (px0 >= a.W || row >= a.H) continue;
    double col[kPix][3];
#pragma unroll
    for (int i = 0; i < kPix; ++i) { col[i][0] = 1.0; col[i][1] = 1.0; col[i][2] = 1.0; }   // glClearColor(1,1,1,1)
    const double py = (double)(a.H - 1 - row) + 0.5;
    for (int ai = 0; ai < na; ++ai) {
      const Prim& p = prims[active[ai]];
      const double pcx = (p.cx + a.cam) * sx, pcy = (p.cy + a.cam) * sy;
      const double rmax = p.r * smax + a.half_width;
      const double rmax2 = rmax * rmax;
      bool any = false, near[kPix];
#pragma unroll
      for (int i = 0; i < kPix; ++i) {
        const double dx = ((double)(px0 + i) + 0.5) - pcx, dy = py - pcy;
        near[i] = (dx * dx + dy * dy <= rmax2);                            // every covered pixel lies in this disc
        any = any || near[i];
      }
      if (!any) continue;
      bool inside[kPix], edge[kPix];
#pragma unroll
      for (int i = 0; i < kPix; ++i) { inside[i] = true; edge[i] = false; }
      double vx0 = ((p.cx + ux[0] * p.r) + a.cam) * sx, vy0 = ((p.cy + uy[0] * p.r) + a.cam) * sy;
      for (int k = 0; k < a.res; ++k) {
        const int k1 = (k + 1 == a.res) ? 0 : k + 1;
        const double vx1 = ((p.cx + ux[k1] * p.r) + a.cam) * sx, vy1 = ((p.cy + uy[k1] * p.r) + a.cam) * sy;
        const double ex = vx1 - vx0, ey = vy1 - vy0;
        const double inv_len2 = 1.0 / (ex * ex + ey * ey);                  // one division per edge, none per pixel
#pragma unroll
        for (int i = 0; i < kPix; ++i) {
          const double qx = ((double)(px0 + i) + 0.5) - vx0, qy = py - vy0;
          const double cross = ex * qy - ey * qx;
          inside[i] = inside[i] && (cross >= 0.0);
          double t = (qx * ex + qy * ey) * inv_len2;
          t = fmin(fmax(t, 0.0), 1.0);
          const double rx = qx - t * ex, ry = qy - t * ey;
          edge[i] = edge[i] || (rx * rx + ry * ry <= hw2);
        }
        vx0 = vx1; vy0 = vy1;
      }
      const double al = p.c[3], al2 = p.c[3] * 0.5;
#pragma unroll
      for (int i = 0; i < kPix; ++i) {
        if (inside[i] && near[i]) {
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) col[i][ch] = quant(p.c[ch] * al + col[i][ch] * (1.0 - al));
        }
        if (edge[i] && near[i]) {
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) col[i][ch] = quant((p.c[ch] * 0.5) * al2 + col[i][ch] * (1.0 - al2));
        }
      }
    }
    // 4 pixels = 12 bytes = three aligned 32-bit words (W is a multiple of 4)
    uint8_t b[12];
  #pragma unroll
    for (int i = 0; i < kPix; ++i) {
  #pragma unroll
      for (int ch = 0; ch < 3; ++ch) b[3 * i + ch] = (uint8_t)(int)floor(col[i][ch] * 255.0 + 0.5);
    }
    uint32_t* dst = reinterpret_cast<uint32_t*>(a.out + ((size_t)img * a.H + row) * (size_t)a.W * 3 + (size_t)px0 * 3);
  #pragma unroll
    for (int w = 0; w < 3; ++w)
      dst[w] = (uint32_t)b[4 * w] | ((uint32_t)b[4 * w + 1] << 8) | ((uint32_t)b[4 * w + 2] << 16) | ((uint32_t)b[4 * w + 3] << 24);
  }
}

}  // namespace
}  // namespace mappo

extern "C" int mappo_env_render(const mappo_env_config* cfg, const mappo_env_state* st, const mappo_render_config* rc,
                                const int32_t* env_ids, const float* entity_size, const float* entity_rgba,
                                const int32_t* comm_dim, const double* unit_circle, uint8_t* rgb_out, void* stream) {
  using namespace mappo;
  MAPPO_REQUIRE(cfg && st && rc && entity_size && entity_rgba && unit_circle && rgb_out, "env_render: null argument");
  MAPPO_REQUIRE(rc->width > 0 && rc->height > 0 && rc->width % 4 == 0, "env_render: width must be a positive multiple of 4");
  MAPPO_REQUIRE(rc->num_images >= 0 && rc->num_images <= 65535, "env_render: num_images %d outside [0, 65535]", rc->num_images);
  MAPPO_REQUIRE(rc->res >= 3 && rc->res <= kMaxRes, "env_render: res %d outside [3, %d]", rc->res, kMaxRes);
  const int L_all = mappo_env_num_landmarks(cfg);
  MAPPO_REQUIRE(rc->num_entities == cfg->num_agents + L_all, "env_render: num_entities must be M + L");
  MAPPO_REQUIRE(rc->num_entities <= kMaxPrim, "env_render: too many entities");
  MAPPO_REQUIRE(rc->cam_range > 0.0 && rc->line_width >= 0.0, "env_render: bad camera range / line width");
  MAPPO_REQUIRE((reinterpret_cast<uintptr_t>(rgb_out) & 3u) == 0, "env_render: rgb_out must be 4-byte aligned");
  MAPPO_REQUIRE((reinterpret_cast<uintptr_t>(entity_rgba) & 15u) == 0, "env_render: entity_rgba must be 16-byte aligned");
  if (rc->num_images == 0) return MAPPO_OK;
  RenderArgs a;
  a.st = *st;
  a.N = cfg->num_envs; a.M = cfg->num_agents; a.L = L_all;
  a.W = rc->width; a.H = rc->height; a.E = rc->num_entities; a.res = rc->res;
  a.cam = rc->cam_range; a.half_width = rc->line_width * 0.5;
  a.env_ids = env_ids; a.size = entity_size; a.rgba = entity_rgba; a.comm_dim = comm_dim; a.unit = unit_circle;
  a.out = rgb_out;
  dim3 grid((unsigned)ceil_div(a.H, kBandH), (unsigned)rc->num_images), block(32, kWarps);
  render_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(a);
  MAPPO_CHECK_LAUNCH("env_render");
  return MAPPO_OK;
}
