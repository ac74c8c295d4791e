// R-MAC region pooling (regions + per-region L2 + sum + final L2) as a persistent, TMA-staged streaming kernel for sm_100a.
//
// Follows the same semantics as region_pool.cu (cvpce/models/classification.py:46-51 generalised to region lists; the
// north-star's "R-MAC kernel staging VGG16 conv5 feature tiles in shared memory, warp-shuffle max and norm reductions").
//
// Why a second kernel: the cluster kernel of region_pool.cu gives every CTA 1/8 of one image (128 KB at the VGG shapes)
// between barriers -- 0.48-0.51 of HBM channels-last, 0.17 NCHW (4-byte loads).  Here ONE persistent CTA per SM owns whole
// images, and the loads are taken off the threads entirely:
//  * a producer warp streams the image's maps, 16 KB at a time, into a shared-memory ring with cp.async.bulk (the TMA
//    engine, mbarrier complete_tx) and runs ahead across map and image boundaries -- up to 8 stages = 128 KB in flight per
//    SM, so the finishing phase of image n (norms, descriptor store) overlaps the loads of image n + 1;
//  * channels-last maps: a consumer thread owns one 16-byte channel group of every pixel (LDS.128 + 4 FMNMX per pixel) and
//    walks the pixels in raster order with CTA-uniform control flow: the running maximum of the current column segment is
//    folded into rs[column segment][channel] when the segment ends, and when a ROW segment ends every region that contains
//    it takes the maximum of its column segments.  Every shared-memory word is touched by one thread only: no atomics, no
//    barrier inside a map;
//  * NCHW maps: a consumer warp owns a plane (two or four planes for narrow maps, lane = (plane, column)): column maxima
//    per row segment -> cell maxima -> region maxima, all out of the staged plane in shared memory;
//  * all channels of an image are in one CTA, so the per-region L2 norms and the final L2 norm are block reductions (no
//    cluster, no workspace, no second launch).
// Every feature value is read from HBM exactly once.  Elementary segments = the intervals between the distinct region
// boundaries (plus 0 and the map size), as in region_pool.cu.
#include "common.cuh"
#include <math.h>
#include <string.h>
#include <algorithm>

namespace cvpce {

constexpr int kRmacConsWarps = 16;                        // consumer warps (channels-last maps use C*elemsize/16 threads of them)
constexpr int kRmacCons = kRmacConsWarps * 32;
constexpr int kRmacThreads = kRmacCons + 32;              // + the producer warp
constexpr int kRmacMaxSegs = 24;                          // elementary segments per axis
constexpr int kRmacMaxStages = 8;
constexpr int kRmacStageBytes = 16384;
constexpr int kRmacCmPitch = 33;
constexpr int kRmacMaxRegions = CVPCE_POOL_MAX_REGIONS;   // per map
constexpr int kRmacMaxRegionsTotal = 64;

struct RmacMap {
  const void* data;
  long long image_bytes;
  int C, H, W;
  int nhwc;          // 1: dense channels-last; 0: dense NCHW
  int n_regions, region_begin, out_offset;
  int nxs, nys;
  int n_units;       // pixels (channels-last) or planes (NCHW) per image
  int unit_bytes;    // bytes of one pixel (all channels) / of one plane
  int upc;           // units per chunk
  int n_chunks;      // per image
  unsigned char xb[kRmacMaxSegs + 2], yb[kRmacMaxSegs + 2];  // segment s = [b[s], b[s+1]), b[0] = 0, b[n] = W / H (one spare entry: the walkers read b[n+1])
};

struct RmacParams {
  RmacMap maps[CVPCE_POOL_MAX_MAPS];
  uchar4 segs[kRmacMaxRegionsTotal];  // (first row segment, end row segment, first column segment, end column segment)
  int n_maps, B;
  int per_region_l2, final_l2;
  float eps;
  float* out;
  long long out_stride;
  int nstages;
  int off_rs, off_vals, off_desc, off_wscr, off_bar;  // byte offsets into dynamic shared memory (the ring is at 0)
  int wscr_floats;                                    // per consumer warp (NCHW maps)
};

__device__ __forceinline__ uint32_t rm_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void rm_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(rm_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void rm_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rm_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void rm_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(rm_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool rm_mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(rm_smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void rm_mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!rm_mbar_try_wait(bar, parity)) {
  }
}
// global -> shared bulk copy by the TMA engine; completion is counted in bytes on `bar`.  The maps are read once: evict-first.
__device__ __forceinline__ void rm_bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(rm_smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(rm_smem_u32(bar)), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void rm_cons_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kRmacCons) : "memory"); }

template <int E>
__device__ __forceinline__ void rm_ld(float (&v)[E], const float* p) {
#pragma unroll
  for (int e = 0; e < E; e += 4) {
    const float4 t = *reinterpret_cast<const float4*>(p + e);
    v[e] = t.x; v[e + 1] = t.y; v[e + 2] = t.z; v[e + 3] = t.w;
  }
}
template <int E>
__device__ __forceinline__ void rm_st(float* p, const float (&v)[E]) {
#pragma unroll
  for (int e = 0; e < E; e += 4) *reinterpret_cast<float4*>(p + e) = make_float4(v[e], v[e + 1], v[e + 2], v[e + 3]);
}

// ---------------------------------------------------------------------------------------------------------------------
// channels-last map: region maxima -> vals[r][C].  `it` is the running chunk counter of the CTA (ring position).
// ---------------------------------------------------------------------------------------------------------------------
template <int DTYPE>
__device__ __forceinline__ void rmac_consume_nhwc(const RmacParams& p, const RmacMap& mp, const unsigned char* ring, uint64_t* full,
                                                  uint64_t* empty, float* rs, float* vals, uint32_t& it, int ctid, int lane) {
  constexpr int E = (DTYPE == CVPCE_F32) ? 4 : 8;  // channels per 16 bytes
  const int V = mp.C / E;
  const bool active = ctid < V;
  const int nst = p.nstages;
  const int W = mp.W, nxs = mp.nxs;
  float* rs_own = rs + ctid * E;
  float* vals_own = vals + ctid * E;
  float acc[E];
#pragma unroll
  for (int e = 0; e < E; ++e) acc[e] = -INFINITY;
  if (active)
    for (int s = 0; s < nxs; ++s) rm_st<E>(rs_own + s * mp.C, acc);
  int x = 0, y = 0, xs = 0, ys = 0;
  int xnext = mp.xb[1], ynext = mp.yb[1];
  for (int k = 0; k < mp.n_chunks; ++k, ++it) {
    const int st = it % nst;
    rm_mbar_wait(&full[st], (it / nst) & 1);
    if (active) {
      const int nu = min(mp.upc, mp.n_units - k * mp.upc);
      const unsigned char* buf = ring + (size_t)st * kRmacStageBytes + ctid * 16;
      for (int u0 = 0; u0 < nu; u0 += 4) {
        uint4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (u0 + j < nu) v[j] = *reinterpret_cast<const uint4*>(buf + (size_t)(u0 + j) * mp.unit_bytes);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (u0 + j < nu) {  // CTA-uniform
            if (DTYPE == CVPCE_F32) {
              acc[0] = fmaxf(acc[0], __uint_as_float(v[j].x));
              acc[1] = fmaxf(acc[1], __uint_as_float(v[j].y));
              acc[2] = fmaxf(acc[2], __uint_as_float(v[j].z));
              acc[3] = fmaxf(acc[3], __uint_as_float(v[j].w));
            } else {
              acc[0] = fmaxf(acc[0], __uint_as_float(v[j].x << 16)); acc[1] = fmaxf(acc[1], __uint_as_float(v[j].x & 0xffff0000u));
              acc[2] = fmaxf(acc[2], __uint_as_float(v[j].y << 16)); acc[3] = fmaxf(acc[3], __uint_as_float(v[j].y & 0xffff0000u));
              acc[4 % E] = fmaxf(acc[4 % E], __uint_as_float(v[j].z << 16)); acc[5 % E] = fmaxf(acc[5 % E], __uint_as_float(v[j].z & 0xffff0000u));
              acc[6 % E] = fmaxf(acc[6 % E], __uint_as_float(v[j].w << 16)); acc[7 % E] = fmaxf(acc[7 % E], __uint_as_float(v[j].w & 0xffff0000u));
            }
            ++x;
            if (x == xnext) {  // CTA-uniform: the column segment ends here (the last one ends with the row)
              float cur[E];
              rm_ld<E>(cur, rs_own + xs * mp.C);
#pragma unroll
              for (int e = 0; e < E; ++e) {
                cur[e] = fmaxf(cur[e], acc[e]);
                acc[e] = -INFINITY;
              }
              rm_st<E>(rs_own + xs * mp.C, cur);
              ++xs;
              if (x == W) {
                x = 0;
                xs = 0;
                ++y;
                if (y == ynext) {  // the row segment ends: every region that contains it takes its column segments
                  for (int r = 0; r < mp.n_regions; ++r) {
                    const uchar4 sg = p.segs[mp.region_begin + r];
                    if (ys >= sg.x && ys < sg.y) {
                      float m[E];
                      rm_ld<E>(m, rs_own + sg.z * mp.C);
                      for (int s = sg.z + 1; s < sg.w; ++s) {
                        float t[E];
                        rm_ld<E>(t, rs_own + s * mp.C);
#pragma unroll
                        for (int e = 0; e < E; ++e) m[e] = fmaxf(m[e], t[e]);
                      }
                      if (ys > sg.x) {
                        float t[E];
                        rm_ld<E>(t, vals_own + r * mp.C);
#pragma unroll
                        for (int e = 0; e < E; ++e) m[e] = fmaxf(m[e], t[e]);
                      }
                      rm_st<E>(vals_own + r * mp.C, m);
                    }
                  }
                  for (int s = 0; s < nxs; ++s) rm_st<E>(rs_own + s * mp.C, acc);  // acc is -inf here
                  ++ys;
                  ynext = mp.yb[ys + 1];
                }
              }
              xnext = mp.xb[xs + 1];
            }
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) rm_mbar_arrive(&empty[st]);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// NCHW map (W <= 32): a warp pass handles 32/Wp staged planes (Wp = W rounded up to 8, 16 or 32), lane = (plane, column).
// ---------------------------------------------------------------------------------------------------------------------
template <int DTYPE>
__device__ __forceinline__ float rm_ld_elem(const unsigned char* base, int idx) {
  if (DTYPE == CVPCE_F32) return reinterpret_cast<const float*>(base)[idx];
  return __uint_as_float((uint32_t)reinterpret_cast<const unsigned short*>(base)[idx] << 16);
}

template <int DTYPE>
__device__ __forceinline__ void rmac_consume_nchw(const RmacParams& p, const RmacMap& mp, const unsigned char* ring, uint64_t* full,
                                                  uint64_t* empty, float* wscr, float* vals, uint32_t& it, int warp, int lane) {
  const int nst = p.nstages;
  const int H = mp.H, W = mp.W, nxs = mp.nxs, nys = mp.nys;
  const int Wp = W > 16 ? 32 : (W > 8 ? 16 : 8);
  const int ppw = 32 / Wp;
  const int x = lane % Wp, psub = lane / Wp;
  const int ncell = nys * nxs;
  const int HW = H * W;
  float* cm = wscr;                          // [nys][kRmacCmPitch]
  float* cells = wscr + nys * kRmacCmPitch;  // [ppw][nys][nxs]
  int rot = 0;
  for (int k = 0; k < mp.n_chunks; ++k, ++it) {
    const int st = it % nst;
    rm_mbar_wait(&full[st], (it / nst) & 1);
    const int u0 = k * mp.upc;
    const int nu = min(mp.upc, mp.n_units - u0);
    const int ngroups = (nu + ppw - 1) / ppw;
    const unsigned char* buf = ring + (size_t)st * kRmacStageBytes;
    // plane group g of this chunk goes to warp (g + rot) % warps, so that chunks of few planes still spread over the warps
    for (int g = (warp - rot + kRmacConsWarps) % kRmacConsWarps; g < ngroups; g += kRmacConsWarps) {
      const int pl = g * ppw + psub;
      const bool live = x < W && pl < nu;
      const int base = pl * HW + x;
      {
        int ysr = 0, next = mp.yb[1];
        float v0 = -INFINITY;
#pragma unroll 4
        for (int y = 0; y < H; ++y) {
          const float e = live ? rm_ld_elem<DTYPE>(buf, base + y * W) : -INFINITY;
          v0 = fmaxf(v0, e);
          if (y + 1 == next) {  // warp-uniform: row-segment boundary
            cm[ysr * kRmacCmPitch + lane] = v0;
            v0 = -INFINITY;
            ++ysr;
            next = mp.yb[ysr + 1];
          }
        }
      }
      __syncwarp();
      for (int c = lane; c < ncell * ppw; c += 32) {
        const int ps = c / ncell, cc = c - ps * ncell;
        const int ysr = cc / nxs, xsr = cc - ysr * nxs;
        const float* row = cm + ysr * kRmacCmPitch + ps * Wp;
        float v0 = -INFINITY, v1 = -INFINITY;
        int xx = mp.xb[xsr];
        const int xe = mp.xb[xsr + 1];
        for (; xx + 1 < xe; xx += 2) {
          v0 = fmaxf(v0, row[xx]);
          v1 = fmaxf(v1, row[xx + 1]);
        }
        if (xx < xe) v0 = fmaxf(v0, row[xx]);
        cells[c] = fmaxf(v0, v1);
      }
      __syncwarp();
      for (int q = lane; q < mp.n_regions * ppw; q += 32) {
        const int ps = q / mp.n_regions, r = q - ps * mp.n_regions;
        const int plr = g * ppw + ps;
        if (plr >= nu) continue;
        const uchar4 sg = p.segs[mp.region_begin + r];
        float v0 = -INFINITY, v1 = -INFINITY;
        for (int ysr = sg.x; ysr < sg.y; ++ysr) {
          const float* crow = cells + ps * ncell + ysr * nxs;
          int xsr = sg.z;
          for (; xsr + 1 < sg.w; xsr += 2) {
            v0 = fmaxf(v0, crow[xsr]);
            v1 = fmaxf(v1, crow[xsr + 1]);
          }
          if (xsr < sg.w) v0 = fmaxf(v0, crow[xsr]);
        }
        vals[r * mp.C + u0 + plr] = fmaxf(v0, v1);
      }
      __syncwarp();  // cm / cells are reused by this warp's next pass
    }
    rot = (rot + ngroups) % kRmacConsWarps;
    __syncwarp();
    if (lane == 0) rm_mbar_arrive(&empty[st]);
  }
}

template <int DTYPE>
__global__ void __launch_bounds__(kRmacThreads, 1) rmac_staged_kernel(const __grid_constant__ RmacParams p) {
  extern __shared__ __align__(128) unsigned char rm_smem[];
  unsigned char* ring = rm_smem;
  float* rs = reinterpret_cast<float*>(rm_smem + p.off_rs);
  float* vals = reinterpret_cast<float*>(rm_smem + p.off_vals);
  float* desc = reinterpret_cast<float*>(rm_smem + p.off_desc);
  float* wscr_all = reinterpret_cast<float*>(rm_smem + p.off_wscr);
  uint64_t* full = reinterpret_cast<uint64_t*>(rm_smem + p.off_bar);
  uint64_t* empty = full + kRmacMaxStages;
  __shared__ float s_part[kRmacMaxRegions][kRmacConsWarps];
  __shared__ float s_denom[kRmacMaxRegions];
  __shared__ float s_wsum[kRmacConsWarps];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nst = p.nstages;
  if (tid == 0) {
    for (int s = 0; s < nst; ++s) {
      rm_mbar_init(&full[s], 1);
      rm_mbar_init(&empty[s], kRmacConsWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kRmacConsWarps) {
    // ---- producer: one thread keeps the ring full, across maps and images -----------------------------------------------
    if (lane == 0) {
      uint64_t policy;
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
      uint32_t it = 0;
      for (int n = blockIdx.x; n < p.B; n += gridDim.x) {
        for (int m = 0; m < p.n_maps; ++m) {
          const RmacMap& mp = p.maps[m];
          const unsigned char* src = reinterpret_cast<const unsigned char*>(mp.data) + (long long)n * mp.image_bytes;
          for (int k = 0; k < mp.n_chunks; ++k, ++it) {
            const int st = it % nst;
            rm_mbar_wait(&empty[st], ((it / nst) & 1) ^ 1);
            const int u0 = k * mp.upc;
            const uint32_t bytes = (uint32_t)min(mp.upc, mp.n_units - u0) * (uint32_t)mp.unit_bytes;
            rm_mbar_expect_tx(&full[st], bytes);
            rm_bulk_load(ring + (size_t)st * kRmacStageBytes, src + (long long)u0 * mp.unit_bytes, bytes, &full[st], policy);
          }
        }
      }
    }
    return;  // the consumers wait for every copy that was issued, so nothing is in flight when the CTA ends
  }

  // ---- consumers -----------------------------------------------------------------------------------------------------------
  const int ctid = tid;
  uint32_t it = 0;
  for (int n = blockIdx.x; n < p.B; n += gridDim.x) {
    float sq = 0.f;
    for (int m = 0; m < p.n_maps; ++m) {
      const RmacMap& mp = p.maps[m];
      if (mp.nhwc) rmac_consume_nhwc<DTYPE>(p, mp, ring, full, empty, rs, vals, it, ctid, lane);
      else rmac_consume_nchw<DTYPE>(p, mp, ring, full, empty, wscr_all + warp * p.wscr_floats, vals, it, warp, lane);
      rm_cons_sync();  // vals[r][C] of this map is complete
      if (p.per_region_l2) {
        for (int r = 0; r < mp.n_regions; ++r) {
          float s = 0.f;
          for (int c = ctid; c < mp.C; c += kRmacCons) {
            const float v = vals[r * mp.C + c];
            s = fmaf(v, v, s);
          }
          s = warp_sum(s);
          if (lane == 0) s_part[r][warp] = s;
        }
        rm_cons_sync();
        if (ctid < mp.n_regions) {
          float t = 0.f;
#pragma unroll
          for (int w = 0; w < kRmacConsWarps; ++w) t += s_part[ctid][w];
          s_denom[ctid] = fmaxf(sqrtf(t), p.eps);
        }
        rm_cons_sync();
      }
      for (int c = ctid; c < mp.C; c += kRmacCons) {
        float d = 0.f;
        for (int r = 0; r < mp.n_regions; ++r) {
          float v = vals[r * mp.C + c];
          if (p.per_region_l2) v = v / s_denom[r];
          d = (r == 0) ? v : d + v;
        }
        desc[mp.out_offset + c] = d;
        sq = fmaf(d, d, sq);
      }
      rm_cons_sync();  // vals, rs and s_denom are reused by the next map
    }
    float denom = 1.f;
    if (p.final_l2) {
      sq = warp_sum(sq);
      if (lane == 0) s_wsum[warp] = sq;
      rm_cons_sync();
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < kRmacConsWarps; ++w) t += s_wsum[w];
      denom = fmaxf(sqrtf(t), p.eps);
    }
    const RmacMap& last = p.maps[p.n_maps - 1];
    const int D = last.out_offset + last.C;
    float* orow = p.out + (long long)n * p.out_stride;
    for (int c = ctid; c < D; c += kRmacCons) orow[c] = p.final_l2 ? desc[c] / denom : desc[c];
    rm_cons_sync();  // desc / s_wsum are reused by the next image
  }
}

// Elementary segments of [0, dim) between the distinct boundaries of the intervals; returns the count (0: too many).
static int rmac_segments(const int* lo, const int* hi, int n, int dim, unsigned char* b) {
  int pts[2 * CVPCE_POOL_MAX_REGIONS + 2];
  int np = 0;
  pts[np++] = 0;
  pts[np++] = dim;
  for (int i = 0; i < n; ++i) {
    pts[np++] = lo[i];
    pts[np++] = hi[i];
  }
  std::sort(pts, pts + np);
  np = (int)(std::unique(pts, pts + np) - pts);
  if (np - 1 > kRmacMaxSegs || dim > 255) return 0;
  for (int i = 0; i < np; ++i) b[i] = (unsigned char)pts[i];
  return np - 1;
}
static int rmac_seg_index(const unsigned char* b, int ns, int v) {
  for (int i = 0; i <= ns; ++i)
    if (b[i] == v) return i;
  return 0;
}

// Launches the staged kernel when it covers the call.  Returns CVPCE_OK (launched), kRmacNotApplicable (the caller falls
// back to the cluster kernel) or an error code.  Arguments are already validated by region_pool_impl.
int rmac_staged_launch(const cvpce_feature_map* maps, int n_maps, int dtype, int B, const int32_t* regions, const int32_t* n_regions,
                       int per_region_l2, int final_l2, float eps, float* out, int64_t out_stride, void* stream) {
  const int mode = tuning("rmac_kernel");  // 0: automatic; 1: never; 2: whenever applicable
  if (mode == 1 || !(dtype == CVPCE_F32 || dtype == CVPCE_BF16) || B <= 0) return kRmacNotApplicable;
  // One CTA per image: below about half the SMs the cluster kernel's 8 CTAs per image fill the GPU better.
  if (mode == 0 && B < sm_count() / 2) return kRmacNotApplicable;
  const int esz = dtype == CVPCE_F32 ? 4 : 2;
  RmacParams p;
  memset(&p, 0, sizeof(p));
  p.n_maps = n_maps;
  p.B = B;
  p.per_region_l2 = per_region_l2 ? 1 : 0;
  p.final_l2 = final_l2 ? 1 : 0;
  p.eps = eps;
  p.out = out;
  p.out_stride = out_stride;
  int rbeg = 0, col = 0;
  size_t rs_floats = 0, vals_floats = 0, wscr_floats = 0;
  const int32_t* rg = regions;
  for (int m = 0; m < n_maps; ++m) {
    const cvpce_feature_map& fm = maps[m];
    RmacMap& pm = p.maps[m];
    pm.data = fm.data;
    pm.C = fm.C; pm.H = fm.H; pm.W = fm.W;
    pm.image_bytes = (long long)fm.stride_n * esz;
    if (!aligned16(fm.data) || pm.image_bytes % 16 != 0 || fm.H > 255 || fm.W > 255) return kRmacNotApplicable;
    const long long pix_bytes = (long long)fm.C * esz, plane_bytes = (long long)fm.H * fm.W * esz;
    const bool dense_nhwc = fm.stride_c == 1 && fm.stride_w == fm.C && fm.stride_h == (int64_t)fm.W * fm.C && pix_bytes % 16 == 0 &&
                            (pix_bytes / 16) % 32 == 0 && pix_bytes / 16 <= kRmacCons && pix_bytes <= kRmacStageBytes;
    const bool dense_nchw = fm.stride_w == 1 && fm.stride_h == fm.W && fm.stride_c == (int64_t)fm.H * fm.W && fm.W <= 32 &&
                            plane_bytes % 16 == 0 && plane_bytes <= kRmacStageBytes;
    if (!dense_nhwc && !dense_nchw) return kRmacNotApplicable;
    pm.nhwc = dense_nhwc ? 1 : 0;
    pm.n_regions = n_regions ? n_regions[m] : 1;
    if (pm.n_regions < 1 || pm.n_regions > kRmacMaxRegions || rbeg + pm.n_regions > kRmacMaxRegionsTotal) return kRmacNotApplicable;
    pm.region_begin = rbeg;
    int x0[CVPCE_POOL_MAX_REGIONS], x1[CVPCE_POOL_MAX_REGIONS], y0[CVPCE_POOL_MAX_REGIONS], y1[CVPCE_POOL_MAX_REGIONS];
    for (int r = 0; r < pm.n_regions; ++r) {
      if (rg) {
        y0[r] = rg[0]; x0[r] = rg[1]; y1[r] = rg[2]; x1[r] = rg[3];
        rg += 4;
      } else {
        y0[r] = 0; x0[r] = 0; y1[r] = fm.H; x1[r] = fm.W;
      }
    }
    pm.nxs = rmac_segments(x0, x1, pm.n_regions, fm.W, pm.xb);
    pm.nys = rmac_segments(y0, y1, pm.n_regions, fm.H, pm.yb);
    if (pm.nxs == 0 || pm.nys == 0) return kRmacNotApplicable;
    for (int r = 0; r < pm.n_regions; ++r)
      p.segs[rbeg + r] = make_uchar4((unsigned char)rmac_seg_index(pm.yb, pm.nys, y0[r]), (unsigned char)rmac_seg_index(pm.yb, pm.nys, y1[r]),
                                     (unsigned char)rmac_seg_index(pm.xb, pm.nxs, x0[r]), (unsigned char)rmac_seg_index(pm.xb, pm.nxs, x1[r]));
    rbeg += pm.n_regions;
    pm.out_offset = col;
    col += fm.C;
    if (pm.nhwc) {
      pm.n_units = fm.H * fm.W;
      pm.unit_bytes = (int)pix_bytes;
      pm.upc = std::max(1, kRmacStageBytes / pm.unit_bytes);
      rs_floats = std::max(rs_floats, (size_t)pm.nxs * fm.C);
    } else {
      const int Wp = fm.W > 16 ? 32 : (fm.W > 8 ? 16 : 8);
      const int ppw = 32 / Wp;
      pm.n_units = fm.C;
      pm.unit_bytes = (int)plane_bytes;
      pm.upc = std::max(1, kRmacStageBytes / pm.unit_bytes);
      if (pm.upc > ppw) pm.upc -= pm.upc % ppw;
      wscr_floats = std::max(wscr_floats, (size_t)pm.nys * kRmacCmPitch + (size_t)ppw * pm.nys * pm.nxs);
    }
    pm.n_chunks = (pm.n_units + pm.upc - 1) / pm.upc;
    vals_floats = std::max(vals_floats, (size_t)pm.n_regions * fm.C);
  }
  auto up128 = [](size_t v) { return (v + 127) & ~(size_t)127; };
  wscr_floats = (wscr_floats + 3) & ~(size_t)3;
  const size_t rest = up128(rs_floats * 4) + up128(vals_floats * 4) + up128((size_t)col * 4) + up128(wscr_floats * 4 * kRmacConsWarps) +
                      up128(2 * kRmacMaxStages * sizeof(uint64_t));
  const size_t budget = 220 * 1024;  // of the 227 KB a CTA may use; the kernel's static shared memory is ~3 KB
  if (rest + 2 * (size_t)kRmacStageBytes > budget) return kRmacNotApplicable;
  int nst = (int)std::min<size_t>(kRmacMaxStages, (budget - rest) / kRmacStageBytes);
  const int forced = tuning("rmac_stages");
  if (forced >= 2 && forced < nst) nst = forced;
  p.nstages = nst;
  size_t off = (size_t)nst * kRmacStageBytes;
  p.off_rs = (int)off;    off += up128(rs_floats * 4);
  p.off_vals = (int)off;  off += up128(vals_floats * 4);
  p.off_desc = (int)off;  off += up128((size_t)col * 4);
  p.off_wscr = (int)off;  off += up128(wscr_floats * 4 * kRmacConsWarps);
  p.off_bar = (int)off;   off += up128(2 * kRmacMaxStages * sizeof(uint64_t));
  p.wscr_floats = (int)wscr_floats;
  auto kern = dtype == CVPCE_F32 ? rmac_staged_kernel<CVPCE_F32> : rmac_staged_kernel<CVPCE_BF16>;
  if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), off)) return rc;
  // equal image counts per CTA: a CTA that alone carries one image more than the others would double the run time
  const int per_cta = (B + sm_count() - 1) / sm_count();
  const int grid = (B + per_cta - 1) / per_cta;
  kern<<<grid, kRmacThreads, off, (cudaStream_t)stream>>>(p);
  CVPCE_LAUNCH_OK("rmac_staged_kernel");
  return CVPCE_OK;
}

}  // namespace cvpce
