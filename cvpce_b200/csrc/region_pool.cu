// Region max-pool embedding (MAC / R-MAC) + L2 normalisation for sm_100a.
//
// Follows cvpce/models/classification.py:46-51 (MACVGG: amax over H,W of relu4_3 and relu5_3, cat, L2 with
// clamp(min=eps)) and :79-85 (MACResNet).  R-MAC (regions + per-region L2 + sum) is the north-star extension.
//
// One thread-block CLUSTER (8 CTAs) per crop: CTA r of the cluster owns the r-th channel slab of every feature map,
// streams it once from HBM, keeps the region maxima in shared memory, and the norms that span all channels (per-region
// L2 and the final L2) are all-reduced across the cluster through distributed shared memory -- no workspace, no second
// launch, the descriptor is written once.
//
// Every feature value is read from HBM exactly once, whatever the number of (overlapping) regions: the host splits the
// rows and the columns of a map into the elementary SEGMENTS between region boundaries, the streaming pass reduces each
// segment once, and a region is then the maximum over the few segments it spans.
//  * NCHW maps: one warp per plane.  MAC: 128-bit streaming loads + warp-shuffle max.  R-MAC: lane = column; the warp
//    walks the rows once (coalesced 128-byte rows) keeping the column maxima of the current ROW segment, parks them in
//    shared memory, then lane r reduces the row segments x column range of region r.
//  * channels-last (NHWC) maps: 4 channels per thread (one 128-bit load per pixel), one row per thread group; the
//    maxima of every COLUMN segment of the row go to shared memory, then one thread per (region, channel) reduces the
//    region's rows x column segments.
#include "common.cuh"
#include <cuda_fp16.h>
#include <cooperative_groups.h>
#include <math.h>
#include <algorithm>
#include <mutex>

namespace cg = cooperative_groups;

namespace cvpce {

constexpr int kPoolThreads = 256;
constexpr int kPoolWarps = kPoolThreads / 32;
constexpr int kPoolCluster = 8;
constexpr int kMaxRegionsTotal = 64;
constexpr int kMaxSegs = 16;   // elementary row / column segments per map handled by the separable paths
constexpr int kCmPitch = 33;   // per-warp column-maxima rows, padded against bank conflicts
constexpr int kWarpScratch = kMaxSegs * kCmPitch + 4 * kMaxSegs * kMaxSegs;  // NCHW R-MAC: column maxima + cell maxima (<= 4 planes) per warp

struct PoolMap {
  const void* data;
  int C, H, W;
  long long sn, sc, sh, sw;
  int region_begin, n_regions;  // into PoolParams::regions
  int out_offset;               // column offset of this map in the descriptor
  int slab;                     // channels per CTA of the cluster
  int vals_offset;              // float offset of this map's [n_regions][slab] block in smem
  int desc_offset;              // float offset of this map's [slab] block in the descriptor smem
  int whole;                    // 1: single region covering the whole map
  int nxs, nys;                 // elementary column / row segments (0: too many -> generic path)
  unsigned char xb[kMaxSegs + 1], yb[kMaxSegs + 1];  // segment s = [b[s], b[s+1])  (maps up to 255 cells per side)
  int nhwc;                     // 1: channels-last fast path applies (unit channel stride, aligned channel quads)
};

struct PoolParams {
  PoolMap maps[CVPCE_POOL_MAX_MAPS];
  int4 regions[kMaxRegionsTotal];  // (y0,x0,y1,x1)
  uchar4 segs[kMaxRegionsTotal];   // (first row segment, end row segment, first column segment, end column segment)
  int n_maps, B;
  int per_region_l2, final_l2;
  float eps;
  float* out;
  long long out_stride;
  int vals_total, desc_total, pairs_total;
  int scratch_floats;  // shared scratch: NHWC cell maxima [nys][nxs][slab] | NCHW [warps][kWarpScratch]
};

template <int DTYPE>
__device__ __forceinline__ float load_elem(const void* base, long long off) {
  if (DTYPE == CVPCE_F32) return __ldg(reinterpret_cast<const float*>(base) + off);
  return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[off]);
}

__device__ __forceinline__ float4 ldg_stream_f4(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ldg_stream_u4(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float ldg_stream_f1(const float* p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ float bf16lo(unsigned u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(unsigned u) { return __uint_as_float(u & 0xffff0000u); }

// 4 consecutive channels of one pixel
template <int DTYPE>
__device__ __forceinline__ float4 load_quad(const char* p) {
  if (DTYPE == CVPCE_F32) return ldg_stream_f4(reinterpret_cast<const float4*>(p));
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return make_float4(bf16lo(r.x), bf16hi(r.x), bf16lo(r.y), bf16hi(r.y));
}
__device__ __forceinline__ void max4(float4& a, const float4& b) {
  a.x = fmaxf(a.x, b.x);
  a.y = fmaxf(a.y, b.y);
  a.z = fmaxf(a.z, b.z);
  a.w = fmaxf(a.w, b.w);
}

// monotone float <-> int mapping so that shared-memory atomicMax (s32) orders floats of either sign
__device__ __forceinline__ int f2ord(float f) {
  const int i = __float_as_int(f);
  return i ^ ((i >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i ^ ((i >> 31) & 0x7fffffff)); }

// ------------------------------------------------------------------------------------------------------------------------
// channels-last phase 1 for one map: region maxima of channels [c_lo, c_lo + nplanes) of crop n -> vals.
// Stage A: thread group g (nplanes/4 threads, 4 channels each) walks rows g, g+G, ... with NB independent 128-bit loads
// in flight; the maximum of every column segment of the row is folded into cells[row segment][column segment][channel]
// with shared-memory atomicMax.  Stage B: one thread per (region, channel) reduces the region's cells.
// MAC is the one-cell, one-region case of the same code.
// ------------------------------------------------------------------------------------------------------------------------
template <int DTYPE, int NB>
__device__ __forceinline__ void pool_map_nhwc(const PoolParams& p, const PoolMap& mp, int n, int c_lo, int nplanes, float* scratch,
                                              float* vals) {
  constexpr int ESZ = (DTYPE == CVPCE_F32) ? 4 : 2;
  const int tid = threadIdx.x;
  const int tpp = nplanes >> 2;  // threads per pixel
  const int groups = kPoolThreads / tpp;
  const int g = tid / tpp, cq = tid - g * tpp;
  const int ns = mp.nxs;
  int* cells = reinterpret_cast<int*>(scratch);  // [nys][ns][nplanes], order-mapped floats
  const int ncell = mp.nys * ns * nplanes;
  const int neg_inf = f2ord(-INFINITY);
  for (int i = tid; i < ncell; i += kPoolThreads) cells[i] = neg_inf;
  __syncthreads();
  if (g < groups) {
    const char* base = reinterpret_cast<const char*>(mp.data) + ((long long)n * mp.sn + c_lo + 4 * cq) * ESZ;
    const long long xstep = mp.sw * ESZ;
    const int x_first = mp.xb[0], x_last = mp.xb[ns];
    for (int y = g; y < mp.H; y += groups) {
      if (y < mp.yb[0] || y >= mp.yb[mp.nys]) continue;  // row outside every region
      int ys = 0;
      while (y >= mp.yb[ys + 1]) ++ys;
      int* dst = cells + (ys * ns) * nplanes + 4 * cq;
      const char* rowp = base + (long long)y * mp.sh * ESZ;
      int sx = 0, next = mp.xb[1];
      float4 acc = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
      for (int xb0 = x_first; xb0 < x_last; xb0 += NB) {
        float4 v[NB];
#pragma unroll
        for (int u = 0; u < NB; ++u)
          if (xb0 + u < x_last) v[u] = load_quad<DTYPE>(rowp + (xb0 + u) * xstep);
#pragma unroll
        for (int u = 0; u < NB; ++u) {
          const int x = xb0 + u;
          if (x < x_last) {
            if (x == next) {  // CTA-uniform: column-segment boundary
              atomicMax(dst, f2ord(acc.x));
              atomicMax(dst + 1, f2ord(acc.y));
              atomicMax(dst + 2, f2ord(acc.z));
              atomicMax(dst + 3, f2ord(acc.w));
              acc = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
              dst += nplanes;
              ++sx;
              next = mp.xb[sx + 1];
            }
            max4(acc, v[u]);
          }
        }
      }
      atomicMax(dst, f2ord(acc.x));
      atomicMax(dst + 1, f2ord(acc.y));
      atomicMax(dst + 2, f2ord(acc.z));
      atomicMax(dst + 3, f2ord(acc.w));
    }
  }
  __syncthreads();
  for (int it = tid; it < mp.n_regions * nplanes; it += kPoolThreads) {
    const int r = it / nplanes, pc = it - r * nplanes;
    const uchar4 sg = p.segs[mp.region_begin + r];
    int v0 = neg_inf, v1 = neg_inf;
    for (int ys = sg.x; ys < sg.y; ++ys) {
      const int* crow = cells + (ys * ns) * nplanes + pc;
      int xs = sg.z;
      for (; xs + 1 < sg.w; xs += 2) {
        v0 = max(v0, crow[xs * nplanes]);
        v1 = max(v1, crow[(xs + 1) * nplanes]);
      }
      if (xs < sg.w) v0 = max(v0, crow[xs * nplanes]);
    }
    vals[mp.vals_offset + r * mp.slab + pc] = ord2f(max(v0, v1));
  }
  __syncthreads();  // the scratch is reused by the next map
}

// ------------------------------------------------------------------------------------------------------------------------
// NCHW, whole-map maximum of one plane by one warp (MAC)
// ------------------------------------------------------------------------------------------------------------------------
template <int DTYPE>
__device__ __forceinline__ float plane_max_whole(const PoolMap& mp, long long poff, int lane) {
  constexpr int VEC = (DTYPE == CVPCE_F32) ? 4 : 8;
  constexpr int ESZ = (DTYPE == CVPCE_F32) ? 4 : 2;
  const int HW = mp.H * mp.W;
  const char* pbase = reinterpret_cast<const char*>(mp.data) + poff * ESZ;
  const bool vec_ok = (mp.sw == 1 && mp.sh == mp.W) && (HW % VEC == 0) && ((reinterpret_cast<uintptr_t>(pbase) & 15u) == 0);
  float v = -INFINITY;
  if (vec_ok) {
    const int nv = HW / VEC;
    if (DTYPE == CVPCE_F32) {
      const float4* src = reinterpret_cast<const float4*>(pbase);
      int i = lane;
      for (; i + 96 < nv; i += 128) {  // 4 independent 128-bit loads in flight per lane
        const float4 a = ldg_stream_f4(src + i), b = ldg_stream_f4(src + i + 32);
        const float4 c = ldg_stream_f4(src + i + 64), d = ldg_stream_f4(src + i + 96);
        v = fmaxf(v, fmaxf(fmaxf(fmaxf(a.x, a.y), fmaxf(a.z, a.w)), fmaxf(fmaxf(b.x, b.y), fmaxf(b.z, b.w))));
        v = fmaxf(v, fmaxf(fmaxf(fmaxf(c.x, c.y), fmaxf(c.z, c.w)), fmaxf(fmaxf(d.x, d.y), fmaxf(d.z, d.w))));
      }
      for (; i < nv; i += 32) {
        const float4 a = ldg_stream_f4(src + i);
        v = fmaxf(v, fmaxf(fmaxf(a.x, a.y), fmaxf(a.z, a.w)));
      }
    } else {
      const uint4* src = reinterpret_cast<const uint4*>(pbase);
      for (int i = lane; i < nv; i += 32) {
        const uint4 a = ldg_stream_u4(src + i);
        v = fmaxf(v, fmaxf(fmaxf(fmaxf(bf16lo(a.x), bf16hi(a.x)), fmaxf(bf16lo(a.y), bf16hi(a.y))),
                           fmaxf(fmaxf(bf16lo(a.z), bf16hi(a.z)), fmaxf(bf16lo(a.w), bf16hi(a.w)))));
      }
    }
  } else {
    for (int i = lane; i < HW; i += 32) {
      const int y = i / mp.W, x = i - y * mp.W;
      v = fmaxf(v, load_elem<DTYPE>(mp.data, poff + y * mp.sh + x * mp.sw));
    }
  }
  return warp_max(v);
}

// ------------------------------------------------------------------------------------------------------------------------
// NCHW, several regions (R-MAC), W <= 32 and H <= 32: a warp handles 32/Wp planes at a time (Wp = W rounded up to 8, 16
// or 32), lane = (plane, column).  Every lane first pulls its whole column into registers (up to 32 independent loads
// in flight), folds it into the maxima of the row segments (boundaries are warp-uniform), parks them in cm[s][lane];
// then one lane per (plane, cell) reduces the cell's columns and one lane per (plane, region) the region's cells.
// ------------------------------------------------------------------------------------------------------------------------
template <int DTYPE>
__device__ __forceinline__ void planes_regions_nchw(const PoolParams& p, const PoolMap& mp, int n, int c_lo, int nplanes, int warp,
                                                    int lane, float* wscratch, float* vals) {
  const int Wp = mp.W > 16 ? 32 : (mp.W > 8 ? 16 : 8);
  const int ppw = 32 / Wp;  // planes per warp pass
  const int x = lane % Wp, psub = lane / Wp;
  const int y_first = mp.yb[0], y_last = mp.yb[mp.nys];
  float* cm = wscratch;                           // [nys][kCmPitch]
  float* cells = wscratch + kMaxSegs * kCmPitch;  // [ppw][nys][nxs]
  const int ncell = mp.nys * mp.nxs;
  for (int pg = warp; pg * ppw < nplanes; pg += kPoolWarps) {
    const int pc = pg * ppw + psub;
    const bool live = x < mp.W && pc < nplanes;
    const long long poff = (long long)n * mp.sn + (long long)(c_lo + pc) * mp.sc + (long long)x * mp.sw;
    float colv[32];
#pragma unroll
    for (int y = 0; y < 32; ++y) {
      colv[y] = -INFINITY;
      if (live && y >= y_first && y < y_last) colv[y] = load_elem<DTYPE>(mp.data, poff + (long long)y * mp.sh);
    }
    int sy = 0, next = mp.yb[1];
    float v = -INFINITY;
#pragma unroll
    for (int y = 0; y < 32; ++y) {
      if (y >= y_first && y < y_last) {  // warp-uniform
        if (y == next) {                 // warp-uniform: row-segment boundary
          cm[sy * kCmPitch + lane] = v;
          v = -INFINITY;
          ++sy;
          next = mp.yb[sy + 1];
        }
        v = fmaxf(v, colv[y]);
      }
    }
    cm[sy * kCmPitch + lane] = v;
    __syncwarp();
    for (int c = lane; c < ncell * ppw; c += 32) {
      const int ps = c / ncell, cc = c - ps * ncell;
      const int ys = cc / mp.nxs, xs = cc - ys * mp.nxs;
      const float* row = cm + ys * kCmPitch + ps * Wp;
      float v0 = -INFINITY, v1 = -INFINITY;
      int xx = mp.xb[xs];
      const int xe = mp.xb[xs + 1];
      for (; xx + 1 < xe; xx += 2) {
        v0 = fmaxf(v0, row[xx]);
        v1 = fmaxf(v1, row[xx + 1]);
      }
      if (xx < xe) v0 = fmaxf(v0, row[xx]);
      cells[c] = fmaxf(v0, v1);
    }
    __syncwarp();
    for (int it = lane; it < mp.n_regions * ppw; it += 32) {
      const int ps = it / mp.n_regions, r = it - ps * mp.n_regions;
      const int pcr = pg * ppw + ps;
      if (pcr >= nplanes) continue;
      const uchar4 sg = p.segs[mp.region_begin + r];
      float v0 = -INFINITY, v1 = -INFINITY;
      for (int ys = sg.x; ys < sg.y; ++ys) {
        const float* crow = cells + ps * ncell + ys * mp.nxs;
        int xs = sg.z;
        for (; xs + 1 < sg.w; xs += 2) {
          v0 = fmaxf(v0, crow[xs]);
          v1 = fmaxf(v1, crow[xs + 1]);
        }
        if (xs < sg.w) v0 = fmaxf(v0, crow[xs]);
      }
      vals[mp.vals_offset + r * mp.slab + pcr] = fmaxf(v0, v1);
    }
    __syncwarp();  // cm / cells are reused by this warp's next pass
  }
}

// GENERAL = false: every map is pooled over its whole extent (MAC, the reference's mode) -- the region code is not
// compiled in, which keeps the streaming kernel at 4 resident CTAs per SM.
// NHWC_ONLY = true: every map takes the channels-last path -- the per-plane NCHW code (and its 32-register column
// buffer) is not compiled in, so the region kernel also fits 4 resident CTAs per SM.
template <int DTYPE, bool GENERAL, bool NHWC_ONLY>
__global__ void __launch_bounds__(kPoolThreads, GENERAL ? (NHWC_ONLY ? 3 : 2) : 4) region_pool_kernel(const __grid_constant__ PoolParams p) {
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned crank = cluster.block_rank();
  const unsigned csize = cluster.num_blocks();
  const int n = blockIdx.x / csize;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  extern __shared__ __align__(16) float smem[];
  float* vals = smem;                         // [vals_total]
  float* desc = vals + p.vals_total;          // [desc_total]
  float* scratch = desc + p.desc_total;       // [scratch_floats]
  __shared__ float part[kMaxRegionsTotal];    // partial sum of squares per (map,region), read by the cluster
  __shared__ float denom[kMaxRegionsTotal];
  __shared__ float fpart;                     // partial sum of squares of the descriptor slab
  __shared__ float wsum[kPoolWarps];

  // ---- phase 1: region maxima of this CTA's channel slabs ----------------------------------------------------------
  for (int m = 0; m < p.n_maps; ++m) {
    const PoolMap& mp = p.maps[m];
    const int c_lo = (int)crank * mp.slab;
    const int nplanes = max(0, min(mp.slab, mp.C - c_lo));
    if (NHWC_ONLY || mp.nhwc) {  // CTA-uniform
      if (m > 0) __syncthreads();  // warps of the previous (NCHW) map may still be reading their part of the scratch
      if (nplanes > 0) pool_map_nhwc<DTYPE, GENERAL ? 8 : 4>(p, mp, n, c_lo, nplanes, scratch, vals);
      continue;
    }
    if (NHWC_ONLY) continue;
    if (GENERAL && !mp.whole && mp.nys > 0 && mp.W <= 32 && mp.H <= 32) {  // CTA-uniform
      planes_regions_nchw<DTYPE>(p, mp, n, c_lo, nplanes, warp, lane, scratch + warp * kWarpScratch, vals);
      continue;
    }
    for (int pc = warp; pc < nplanes; pc += kPoolWarps) {
      const long long poff = (long long)n * mp.sn + (long long)(c_lo + pc) * mp.sc;
      float* vout = vals + mp.vals_offset + pc;
      if (!GENERAL || mp.whole) {
        const float v = plane_max_whole<DTYPE>(mp, poff, lane);
        if (lane == 0) vout[0] = v;
      } else {
        for (int r = 0; r < mp.n_regions; ++r) {  // generic: any strides, any size, any region list
          const int4 rg = p.regions[mp.region_begin + r];
          const int rw = rg.w - rg.y, rh = rg.z - rg.x;
          float v = -INFINITY;
          for (int i = lane; i < rw * rh; i += 32) {
            const int y = rg.x + i / rw, x = rg.y + i % rw;
            v = fmaxf(v, load_elem<DTYPE>(mp.data, poff + y * mp.sh + x * mp.sw));
          }
          v = warp_max(v);
          if (lane == 0) vout[r * mp.slab] = v;
        }
      }
    }
  }
  __syncthreads();

  // ---- phase 2: per-region L2 norms over ALL channels (cluster all-reduce through DSMEM) -----------------------------
  if (GENERAL && p.per_region_l2) {
    for (int pair = warp; pair < p.pairs_total; pair += kPoolWarps) {
      // locate the map of this (map,region) pair
      int m = 0;
      while (m + 1 < p.n_maps && pair >= p.maps[m + 1].region_begin) ++m;
      const PoolMap& mp = p.maps[m];
      const int r = pair - mp.region_begin;
      const int nplanes = max(0, min(mp.slab, mp.C - (int)crank * mp.slab));
      float s = 0.f;
      for (int pc = lane; pc < nplanes; pc += 32) {
        const float v = vals[mp.vals_offset + r * mp.slab + pc];
        s = fmaf(v, v, s);
      }
      s = warp_sum(s);
      if (lane == 0) part[pair] = s;
    }
    cluster.sync();
    for (int pair = tid; pair < p.pairs_total; pair += kPoolThreads) {
      float tot = 0.f;
      for (unsigned rk = 0; rk < csize; ++rk) tot += cluster.map_shared_rank(part, rk)[pair];
      denom[pair] = fmaxf(sqrtf(tot), p.eps);
    }
    __syncthreads();
  }

  // ---- phase 3: descriptor slab = sum over regions, partial norm ------------------------------------------------------
  float sq = 0.f;
  for (int m = 0; m < p.n_maps; ++m) {
    const PoolMap& mp = p.maps[m];
    const int nplanes = max(0, min(mp.slab, mp.C - (int)crank * mp.slab));
    for (int pc = tid; pc < nplanes; pc += kPoolThreads) {
      float d = 0.f;
      for (int r = 0; r < mp.n_regions; ++r) {
        float v = vals[mp.vals_offset + r * mp.slab + pc];
        if (GENERAL && p.per_region_l2) v = v / denom[mp.region_begin + r];
        d = (r == 0) ? v : d + v;
      }
      desc[mp.desc_offset + pc] = d;
      sq = fmaf(d, d, sq);
    }
  }
  float total_denom = 1.f;
  if (p.final_l2) {
    sq = warp_sum(sq);
    if (lane == 0) wsum[warp] = sq;
    __syncthreads();
    if (tid == 0) {
      float s = 0.f;
      for (int w = 0; w < kPoolWarps; ++w) s += wsum[w];
      fpart = s;
    }
    cluster.sync();
    float tot = 0.f;
    for (unsigned rk = 0; rk < csize; ++rk) tot += *cluster.map_shared_rank(&fpart, rk);
    total_denom = fmaxf(sqrtf(tot), p.eps);
  } else {
    __syncthreads();
  }
  for (int m = 0; m < p.n_maps; ++m) {
    const PoolMap& mp = p.maps[m];
    const int c_lo = (int)crank * mp.slab;
    const int nplanes = max(0, min(mp.slab, mp.C - c_lo));
    float* orow = p.out + (long long)n * p.out_stride + mp.out_offset + c_lo;
    for (int pc = tid; pc < nplanes; pc += kPoolThreads) {
      const float d = desc[mp.desc_offset + pc];
      orow[pc] = p.final_l2 ? d / total_denom : d;
    }
  }
  // a CTA's shared memory must stay alive until every peer has read its partials
  cluster.sync();
}

// ---------------------------------------------------------------------------------------------------------------------
// MAC over dense channels-last maps (what the fast trunk produces; the reference's mode): PIXEL-split across the cluster.
// The channel-split kernel above reads, per CTA, a 256-byte channel slab of every pixel (2 KB apart): 65-70 % of HBM.
// Here CTA r of the 8-CTA cluster streams the r-th eighth of each map's pixels -- one contiguous block of memory, 128-bit
// loads, 8 in flight per thread, a thread always sees the same 16-byte channel group so its running maxima stay in
// registers (no shared-memory atomics) -- and the per-CTA maxima are max-reduced across the cluster through distributed
// shared memory; the channel slab r of the descriptor is then finished (sum of squares all-reduce, scale, store) by CTA r.
// Requirements (host-checked): whole-map regions, no per-region L2, sc == 1, sw == C, sh == W*C, 16-byte aligned images,
// C*elemsize/16 divides 256.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kMacMaxCols = 4096;  // sum of C over the maps

#ifndef CVPCE_MAC_MIN_CTAS
#define CVPCE_MAC_MIN_CTAS 1
#endif
#ifndef CVPCE_MAC_UNROLL
#define CVPCE_MAC_UNROLL 8
#endif
#define CVPCE_PRAGMA(x) _Pragma(#x)
#define CVPCE_UNROLL(n) CVPCE_PRAGMA(unroll n)

template <int DT>
__global__ void __launch_bounds__(kPoolThreads, CVPCE_MAC_MIN_CTAS) mac_nhwc_pixelsplit_kernel(const PoolParams p) {
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned crank = cluster.block_rank(), csize = cluster.num_blocks();
  const int n = blockIdx.x / csize;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr bool BF16 = DT != CVPCE_F32;  // any 16-bit element type
  constexpr int E = BF16 ? 8 : 4;  // elements per 16-byte vector
  __shared__ float pmax[kMacMaxCols];          // this CTA's maxima over its pixels, every channel of every map
  __shared__ float red[kPoolThreads * E];
  __shared__ float own[kMacMaxCols / 2];  // finished maxima of this CTA's channel slab (cluster size >= 2)
  __shared__ float wsum[kPoolWarps];
  __shared__ float fpart;

  int col = 0;
  for (int m = 0; m < p.n_maps; ++m) {
    const PoolMap& mp = p.maps[m];
    const int V = mp.C / E;
    const int npix = mp.H * mp.W;
    const int lo = (int)((long long)npix * crank / csize), hi = (int)((long long)npix * (crank + 1) / csize);
    const uint4* base = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(mp.data) + (size_t)n * mp.sn * (BF16 ? 2 : 4)) +
                        (size_t)lo * V;
    const int total = (hi - lo) * V;
    float acc[E];
#pragma unroll
    for (int e = 0; e < E; ++e) acc[e] = -INFINITY;
    CVPCE_UNROLL(CVPCE_MAC_UNROLL)
    for (int i = tid; i < total; i += kPoolThreads) {
      const uint4 x = ldg_stream_u4(base + i);
      if (DT == CVPCE_F16) {
        const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&x.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&x.y));
        const float2 c = __half22float2(*reinterpret_cast<const __half2*>(&x.z)), d = __half22float2(*reinterpret_cast<const __half2*>(&x.w));
        acc[0] = fmaxf(acc[0], a.x); acc[1] = fmaxf(acc[1], a.y);
        acc[2] = fmaxf(acc[2], b.x); acc[3] = fmaxf(acc[3], b.y);
        acc[4 % E] = fmaxf(acc[4 % E], c.x); acc[5 % E] = fmaxf(acc[5 % E], c.y);
        acc[6 % E] = fmaxf(acc[6 % E], d.x); acc[7 % E] = fmaxf(acc[7 % E], d.y);
      } else if (BF16) {
        acc[0] = fmaxf(acc[0], bf16lo(x.x)); acc[1] = fmaxf(acc[1], bf16hi(x.x));
        acc[2] = fmaxf(acc[2], bf16lo(x.y)); acc[3] = fmaxf(acc[3], bf16hi(x.y));
        acc[4 % E] = fmaxf(acc[4 % E], bf16lo(x.z)); acc[5 % E] = fmaxf(acc[5 % E], bf16hi(x.z));
        acc[6 % E] = fmaxf(acc[6 % E], bf16lo(x.w)); acc[7 % E] = fmaxf(acc[7 % E], bf16hi(x.w));
      } else {
        acc[0] = fmaxf(acc[0], __uint_as_float(x.x)); acc[1] = fmaxf(acc[1], __uint_as_float(x.y));
        acc[2] = fmaxf(acc[2], __uint_as_float(x.z)); acc[3] = fmaxf(acc[3], __uint_as_float(x.w));
      }
    }
#pragma unroll
    for (int e = 0; e < E; ++e) red[tid * E + e] = acc[e];
    __syncthreads();
    if (tid < V) {  // threads tid, tid+V, tid+2V, .. saw the same channel group (V divides the block size)
      for (int g = tid + V; g < kPoolThreads; g += V) {
#pragma unroll
        for (int e = 0; e < E; ++e) acc[e] = fmaxf(acc[e], red[g * E + e]);
      }
#pragma unroll
      for (int e = 0; e < E; ++e) pmax[col + tid * E + e] = acc[e];
    }
    __syncthreads();
    col += mp.C;
  }
  cluster.sync();
  // channel slab `crank` of every map: max over the cluster, then the descriptor's L2 norm over all channels
  float sq = 0.f;
  int ocol = 0, mcol = 0;
  for (int m = 0; m < p.n_maps; ++m) {
    const PoolMap& mp = p.maps[m];
    const int c_lo = (int)crank * mp.slab;
    const int nch = max(0, min(mp.slab, mp.C - c_lo));
    for (int c = tid; c < nch; c += kPoolThreads) {
      float v = -INFINITY;
      for (unsigned rk = 0; rk < csize; ++rk) v = fmaxf(v, cluster.map_shared_rank(pmax, rk)[mcol + c_lo + c]);
      own[ocol + c] = v;
      sq = fmaf(v, v, sq);
    }
    ocol += mp.slab;
    mcol += mp.C;
  }
  float denom = 1.f;
  if (p.final_l2) {
    sq = warp_sum(sq);
    if (lane == 0) wsum[warp] = sq;
    __syncthreads();
    if (tid == 0) {
      float t = 0.f;
      for (int w = 0; w < kPoolWarps; ++w) t += wsum[w];
      fpart = t;
    }
    cluster.sync();
    float tot = 0.f;
    for (unsigned rk = 0; rk < csize; ++rk) tot += *cluster.map_shared_rank(&fpart, rk);
    denom = fmaxf(sqrtf(tot), p.eps);
  } else {
    __syncthreads();
  }
  ocol = 0;
  for (int m = 0; m < p.n_maps; ++m) {
    const PoolMap& mp = p.maps[m];
    const int c_lo = (int)crank * mp.slab;
    const int nch = max(0, min(mp.slab, mp.C - c_lo));
    float* orow = p.out + (long long)n * p.out_stride + mp.out_offset + c_lo;
    for (int c = tid; c < nch; c += kPoolThreads) orow[c] = p.final_l2 ? own[ocol + c] / denom : own[ocol + c];
    ocol += mp.slab;
  }
  cluster.sync();  // peers may still be reading pmax / fpart
}

// ---------------------------------------------------------------------------------------------------------------------
// MAC over dense channels-last maps, balanced streaming form (used when the caller provides a workspace).
// The cluster kernel above gives every image its own cluster: B = 256 images are 512 CTAs on 148 SMs -- 3.46 per SM, the
// slowest SM carries 4 -- and every cluster ends in two cluster barriers.  Here the whole batch is one linear space of
// CHUNKS (16 KB runs of whole pixels of one map of one image, images outermost) cut into gridDim.x equal contiguous ranges,
// gridDim.x = a multiple of the SM count: every CTA streams the same number of bytes whatever B is.  A CTA's range crosses
// at most a few (image, map) boundaries; for each piece (a SEGMENT) it writes its per-channel maxima to a workspace row, and
// the CTA that completes an image's LAST segment (a counter per image, release / acquire through __threadfence) reduces
// that image's rows from L2, takes the L2 norm and writes the descriptor.  No CTA ever waits for another one.
// The counters are left at zero, so a workspace zeroed once can be reused by every later launch on the same stream.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kMacChunkVecs = 1024;  // 16-byte vectors per chunk (a multiple of the block size: a thread keeps its channel group)

struct MacStreamPlan {
  long long total_chunks;                  // B * chunks_per_image
  int chunks_per_image;
  int chunk_pix[CVPCE_POOL_MAX_MAPS];      // pixels per chunk
  int chunks[CVPCE_POOL_MAX_MAPS];         // chunks per (image, map)
  int slots[CVPCE_POOL_MAX_MAPS];          // workspace rows per (image, map): an upper bound of its segment count
  int slot_offset[CVPCE_POOL_MAX_MAPS];    // float offset of a map's rows inside an image's block
  long long image_stride;                  // floats per image in the workspace
  unsigned* counters;                      // [B], zero between launches
  float* partial;                          // [B][image_stride]
};

// the CTA whose range [c*T/G, (c+1)*T/G) holds chunk x
__device__ __forceinline__ long long mac_owner(long long x, long long G, long long T) { return ((x + 1) * G - 1) / T; }

#ifndef CVPCE_MAC_STREAM_MIN_CTAS
#define CVPCE_MAC_STREAM_MIN_CTAS 4
#endif
template <int DT>
__global__ void __launch_bounds__(kPoolThreads, CVPCE_MAC_STREAM_MIN_CTAS) mac_nhwc_stream_kernel(const __grid_constant__ PoolParams p, const __grid_constant__ MacStreamPlan w) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr bool BF16 = DT != CVPCE_F32;
  constexpr int E = BF16 ? 8 : 4;
  __shared__ float red[kPoolThreads * E];
  __shared__ float own[kMacMaxCols];
  __shared__ float wsum[kPoolWarps];
  __shared__ int s_last;
  const long long G = gridDim.x, T = w.total_chunks;
  long long x = (long long)blockIdx.x * T / G;
  const long long x_end = ((long long)blockIdx.x + 1) * T / G;
  while (x < x_end) {
    const int n = (int)(x / w.chunks_per_image);
    int j = (int)(x - (long long)n * w.chunks_per_image);
    int m = 0;
    while (j >= w.chunks[m]) j -= w.chunks[m++];
    const PoolMap& mp = p.maps[m];
    const int j_end = (int)min((long long)w.chunks[m], (long long)j + (x_end - x));
    const long long map_first = x - j;  // chunk index of this (image, map)'s first chunk
    const int V = mp.C / E;
    const int npix = mp.H * mp.W;
    const int lo = j * w.chunk_pix[m], hi = min(npix, j_end * w.chunk_pix[m]);
    const uint4* base = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(mp.data) + (size_t)n * mp.sn * (BF16 ? 2 : 4)) + (size_t)lo * V;
    const int total = (hi - lo) * V;
    float acc[E];
#pragma unroll
    for (int e = 0; e < E; ++e) acc[e] = -INFINITY;
    CVPCE_UNROLL(CVPCE_MAC_UNROLL)
    for (int i = tid; i < total; i += kPoolThreads) {
      const uint4 v = ldg_stream_u4(base + i);
      if (DT == CVPCE_F16) {
        const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&v.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
        const float2 c = __half22float2(*reinterpret_cast<const __half2*>(&v.z)), d = __half22float2(*reinterpret_cast<const __half2*>(&v.w));
        acc[0] = fmaxf(acc[0], a.x); acc[1] = fmaxf(acc[1], a.y);
        acc[2] = fmaxf(acc[2], b.x); acc[3] = fmaxf(acc[3], b.y);
        acc[4 % E] = fmaxf(acc[4 % E], c.x); acc[5 % E] = fmaxf(acc[5 % E], c.y);
        acc[6 % E] = fmaxf(acc[6 % E], d.x); acc[7 % E] = fmaxf(acc[7 % E], d.y);
      } else if (BF16) {
        acc[0] = fmaxf(acc[0], bf16lo(v.x)); acc[1] = fmaxf(acc[1], bf16hi(v.x));
        acc[2] = fmaxf(acc[2], bf16lo(v.y)); acc[3] = fmaxf(acc[3], bf16hi(v.y));
        acc[4 % E] = fmaxf(acc[4 % E], bf16lo(v.z)); acc[5 % E] = fmaxf(acc[5 % E], bf16hi(v.z));
        acc[6 % E] = fmaxf(acc[6 % E], bf16lo(v.w)); acc[7 % E] = fmaxf(acc[7 % E], bf16hi(v.w));
      } else {
        acc[0] = fmaxf(acc[0], __uint_as_float(v.x)); acc[1] = fmaxf(acc[1], __uint_as_float(v.y));
        acc[2] = fmaxf(acc[2], __uint_as_float(v.z)); acc[3] = fmaxf(acc[3], __uint_as_float(v.w));
      }
    }
#pragma unroll
    for (int e = 0; e < E; ++e) red[tid * E + e] = acc[e];
    __syncthreads();
    // segment row of this CTA = its ordinal among the CTAs that touch this (image, map)
    const int slot = (int)((long long)blockIdx.x - mac_owner(map_first, G, T));
    if (tid < V) {  // threads tid, tid+V, .. saw the same channel group (segments start at multiples of the block size)
      for (int g = tid + V; g < kPoolThreads; g += V) {
#pragma unroll
        for (int e = 0; e < E; ++e) acc[e] = fmaxf(acc[e], red[g * E + e]);
      }
      float* row = w.partial + (long long)n * w.image_stride + w.slot_offset[m] + (long long)slot * mp.C + tid * E;
#pragma unroll
      for (int e = 0; e < E; e += 4) __stcg(reinterpret_cast<float4*>(row + e), make_float4(acc[e], acc[e + 1], acc[e + 2], acc[e + 3]));
      __threadfence();  // release: the row is visible device-wide before the counter moves
    }
    __syncthreads();
    x += j_end - j;
    if (tid == 0) {
      // segments of image n, from the same arithmetic every CTA uses
      unsigned expect = 0;
      long long first = (long long)n * w.chunks_per_image;
      for (int q = 0; q < p.n_maps; ++q) {
        expect += (unsigned)(mac_owner(first + w.chunks[q] - 1, G, T) - mac_owner(first, G, T) + 1);
        first += w.chunks[q];
      }
      const unsigned prev = atomicAdd(&w.counters[n], 1u);
      s_last = prev + 1 == expect;
      if (s_last) {
        w.counters[n] = 0;  // nobody else touches this counter any more in this launch: leave it clean for the next one
        __threadfence();    // acquire: the other CTAs' rows, released before their increments, are visible below
      }
    }
    __syncthreads();
    if (s_last) {
      // ---- finish image n: maxima over its segments, L2 norm over all channels, store --------------------------------------
      float sq = 0.f;
      int col = 0;
      long long first = (long long)n * w.chunks_per_image;
      for (int q = 0; q < p.n_maps; ++q) {
        const PoolMap& mq = p.maps[q];
        const int nseg = (int)(mac_owner(first + w.chunks[q] - 1, G, T) - mac_owner(first, G, T) + 1);
        const float* rows = w.partial + (long long)n * w.image_stride + w.slot_offset[q];
        for (int c = tid; c < mq.C; c += kPoolThreads) {
          float v = -INFINITY;
          for (int sg = 0; sg < nseg; ++sg) v = fmaxf(v, __ldcg(rows + (long long)sg * mq.C + c));
          own[col + c] = v;
          sq = fmaf(v, v, sq);
        }
        col += mq.C;
        first += w.chunks[q];
      }
      float denom = 1.f;
      if (p.final_l2) {
        sq = warp_sum(sq);
        if (lane == 0) wsum[warp] = sq;
        __syncthreads();
        float t = 0.f;
        for (int k = 0; k < kPoolWarps; ++k) t += wsum[k];
        denom = fmaxf(sqrtf(t), p.eps);
      }
      col = 0;
      for (int q = 0; q < p.n_maps; ++q) {
        const PoolMap& mq = p.maps[q];
        float* orow = p.out + (long long)n * p.out_stride + mq.out_offset;
        for (int c = tid; c < mq.C; c += kPoolThreads) orow[c] = p.final_l2 ? own[col + c] / denom : own[col + c];
        col += mq.C;
      }
      __syncthreads();  // own / wsum / s_last are reused by the next segment
    }
  }
}

__global__ void __launch_bounds__(256) l2norm_rows_kernel(float* x, int B, int D, long long stride, float eps) {
  __shared__ float wsum[8];
  __shared__ float s_denom;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int row = blockIdx.x; row < B; row += gridDim.x) {
    float* r = x + (long long)row * stride;
    float s = 0.f;
    for (int i = tid; i < D; i += 256) {
      const float v = r[i];
      s = fmaf(v, v, s);
    }
    s = warp_sum(s);
    if (lane == 0) wsum[warp] = s;
    __syncthreads();
    if (tid == 0) {
      float t = 0.f;
      for (int w = 0; w < 8; ++w) t += wsum[w];
      s_denom = fmaxf(sqrtf(t), eps);
    }
    __syncthreads();
    const float dn = s_denom;
    for (int i = tid; i < D; i += 256) r[i] = r[i] / dn;
    __syncthreads();
  }
}

}  // namespace cvpce

extern "C" int cvpce_rmac_regions(int H, int W, int levels, int32_t* regions, int max_regions) {
  // Tolias, Sicre, Jegou 2016 (R-MAC), square-ish grid per scale l = 1..levels:
  // region side = floor(2*min(H,W)/(l+1)); along each axis the regions are spread uniformly over [0, dim - side].
  int count = 0;
  const int w = H < W ? H : W;
  for (int l = 1; l <= levels; ++l) {
    int side = (2 * w) / (l + 1);
    if (side < 1) side = 1;
    // number of regions along each axis: l on the short side, enough to keep ~40% overlap on the long side
    const int ny = (H == w) ? l : l + (int)((double)(H - w) / (double)w * (l + 1) / 2.0 + 0.5);
    const int nx = (W == w) ? l : l + (int)((double)(W - w) / (double)w * (l + 1) / 2.0 + 0.5);
    for (int iy = 0; iy < ny; ++iy) {
      const int y0 = ny > 1 ? (int)((double)(H - side) * iy / (ny - 1)) : (H - side) / 2;
      for (int ix = 0; ix < nx; ++ix) {
        const int x0 = nx > 1 ? (int)((double)(W - side) * ix / (nx - 1)) : (W - side) / 2;
        if (count < max_regions && regions) {
          regions[4 * count + 0] = y0;
          regions[4 * count + 1] = x0;
          regions[4 * count + 2] = y0 + side;
          regions[4 * count + 3] = x0 + side;
        }
        ++count;
      }
    }
  }
  return count;
}

// Elementary segments between the distinct boundaries of a list of [lo,hi) intervals: b[0..ns]; returns ns (0 if more
// than kMaxSegs or a boundary does not fit a byte).
namespace cvpce {
static int make_segments(const int* lo, const int* hi, int n, unsigned char* b) {
  int pts[2 * CVPCE_POOL_MAX_REGIONS];
  int np = 0;
  for (int i = 0; i < n; ++i) {
    pts[np++] = lo[i];
    pts[np++] = hi[i];
  }
  std::sort(pts, pts + np);
  np = (int)(std::unique(pts, pts + np) - pts);
  if (np - 1 > kMaxSegs || pts[np - 1] > 255) return 0;
  for (int i = 0; i < np; ++i) b[i] = (unsigned char)pts[i];
  return np - 1;
}
static int seg_index(const unsigned char* b, int ns, int v) {
  for (int i = 0; i <= ns; ++i)
    if (b[i] == v) return i;
  return 0;
}
}  // namespace cvpce

namespace cvpce {

// Whether the dense channels-last MAC kernels apply to these maps (whole-map regions and no per-region L2 are checked by the caller).
static bool mac_dense_nhwc(const cvpce_feature_map* maps, int n_maps, int dtype) {
  const int E = dtype == CVPCE_F32 ? 4 : 8;
  int col = 0;
  for (int m = 0; m < n_maps; ++m) {
    const cvpce_feature_map& fm = maps[m];
    if (!fm.data || fm.C <= 0 || fm.H <= 0 || fm.W <= 0) return false;
    const int V = fm.C / E;
    if (!(fm.C % E == 0 && V >= 1 && V <= kPoolThreads && kPoolThreads % V == 0 && fm.stride_c == 1 && fm.stride_w == fm.C &&
          fm.stride_h == (int64_t)fm.W * fm.C && fm.stride_n % E == 0 && (reinterpret_cast<uintptr_t>(fm.data) & 15u) == 0))
      return false;
    col += fm.C;
  }
  return col <= kMacMaxCols;
}

// Chunking of the batch for mac_nhwc_stream_kernel and the workspace it needs; returns the grid size.
static int mac_stream_plan(const cvpce_feature_map* maps, int n_maps, int dtype, int B, MacStreamPlan* w, size_t* counter_bytes, size_t* total_bytes) {
  const int E = dtype == CVPCE_F32 ? 4 : 8;
  int cpi = 0;
  for (int m = 0; m < n_maps; ++m) {
    const int V = maps[m].C / E;
    w->chunk_pix[m] = kMacChunkVecs / V;
    w->chunks[m] = (maps[m].H * maps[m].W + w->chunk_pix[m] - 1) / w->chunk_pix[m];
    cpi += w->chunks[m];
  }
  w->chunks_per_image = cpi;
  w->total_chunks = (long long)B * cpi;
  const int per_sm = tuning("mac_ctas_per_sm") > 0 ? tuning("mac_ctas_per_sm") : 2;  // measured (tools/mac_ab.py, same box): 2 beats 1, 3, 4, 6
  const long long G = std::max(1ll, std::min<long long>((long long)sm_count() * per_sm, w->total_chunks));
  long long off = 0;
  for (int m = 0; m < n_maps; ++m) {
    const long long bound = ((long long)w->chunks[m] * G + w->total_chunks - 1) / w->total_chunks + 1;
    w->slots[m] = (int)std::min<long long>(w->chunks[m], bound);
    w->slot_offset[m] = (int)off;
    off += (long long)w->slots[m] * maps[m].C;
  }
  w->image_stride = off;
  // a fixed-size counter block for every realistic batch, so that one zeroed workspace serves launches of different B
  *counter_bytes = std::max<size_t>(65536, (((size_t)B * sizeof(unsigned)) + 255) & ~(size_t)255);
  *total_bytes = *counter_bytes + (size_t)B * (size_t)off * sizeof(float);
  return (int)G;
}

}  // namespace cvpce

extern "C" int64_t cvpce_region_pool_workspace_bytes(const cvpce_feature_map* maps, int n_maps, int dtype, int B) {
  using namespace cvpce;
  if (!maps || n_maps < 1 || n_maps > CVPCE_POOL_MAX_MAPS || B <= 0) return 0;
  if (!(dtype == CVPCE_F32 || dtype == CVPCE_BF16 || dtype == CVPCE_F16) || !mac_dense_nhwc(maps, n_maps, dtype)) return 0;
  MacStreamPlan w;
  size_t cb = 0, tb = 0;
  mac_stream_plan(maps, n_maps, dtype, B, &w, &cb, &tb);
  return (int64_t)tb;
}

static int region_pool_impl(const cvpce_feature_map* maps, int n_maps, int dtype, int B,
                            const int32_t* regions, const int32_t* n_regions, int per_region_l2,
                            int final_l2, float eps, float* out, int64_t out_stride, void* workspace, size_t workspace_bytes, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(maps && n_maps >= 1 && n_maps <= CVPCE_POOL_MAX_MAPS, CVPCE_ERR_INVALID_ARG, "region_pool: n_maps=%d out of [1,%d]", n_maps, CVPCE_POOL_MAX_MAPS);
  CVPCE_REQUIRE(dtype == CVPCE_F32 || dtype == CVPCE_BF16 || dtype == CVPCE_F16, CVPCE_ERR_INVALID_ARG, "region_pool: bad dtype %d", dtype);
  CVPCE_REQUIRE(B >= 0, CVPCE_ERR_INVALID_ARG, "region_pool: B=%d", B);
  CVPCE_REQUIRE((regions == nullptr) == (n_regions == nullptr), CVPCE_ERR_INVALID_ARG, "region_pool: regions and n_regions must both be given or both null");
  if (B == 0) return CVPCE_OK;
  CVPCE_REQUIRE(out, CVPCE_ERR_INVALID_ARG, "region_pool: out is null");
  if (int rc = require_sm100()) return rc;

  PoolParams p;
  p.n_maps = n_maps;
  p.B = B;
  p.per_region_l2 = per_region_l2 ? 1 : 0;
  p.final_l2 = final_l2 ? 1 : 0;
  p.eps = eps;
  p.out = out;
  p.out_stride = out_stride;
  int rbeg = 0, col = 0, vals = 0, desc = 0, scratch = kPoolWarps * kWarpScratch;
  bool all_whole = true;
  const int32_t* rg = regions;
  for (int m = 0; m < n_maps; ++m) {
    const cvpce_feature_map& fm = maps[m];
    CVPCE_REQUIRE(fm.data && fm.C > 0 && fm.H > 0 && fm.W > 0, CVPCE_ERR_INVALID_ARG, "region_pool: map %d is null or has an empty dimension", m);
    PoolMap& pm = p.maps[m];
    pm.data = fm.data;
    pm.C = fm.C; pm.H = fm.H; pm.W = fm.W;
    pm.sn = fm.stride_n; pm.sc = fm.stride_c; pm.sh = fm.stride_h; pm.sw = fm.stride_w;
    pm.region_begin = rbeg;
    pm.n_regions = n_regions ? n_regions[m] : 1;
    CVPCE_REQUIRE(pm.n_regions >= 1 && pm.n_regions <= CVPCE_POOL_MAX_REGIONS && rbeg + pm.n_regions <= kMaxRegionsTotal,
                  CVPCE_ERR_UNSUPPORTED, "region_pool: map %d has %d regions (per map <= %d, total <= %d)", m, pm.n_regions, CVPCE_POOL_MAX_REGIONS, kMaxRegionsTotal);
    pm.whole = 0;
    int x0[CVPCE_POOL_MAX_REGIONS], x1[CVPCE_POOL_MAX_REGIONS], y0[CVPCE_POOL_MAX_REGIONS], y1[CVPCE_POOL_MAX_REGIONS];
    for (int r = 0; r < pm.n_regions; ++r) {
      int4 q = rg ? make_int4(rg[0], rg[1], rg[2], rg[3]) : make_int4(0, 0, fm.H, fm.W);
      if (rg) rg += 4;
      CVPCE_REQUIRE(q.x >= 0 && q.y >= 0 && q.z > q.x && q.w > q.y && q.z <= fm.H && q.w <= fm.W, CVPCE_ERR_INVALID_ARG,
                    "region_pool: map %d region %d = (%d,%d,%d,%d) is empty or outside the %dx%d map", m, r, q.x, q.y, q.z, q.w, fm.H, fm.W);
      p.regions[rbeg + r] = q;
      y0[r] = q.x; x0[r] = q.y; y1[r] = q.z; x1[r] = q.w;
      if (pm.n_regions == 1 && q.x == 0 && q.y == 0 && q.z == fm.H && q.w == fm.W) pm.whole = 1;
    }
    if (!pm.whole) all_whole = false;
    // elementary row / column segments between region boundaries (each is reduced once by the streaming pass)
    pm.nxs = make_segments(x0, x1, pm.n_regions, pm.xb);
    pm.nys = make_segments(y0, y1, pm.n_regions, pm.yb);
    if (pm.nxs == 0 || pm.nys == 0) pm.nxs = pm.nys = 0;
    for (int r = 0; r < pm.n_regions; ++r)
      p.segs[rbeg + r] = pm.nxs ? make_uchar4((unsigned char)seg_index(pm.yb, pm.nys, y0[r]), (unsigned char)seg_index(pm.yb, pm.nys, y1[r]),
                                              (unsigned char)seg_index(pm.xb, pm.nxs, x0[r]), (unsigned char)seg_index(pm.xb, pm.nxs, x1[r]))
                                : make_uchar4(0, 0, 0, 0);
    rbeg += pm.n_regions;
    pm.out_offset = col;
    col += fm.C;
    pm.slab = (fm.C + kPoolCluster - 1) / kPoolCluster;
    pm.vals_offset = vals;
    vals += pm.n_regions * pm.slab;
    pm.desc_offset = desc;
    desc += pm.slab;
    // channels-last fast path: unit channel stride, every CTA's slab made of whole, 16-byte (f32) / 8-byte (bf16) aligned quads
    const int esz = dtype == CVPCE_F32 ? 4 : 2;
    const uintptr_t align_mask = (uintptr_t)(4 * esz - 1);
    pm.nhwc = (fm.stride_c == 1 && fm.C > 1 && fm.stride_w >= fm.C && fm.C % (4 * kPoolCluster) == 0 && pm.slab <= 4 * kPoolThreads &&
               pm.nxs > 0 && (reinterpret_cast<uintptr_t>(fm.data) & align_mask) == 0 && fm.stride_n % 4 == 0 &&
               fm.stride_h % 4 == 0 && fm.stride_w % 4 == 0)
                  ? 1 : 0;
    if (pm.nhwc) {
      const long long need = (long long)pm.nys * pm.nxs * pm.slab;  // cell maxima
      if (need > 24 * 1024) pm.nhwc = 0;  // segment-maxima scratch capped at 96 KB
      else if (need > scratch) scratch = (int)need;
    }
  }
  CVPCE_REQUIRE(out_stride >= col, CVPCE_ERR_INVALID_ARG, "region_pool: out_stride %lld < descriptor size %d", (long long)out_stride, col);
  p.vals_total = (vals + 3) & ~3;
  p.desc_total = (desc + 3) & ~3;
  p.pairs_total = rbeg;
  p.scratch_floats = (scratch + 3) & ~3;
  const size_t smem = (size_t)(p.vals_total + p.desc_total + p.scratch_floats) * sizeof(float);
  CVPCE_REQUIRE(smem <= 200 * 1024, CVPCE_ERR_UNSUPPORTED, "region_pool: %zu bytes of shared memory needed (regions x channels too large)", smem);
  CVPCE_REQUIRE((long long)B * kPoolCluster < 2147483647ll, CVPCE_ERR_UNSUPPORTED, "region_pool: B too large");

  const bool general = !all_whole || per_region_l2;
  if (general) {
    // region lists / per-region L2 (R-MAC): the persistent TMA-staged kernel (rmac_staged.cu) when the maps are dense and the
    // batch is large enough to give every SM whole images; the cluster kernel below otherwise
    const int rc = rmac_staged_launch(maps, n_maps, dtype, B, regions, n_regions, per_region_l2, final_l2, eps, out, out_stride, stream);
    if (rc != kRmacNotApplicable) return rc;
  }
  bool all_nhwc = true;
  for (int m = 0; m < n_maps; ++m) all_nhwc = all_nhwc && p.maps[m].nhwc;
  // dense channels-last MAC: the pixel-split streaming kernel
  bool pixsplit = !general && col <= kMacMaxCols;
  for (int m = 0; m < n_maps && pixsplit; ++m) {
    const cvpce_feature_map& fm = maps[m];
    const int E = dtype == CVPCE_F32 ? 4 : 8;
    const int V = fm.C / E;
    pixsplit = fm.C % E == 0 && V >= 1 && V <= kPoolThreads && kPoolThreads % V == 0 && fm.stride_c == 1 && fm.stride_w == fm.C &&
               fm.stride_h == (int64_t)fm.W * fm.C && fm.stride_n % E == 0 && (reinterpret_cast<uintptr_t>(fm.data) & 15u) == 0 &&
               p.maps[m].slab * kPoolCluster == fm.C;
  }
  if (pixsplit && workspace && tuning("mac_kernel") != 1) {
    // balanced streaming kernel (needs the caller's workspace; its counters are zero on entry and on exit)
    MacStreamPlan w;
    size_t cb = 0, tb = 0;
    const int G = mac_stream_plan(maps, n_maps, dtype, B, &w, &cb, &tb);
    CVPCE_REQUIRE(workspace_bytes >= tb, CVPCE_ERR_INVALID_ARG, "region_pool: workspace of %zu bytes, %zu needed (cvpce_region_pool_workspace_bytes)", workspace_bytes, tb);
    CVPCE_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255u) == 0, CVPCE_ERR_INVALID_ARG, "region_pool: workspace must be 256-byte aligned");
    w.counters = reinterpret_cast<unsigned*>(workspace);
    w.partial = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + cb);
    auto sk = dtype == CVPCE_F32 ? mac_nhwc_stream_kernel<CVPCE_F32>
                                 : (dtype == CVPCE_BF16 ? mac_nhwc_stream_kernel<CVPCE_BF16> : mac_nhwc_stream_kernel<CVPCE_F16>);
    sk<<<G, kPoolThreads, 0, (cudaStream_t)stream>>>(p, w);
    CVPCE_CUDA_OK(cudaGetLastError());
    return CVPCE_OK;
  }
  void (*kern)(PoolParams);
  size_t smem_launch = smem;
  int cluster_size = kPoolCluster;
  if (pixsplit) {
    // The pixel-split kernel works with any cluster size that divides every C.  Measured on B200 (tools/kbench.py pool,
    // two 512 x 16 x 16 maps per image): about 512 CTAs -- one resident wave, ~0.5 MB streamed per CTA before its
    // reduction tail -- is the sweet spot: B = 256: 0.77 of HBM with clusters of 2 against 0.65 with 4 or 8 (a ragged
    // second wave); B = 128: 0.62 with 4 against 0.52 (8) and 0.50 (2, too few CTAs).  cvpce_tuning_set("mac_cluster", n) overrides.
    const int forced = tuning("mac_cluster");
    const int want = forced ? forced : (B >= 256 ? 2 : (B >= 128 ? 4 : kPoolCluster));
    if (want == 2 || want == 4) {
      bool ok = true;
      for (int m = 0; m < n_maps; ++m) ok = ok && maps[m].C % want == 0;
      if (ok) {
        cluster_size = want;
        for (int m = 0; m < n_maps; ++m) p.maps[m].slab = maps[m].C / want;
      }
    }
    kern = dtype == CVPCE_F32 ? mac_nhwc_pixelsplit_kernel<CVPCE_F32>
                              : (dtype == CVPCE_BF16 ? mac_nhwc_pixelsplit_kernel<CVPCE_BF16> : mac_nhwc_pixelsplit_kernel<CVPCE_F16>);
    smem_launch = 0;
  } else if (dtype == CVPCE_F16) {
    return set_err(CVPCE_ERR_UNSUPPORTED, "region_pool: float16 maps are supported for dense channels-last MAC only (whole-map regions, "
                                          "no per-region L2, C*2/16 dividing 256)");
  } else if (dtype == CVPCE_F32) {
    kern = all_nhwc ? (general ? region_pool_kernel<CVPCE_F32, true, true> : region_pool_kernel<CVPCE_F32, false, true>)
                    : (general ? region_pool_kernel<CVPCE_F32, true, false> : region_pool_kernel<CVPCE_F32, false, false>);
  } else {
    kern = all_nhwc ? (general ? region_pool_kernel<CVPCE_BF16, true, true> : region_pool_kernel<CVPCE_BF16, false, true>)
                    : (general ? region_pool_kernel<CVPCE_BF16, true, false> : region_pool_kernel<CVPCE_BF16, false, false>);
  }
  // the default 48 KB limit covers static + dynamic shared memory: opt in with a margin for the static part
  if (smem_launch > 40 * 1024)
    if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), smem_launch)) return rc;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(B * cluster_size));
  cfg.blockDim = dim3(kPoolThreads);
  cfg.dynamicSmemBytes = smem_launch;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cluster_size;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  CVPCE_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, p));
  return CVPCE_OK;
}

extern "C" int cvpce_region_pool_embed(const cvpce_feature_map* maps, int n_maps, int dtype, int B,
                                       const int32_t* regions, const int32_t* n_regions, int per_region_l2,
                                       int final_l2, float eps, float* out, int64_t out_stride, void* stream) {
  return region_pool_impl(maps, n_maps, dtype, B, regions, n_regions, per_region_l2, final_l2, eps, out, out_stride, nullptr, 0, stream);
}

extern "C" int cvpce_region_pool_embed_ws(const cvpce_feature_map* maps, int n_maps, int dtype, int B,
                                          const int32_t* regions, const int32_t* n_regions, int per_region_l2,
                                          int final_l2, float eps, float* out, int64_t out_stride,
                                          void* workspace, int64_t workspace_bytes, void* stream) {
  return region_pool_impl(maps, n_maps, dtype, B, regions, n_regions, per_region_l2, final_l2, eps, out, out_stride, workspace,
                          workspace_bytes > 0 ? (size_t)workspace_bytes : 0, stream);
}

// ---------------------------------------------------------------------------------------------------------------------
// Allocation-free form of the balanced streaming MAC kernel: its scratch (per-image counters + per-segment maxima) lives in a
// file-scope device array that every GPU loading the library holds until exit, so that callers which cannot call cudaMalloc
// (CUDA graph capture, serving loops, harnesses that forbid it) need no workspace.  Only host code needs the array's
// ADDRESS: the name below is a host-side stand-in there, so the device address is looked up once per device with
// cudaGetSymbolAddress (that device current) and cached.
// ---------------------------------------------------------------------------------------------------------------------
namespace cvpce {

static __device__ __align__(256) unsigned char g_mac_static_scratch[CVPCE_POOL_STATIC_SCRATCH_BYTES];

static std::mutex g_mac_static_mutex;
static void* g_mac_static_addr[64];  // device address of the array per device ordinal (nullptr: not looked up yet)

static int mac_static_scratch(void** out) {
  int dev = 0;
  CVPCE_CUDA_OK(cudaGetDevice(&dev));
  CVPCE_REQUIRE(dev >= 0 && dev < 64, CVPCE_ERR_UNSUPPORTED, "region_pool_static: device ordinal %d outside [0,64)", dev);
  std::lock_guard<std::mutex> lock(g_mac_static_mutex);
  if (!g_mac_static_addr[dev]) {
    void* ptr = nullptr;
    CVPCE_CUDA_OK(cudaGetSymbolAddress(&ptr, g_mac_static_scratch));  // the current device's instance of the array
    CVPCE_REQUIRE(ptr != nullptr, CVPCE_ERR_CUDA, "region_pool_static: cudaGetSymbolAddress returned a null address");
    g_mac_static_addr[dev] = ptr;
  }
  *out = g_mac_static_addr[dev];
  return CVPCE_OK;
}

}  // namespace cvpce

extern "C" int cvpce_region_pool_embed_static(const cvpce_feature_map* maps, int n_maps, int dtype, int B, int final_l2, float eps,
                                              float* out, int64_t out_stride, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(maps && n_maps >= 1 && n_maps <= CVPCE_POOL_MAX_MAPS, CVPCE_ERR_INVALID_ARG, "region_pool_static: n_maps=%d out of [1,%d]", n_maps, CVPCE_POOL_MAX_MAPS);
  CVPCE_REQUIRE(dtype == CVPCE_F32 || dtype == CVPCE_BF16 || dtype == CVPCE_F16, CVPCE_ERR_INVALID_ARG, "region_pool_static: bad dtype %d", dtype);
  CVPCE_REQUIRE(B >= 0, CVPCE_ERR_INVALID_ARG, "region_pool_static: B=%d", B);
  if (B == 0) return CVPCE_OK;
  CVPCE_REQUIRE(mac_dense_nhwc(maps, n_maps, dtype), CVPCE_ERR_UNSUPPORTED,
                "region_pool_static: only dense channels-last MAC maps (unit channel stride, stride_w == C, stride_h == W*C, 16-byte aligned, "
                "C*elemsize/16 dividing 256, sum of C <= %d) run the streaming kernel this entry point exists for", kMacMaxCols);
  MacStreamPlan w;
  size_t counter_bytes = 0, total_bytes = 0;
  mac_stream_plan(maps, n_maps, dtype, B, &w, &counter_bytes, &total_bytes);
  CVPCE_REQUIRE(total_bytes <= (size_t)CVPCE_POOL_STATIC_SCRATCH_BYTES, CVPCE_ERR_UNSUPPORTED,
                "region_pool_static: this batch needs %zu bytes of scratch, the static array holds %d (CVPCE_POOL_STATIC_SCRATCH_BYTES): "
                "split the batch or use cvpce_region_pool_embed_ws", total_bytes, CVPCE_POOL_STATIC_SCRATCH_BYTES);
  if (int rc = require_sm100()) return rc;
  void* scratch = nullptr;
  if (int rc = mac_static_scratch(&scratch)) return rc;
  // the array keeps the last call's contents: the counters must start at zero (a call that failed half-way, or another
  // process-wide user, may have left them otherwise); the segment rows are written before they are read
  CVPCE_CUDA_OK(cudaMemsetAsync(scratch, 0, (size_t)B * sizeof(unsigned), (cudaStream_t)stream));
  return region_pool_impl(maps, n_maps, dtype, B, nullptr, nullptr, 0, final_l2, eps, out, out_stride, scratch,
                          (size_t)CVPCE_POOL_STATIC_SCRATCH_BYTES, stream);
}

extern "C" int cvpce_l2norm_rows(float* x, int B, int D, int64_t stride, float eps, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(B >= 0 && D >= 0, CVPCE_ERR_INVALID_ARG, "l2norm_rows: negative size");
  if (B == 0 || D == 0) return CVPCE_OK;
  CVPCE_REQUIRE(x && stride >= D, CVPCE_ERR_INVALID_ARG, "l2norm_rows: null pointer or stride < D");
  if (int rc = require_sm100()) return rc;
  const int grid = B < 148 * 8 ? B : 148 * 8;
  l2norm_rows_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, B, D, stride, eps);
  CVPCE_LAUNCH_OK("l2norm_rows_kernel");
  return CVPCE_OK;
}
