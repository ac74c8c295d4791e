// Region max-pool embedding (MAC / R-MAC) + L2 normalisation for sm_100a.
//
// Follows cvpce/models/classification.py:46-51 (MACVGG: amax over H,W of relu4_3 and relu5_3, cat, L2 with
// clamp(min=eps)) and :79-85 (MACResNet).  R-MAC (regions + per-region L2 + sum) is the north-star extension.
//
// One thread-block CLUSTER (8 CTAs) per crop: CTA r of the cluster owns the r-th channel slab of every feature map,
// streams those planes once from HBM (128-bit loads, one warp per plane, warp-shuffle max), keeps the region maxima
// in shared memory, and the norms that span all channels (per-region L2 and the final L2) are all-reduced across the
// cluster through distributed shared memory -- no workspace, no second launch, the descriptor is written once.
// With several regions per map the plane is staged in shared memory and every region is evaluated from there.
#include "common.cuh"
#include <cooperative_groups.h>
#include <math.h>

namespace cg = cooperative_groups;

namespace cvpce {

constexpr int kPoolThreads = 256;
constexpr int kPoolWarps = kPoolThreads / 32;
constexpr int kPlaneCap = 1024;  // floats staged per warp (32x32 map)
constexpr int kPoolCluster = 8;
constexpr int kMaxRegionsTotal = 64;

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
};

struct PoolParams {
  PoolMap maps[CVPCE_POOL_MAX_MAPS];
  int4 regions[kMaxRegionsTotal];  // (y0,x0,y1,x1)
  int n_maps, B;
  int per_region_l2, final_l2;
  float eps;
  float* out;
  long long out_stride;
  int vals_total, desc_total, pairs_total;
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
__device__ __forceinline__ float bf16lo(unsigned u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(unsigned u) { return __uint_as_float(u & 0xffff0000u); }

template <int DTYPE>
__global__ void __launch_bounds__(kPoolThreads) region_pool_kernel(const __grid_constant__ PoolParams p) {
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned crank = cluster.block_rank();
  const unsigned csize = cluster.num_blocks();
  const int n = blockIdx.x / csize;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  extern __shared__ __align__(16) float smem[];
  float* vals = smem;                         // [vals_total]
  float* desc = vals + p.vals_total;          // [desc_total]
  float* planes = desc + p.desc_total;        // [kPoolWarps][kPlaneCap]
  __shared__ float part[kMaxRegionsTotal];    // partial sum of squares per (map,region), read by the cluster
  __shared__ float denom[kMaxRegionsTotal];
  __shared__ float fpart;                     // partial sum of squares of the descriptor slab
  __shared__ float wsum[kPoolWarps];

  // ---- phase 1: region maxima of this CTA's channel slabs ----------------------------------------------------------
  for (int m = 0; m < p.n_maps; ++m) {
    const PoolMap& mp = p.maps[m];
    const int c_lo = (int)crank * mp.slab;
    const int nplanes = max(0, min(mp.slab, mp.C - c_lo));
    const int HW = mp.H * mp.W;
    const bool contiguous = (mp.sw == 1 && mp.sh == mp.W);
    constexpr int VEC = (DTYPE == CVPCE_F32) ? 4 : 8;
    constexpr int ESZ = (DTYPE == CVPCE_F32) ? 4 : 2;
    for (int pc = warp; pc < nplanes; pc += kPoolWarps) {
      const long long poff = (long long)n * mp.sn + (long long)(c_lo + pc) * mp.sc;
      const char* pbase = reinterpret_cast<const char*>(mp.data) + poff * ESZ;
      const bool vec_ok = contiguous && (HW % VEC == 0) && ((reinterpret_cast<uintptr_t>(pbase) & 15u) == 0);
      float* vout = vals + mp.vals_offset + pc;
      if (mp.whole) {
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
        v = warp_max(v);
        if (lane == 0) vout[0] = v;
      } else {
        const bool staged = HW <= kPlaneCap;
        float* pl = planes + warp * kPlaneCap;
        if (staged) {
          if (vec_ok && DTYPE == CVPCE_F32) {
            const float4* src = reinterpret_cast<const float4*>(pbase);
            for (int i = lane; i < HW / 4; i += 32) reinterpret_cast<float4*>(pl)[i] = ldg_stream_f4(src + i);
          } else {
            for (int i = lane; i < HW; i += 32) {
              const int y = i / mp.W, x = i - y * mp.W;
              pl[i] = load_elem<DTYPE>(mp.data, poff + y * mp.sh + x * mp.sw);
            }
          }
          __syncwarp();
        }
        for (int r = 0; r < mp.n_regions; ++r) {
          const int4 rg = p.regions[mp.region_begin + r];
          const int rw = rg.w - rg.y, rh = rg.z - rg.x;
          float v = -INFINITY;
          for (int i = lane; i < rw * rh; i += 32) {
            const int y = rg.x + i / rw, x = rg.y + i % rw;
            v = fmaxf(v, staged ? pl[y * mp.W + x] : load_elem<DTYPE>(mp.data, poff + y * mp.sh + x * mp.sw));
          }
          v = warp_max(v);
          if (lane == 0) vout[r * mp.slab] = v;
        }
        __syncwarp();
      }
    }
  }
  __syncthreads();

  // ---- phase 2: per-region L2 norms over ALL channels (cluster all-reduce through DSMEM) -----------------------------
  if (p.per_region_l2) {
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
        if (p.per_region_l2) v = v / denom[mp.region_begin + r];
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

extern "C" int cvpce_region_pool_embed(const cvpce_feature_map* maps, int n_maps, int dtype, int B,
                                       const int32_t* regions, const int32_t* n_regions, int per_region_l2,
                                       int final_l2, float eps, float* out, int64_t out_stride, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(maps && n_maps >= 1 && n_maps <= CVPCE_POOL_MAX_MAPS, CVPCE_ERR_INVALID_ARG, "region_pool: n_maps=%d out of [1,%d]", n_maps, CVPCE_POOL_MAX_MAPS);
  CVPCE_REQUIRE(dtype == CVPCE_F32 || dtype == CVPCE_BF16, CVPCE_ERR_INVALID_ARG, "region_pool: bad dtype %d", dtype);
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
  int rbeg = 0, col = 0, vals = 0, desc = 0;
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
    for (int r = 0; r < pm.n_regions; ++r) {
      int4 q = rg ? make_int4(rg[0], rg[1], rg[2], rg[3]) : make_int4(0, 0, fm.H, fm.W);
      if (rg) rg += 4;
      CVPCE_REQUIRE(q.x >= 0 && q.y >= 0 && q.z > q.x && q.w > q.y && q.z <= fm.H && q.w <= fm.W, CVPCE_ERR_INVALID_ARG,
                    "region_pool: map %d region %d = (%d,%d,%d,%d) is empty or outside the %dx%d map", m, r, q.x, q.y, q.z, q.w, fm.H, fm.W);
      p.regions[rbeg + r] = q;
      if (pm.n_regions == 1 && q.x == 0 && q.y == 0 && q.z == fm.H && q.w == fm.W) pm.whole = 1;
    }
    rbeg += pm.n_regions;
    pm.out_offset = col;
    col += fm.C;
    pm.slab = (fm.C + kPoolCluster - 1) / kPoolCluster;
    pm.vals_offset = vals;
    vals += pm.n_regions * pm.slab;
    pm.desc_offset = desc;
    desc += pm.slab;
  }
  CVPCE_REQUIRE(out_stride >= col, CVPCE_ERR_INVALID_ARG, "region_pool: out_stride %lld < descriptor size %d", (long long)out_stride, col);
  p.vals_total = (vals + 3) & ~3;
  p.desc_total = (desc + 3) & ~3;
  p.pairs_total = rbeg;
  const size_t smem = (size_t)(p.vals_total + p.desc_total + kPoolWarps * kPlaneCap) * sizeof(float);
  CVPCE_REQUIRE(smem <= 200 * 1024, CVPCE_ERR_UNSUPPORTED, "region_pool: %zu bytes of shared memory needed (regions x channels too large)", smem);
  CVPCE_REQUIRE((long long)B * kPoolCluster < 2147483647ll, CVPCE_ERR_UNSUPPORTED, "region_pool: B too large");

  auto kern = dtype == CVPCE_F32 ? region_pool_kernel<CVPCE_F32> : region_pool_kernel<CVPCE_BF16>;
  if (smem > 48 * 1024) CVPCE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(B * kPoolCluster));
  cfg.blockDim = dim3(kPoolThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kPoolCluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  CVPCE_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, p));
  return CVPCE_OK;
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
