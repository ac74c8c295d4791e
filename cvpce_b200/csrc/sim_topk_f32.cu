// Cosine similarity + per-row top-k, CUDA-core fp32 path, operand preparation and shard merge (sm_100a).
//
// Follows cvpce/models/classification.py:87-95 (distance = 1 - cosine_similarity; nearest_neighbors = argsort of the
// distances, first k).  Ordering contract everywhere: (similarity descending, gallery index ascending).
// This file is the exact-fp32 path used for tiny shapes, unnormalised inputs and as the on-device cross-check of the
// tcgen05 kernel in sim_topk_sm100.cu; the similarity matrix is never written to HBM here either.
#include "common.cuh"
#include <math.h>

namespace cvpce {

// ------------------------------------------------------------------------------------------------------------------
// rows_prepare: fp32 rows -> bf16 rows (optionally L2-normalised) + inverse norms.  One warp per row.
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) rows_prepare_kernel(const float* __restrict__ x, long long R, int D, int normalize,
                                                           float eps, __nv_bfloat16* __restrict__ xb,
                                                           float* __restrict__ inv_norm) {
  const int lane = threadIdx.x & 31;
  const long long warps_total = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < R; row += warps_total) {
    const float* r = x + row * D;
    float inv = 1.f;
    if (normalize) {
      float s = 0.f;
      for (int i = lane; i < D; i += 32) {
        const float v = __ldg(r + i);
        s = fmaf(v, v, s);
      }
      s = warp_sum(s);
      inv = 1.f / fmaxf(sqrtf(s), eps);
    }
    if (inv_norm && lane == 0) inv_norm[row] = inv;
    if (xb) {
      __nv_bfloat16* o = xb + row * D;
      if ((D & 1) == 0) {
        for (int i = lane * 2; i < D; i += 64) {
          const float2 v = *reinterpret_cast<const float2*>(r + i);
          *reinterpret_cast<__nv_bfloat162*>(o + i) = __floats2bfloat162_rn(v.x * inv, v.y * inv);
        }
      } else {
        for (int i = lane; i < D; i += 32) o[i] = __float2bfloat16_rn(r[i] * inv);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------
// rows_prepare_scatter: the same preparation of THIS rank's block of query rows, with the three results (fp32 rows for
// the exact re-score, unit bf16 rows for the tensor cores, inverse norms) stored into the same row slot of EVERY rank's
// query buffers (peer memory over NVLink): "prepare, then all-gather" as one kernel whose stores are the transfer.
// One warp per row, 16-byte loads, 16-byte (fp32) and 8-byte (bf16) stores: a warp writes 512 / 256 contiguous bytes.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kMaxPrepPeers = 16;
struct PrepTargets {
  int n;
  float* f32[kMaxPrepPeers];
  __nv_bfloat16* bf16[kMaxPrepPeers];
  float* inv[kMaxPrepPeers];
};

__global__ void __launch_bounds__(256) rows_prepare_scatter_kernel(const float* __restrict__ x, long long R, int D, int normalize,
                                                                   float eps, const PrepTargets t) {
  const int lane = threadIdx.x & 31;
  const long long warps_total = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < R; row += warps_total) {
    const float4* r = reinterpret_cast<const float4*>(x + row * D);
    float inv = 1.f;
    if (normalize) {
      // the same summation order as rows_prepare_kernel (lane i takes elements i, i + 32, ...): the inverse norm of a row,
      // and with it every re-scored similarity, is bit-identical whether the queries were prepared locally or scattered
      const float* rs = x + row * D;
      float s = 0.f;
      for (int i = lane; i < D; i += 32) {
        const float v = __ldg(rs + i);
        s = fmaf(v, v, s);
      }
      s = warp_sum(s);
      inv = 1.f / fmaxf(sqrtf(s), eps);
    }
    if (lane == 0)
      for (int p = 0; p < t.n; ++p)
        if (t.inv[p]) t.inv[p][row] = inv;
    for (int i = lane; i < (D >> 2); i += 32) {
      const float4 v = __ldg(r + i);  // second pass: L1 / L2 hit
      const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x * inv, v.y * inv), hi = __floats2bfloat162_rn(v.z * inv, v.w * inv);
      const uint2 b = make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
      for (int p = 0; p < t.n; ++p) {
        if (t.f32[p]) reinterpret_cast<float4*>(t.f32[p] + row * D)[i] = v;
        if (t.bf16[p]) reinterpret_cast<uint2*>(t.bf16[p] + row * D)[i] = b;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------
// fp32 similarity + top-k.  CTA = 64 query rows, streams the gallery in 64-row tiles, 4x4 register micro-tiles.
// ------------------------------------------------------------------------------------------------------------------
constexpr int F_BM = 64, F_BN = 64, F_BK = 16, F_THREADS = 256, F_KMAX = 64;

// insert (s, idx) into a descending list of length k kept in shared memory; caller guarantees s > list[k-1]
// or the list is not yet full.  Equal scores keep the earlier (lower) index first because columns arrive ascending.
__device__ __forceinline__ void topk_insert(float* ls, int* li, int k, float s, int idx) {
  int pos = k - 1;
  while (pos > 0 && !(ls[pos - 1] >= s)) {  // NaN-safe: move while previous is strictly smaller (or NaN)
    ls[pos] = ls[pos - 1];
    li[pos] = li[pos - 1];
    --pos;
  }
  ls[pos] = s;
  li[pos] = idx;
}

template <bool NORMALIZE>
__global__ void __launch_bounds__(F_THREADS) sim_topk_f32_kernel(const float* __restrict__ q, const float* __restrict__ g,
                                                                 int Q, int G, int D, int k, long long idx_base, float eps,
                                                                 float* __restrict__ out_s, long long* __restrict__ out_i) {
  __shared__ float As[F_BK][F_BM + 4];
  __shared__ float Bs[F_BK][F_BN + 4];
  __shared__ float Ss[F_BM][F_BN + 1];
  extern __shared__ float dyn[];  // [F_BM][k] scores, then [F_BM][k] indices
  float* tks = dyn;
  int* tki = reinterpret_cast<int*>(dyn + F_BM * k);

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int q0 = blockIdx.x * F_BM;
  for (int i = tid; i < F_BM * k; i += F_THREADS) {
    tks[i] = -INFINITY;
    tki[i] = -1;
  }
  __syncthreads();

  for (int g0 = 0; g0 < G; g0 += F_BN) {
    float acc[4][4] = {};
    float na[4] = {}, nb[4] = {};
    for (int d0 = 0; d0 < D; d0 += F_BK) {
      // load 64x16 of Q and of G, transposed into [k][row]
      for (int i = tid; i < F_BM * F_BK; i += F_THREADS) {
        const int r = i / F_BK, c = i % F_BK;
        const int qr = q0 + r, gr = g0 + r, d = d0 + c;
        As[c][r] = (qr < Q && d < D) ? __ldg(q + (long long)qr * D + d) : 0.f;
        Bs[c][r] = (gr < G && d < D) ? __ldg(g + (long long)gr * D + d) : 0.f;
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < F_BK; ++kk) {
        const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
        const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
        const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
          if (NORMALIZE) {
            na[i] = fmaf(av[i], av[i], na[i]);
            nb[i] = fmaf(bv[i], bv[i], nb[i]);
          }
        }
      }
      __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float s = acc[i][j];
        if (NORMALIZE) s = s / (fmaxf(sqrtf(na[i]), eps) * fmaxf(sqrtf(nb[j]), eps));
        Ss[ty * 4 + i][tx * 4 + j] = s;
      }
    __syncthreads();
    if (tid < F_BM && q0 + tid < Q) {
      float* ls = tks + tid * k;
      int* li = tki + tid * k;
      const int ncol = min(F_BN, G - g0);
      float kth = ls[k - 1];
      for (int c = 0; c < ncol; ++c) {
        const float s = Ss[tid][c];
        if (s > kth || (li[k - 1] < 0 && !(s < kth))) {  // second clause: list not full yet (kth == -inf, s == -inf / NaN)
          topk_insert(ls, li, k, s, g0 + c);
          kth = ls[k - 1];
        }
      }
    }
    __syncthreads();
  }
  for (int i = tid; i < F_BM * k; i += F_THREADS) {
    const int r = i / k, c = i % k;
    if (q0 + r < Q) {
      const int gi = tki[i];
      out_s[(long long)(q0 + r) * k + c] = tks[i];
      out_i[(long long)(q0 + r) * k + c] = gi < 0 ? -1ll : idx_base + gi;
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------
// topk_merge: [W,Q,k] shard results -> [Q,k].  One warp per row, rank-by-counting over the W*k candidates.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kMergeMaxCand = 1024;

__device__ __forceinline__ bool cand_better(float sa, long long ia, float sb, long long ib) {
  // strict total order: higher score first; on equal score lower index first; padding (idx < 0) last
  if ((ia < 0) != (ib < 0)) return ib < 0;
  if (ia < 0) return false;  // both padding: caller breaks the tie by position
  if (sa > sb) return true;
  if (sa < sb) return false;
  return ia < ib;
}

__global__ void __launch_bounds__(128) topk_merge_kernel(const float* __restrict__ scores, const long long* __restrict__ idx,
                                                         int W, int Q, int k, float* __restrict__ out_s,
                                                         long long* __restrict__ out_i) {
  extern __shared__ unsigned char mraw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nc = W * k;
  long long* ci = reinterpret_cast<long long*>(mraw) + (size_t)warp * nc;
  float* cs = reinterpret_cast<float*>(reinterpret_cast<long long*>(mraw) + (size_t)(blockDim.x >> 5) * nc) + (size_t)warp * nc;
  for (int row = blockIdx.x * (blockDim.x >> 5) + warp; row < Q; row += gridDim.x * (blockDim.x >> 5)) {
    for (int c = lane; c < nc; c += 32) {
      const int w = c / k, j = c - w * k;
      const size_t o = ((size_t)w * Q + row) * k + j;
      cs[c] = scores[o];
      ci[c] = idx[o];
    }
    __syncwarp();
    for (int c = lane; c < nc; c += 32) {
      const float s = cs[c];
      const long long i = ci[c];
      int rank = 0;
      for (int o = 0; o < nc; ++o) {
        if (o == c) continue;
        const bool ahead = cand_better(cs[o], ci[o], s, i) || (o < c && !cand_better(s, i, cs[o], ci[o]));
        rank += ahead ? 1 : 0;
      }
      if (rank < k) {
        out_s[(size_t)row * k + rank] = s;
        out_i[(size_t)row * k + rank] = i;
      }
    }
    __syncwarp();
  }
}

}  // namespace cvpce

extern "C" int cvpce_rows_prepare(const float* x, int64_t R, int D, int normalize, float eps, void* x_bf16,
                                  float* inv_norm, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(R >= 0 && D > 0, CVPCE_ERR_INVALID_ARG, "rows_prepare: bad shape R=%lld D=%d", (long long)R, D);
  if (R == 0) return CVPCE_OK;
  CVPCE_REQUIRE(x && (x_bf16 || inv_norm), CVPCE_ERR_INVALID_ARG, "rows_prepare: null pointer");
  CVPCE_REQUIRE((D & 1) || ((reinterpret_cast<uintptr_t>(x) & 7u) == 0 && (reinterpret_cast<uintptr_t>(x_bf16) & 3u) == 0),
                CVPCE_ERR_INVALID_ARG, "rows_prepare: x must be 8-byte and x_bf16 4-byte aligned");
  if (int rc = require_sm100()) return rc;
  long long blocks = (R + 7) / 8;
  if (blocks > 148 * 16) blocks = 148 * 16;
  rows_prepare_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, R, D, normalize, eps,
                                                                          reinterpret_cast<__nv_bfloat16*>(x_bf16), inv_norm);
  CVPCE_LAUNCH_OK("rows_prepare_kernel");
  return CVPCE_OK;
}

extern "C" int cvpce_rows_prepare_scatter(const float* x, int64_t R, int D, int normalize, float eps, float* const* peer_f32,
                                          void* const* peer_bf16, float* const* peer_inv_norm, int n_peers, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(R >= 0 && D > 0, CVPCE_ERR_INVALID_ARG, "rows_prepare_scatter: bad shape R=%lld D=%d", (long long)R, D);
  CVPCE_REQUIRE(D % 4 == 0, CVPCE_ERR_UNSUPPORTED, "rows_prepare_scatter: D=%d must be a multiple of 4", D);
  CVPCE_REQUIRE(n_peers >= 1 && n_peers <= kMaxPrepPeers, CVPCE_ERR_INVALID_ARG, "rows_prepare_scatter: n_peers=%d outside [1,%d]",
                n_peers, kMaxPrepPeers);
  if (R == 0) return CVPCE_OK;
  CVPCE_REQUIRE(x && aligned16(x), CVPCE_ERR_INVALID_ARG, "rows_prepare_scatter: x null or not 16-byte aligned");
  PrepTargets t;
  t.n = n_peers;
  for (int p = 0; p < n_peers; ++p) {
    t.f32[p] = peer_f32 ? peer_f32[p] : nullptr;
    t.bf16[p] = peer_bf16 ? reinterpret_cast<__nv_bfloat16*>(peer_bf16[p]) : nullptr;
    t.inv[p] = peer_inv_norm ? peer_inv_norm[p] : nullptr;
    CVPCE_REQUIRE(aligned16(t.f32[p]) && (reinterpret_cast<uintptr_t>(t.bf16[p]) & 7u) == 0, CVPCE_ERR_INVALID_ARG,
                  "rows_prepare_scatter: peer %d buffers must be 16-byte (fp32) / 8-byte (bf16) aligned", p);
  }
  if (int rc = require_sm100()) return rc;
  long long blocks = (R + 7) / 8;
  if (blocks > 148 * 16) blocks = 148 * 16;
  rows_prepare_scatter_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, R, D, normalize, eps, t);
  CVPCE_LAUNCH_OK("rows_prepare_scatter_kernel");
  return CVPCE_OK;
}

extern "C" int cvpce_sim_topk_f32(const float* q, const float* g, int Q, int G, int D, int k, int64_t idx_base,
                                  int normalize, float eps, float* out_scores, int64_t* out_idx, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(Q >= 0 && G >= 0 && D > 0, CVPCE_ERR_INVALID_ARG, "sim_topk_f32: bad shape Q=%d G=%d D=%d", Q, G, D);
  CVPCE_REQUIRE(k >= 1 && k <= F_KMAX, CVPCE_ERR_UNSUPPORTED, "sim_topk_f32: k=%d outside [1,%d]", k, F_KMAX);
  if (Q == 0) return CVPCE_OK;
  CVPCE_REQUIRE(q && (g || G == 0) && out_scores && out_idx, CVPCE_ERR_INVALID_ARG, "sim_topk_f32: null pointer");
  if (int rc = require_sm100()) return rc;
  const size_t smem = (size_t)F_BM * k * 8;
  const int grid = (Q + F_BM - 1) / F_BM;
  if (smem > 16 * 1024) {
    if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(&sim_topk_f32_kernel<true>), 64 * 1024)) return rc;
    if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(&sim_topk_f32_kernel<false>), 64 * 1024)) return rc;
  }
  if (normalize)
    sim_topk_f32_kernel<true><<<grid, F_THREADS, smem, (cudaStream_t)stream>>>(q, g, Q, G, D, k, idx_base, eps, out_scores,
                                                                              reinterpret_cast<long long*>(out_idx));
  else
    sim_topk_f32_kernel<false><<<grid, F_THREADS, smem, (cudaStream_t)stream>>>(q, g, Q, G, D, k, idx_base, eps, out_scores,
                                                                               reinterpret_cast<long long*>(out_idx));
  CVPCE_LAUNCH_OK("sim_topk_f32_kernel");
  return CVPCE_OK;
}

extern "C" int cvpce_topk_merge(const float* scores, const int64_t* idx, int W, int Q, int k, float* out_scores,
                                int64_t* out_idx, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(W >= 1 && Q >= 0 && k >= 1, CVPCE_ERR_INVALID_ARG, "topk_merge: bad shape W=%d Q=%d k=%d", W, Q, k);
  CVPCE_REQUIRE((long long)W * k <= kMergeMaxCand, CVPCE_ERR_UNSUPPORTED, "topk_merge: W*k=%d exceeds %d", W * k, kMergeMaxCand);
  if (Q == 0) return CVPCE_OK;
  CVPCE_REQUIRE(scores && idx && out_scores && out_idx, CVPCE_ERR_INVALID_ARG, "topk_merge: null pointer");
  if (int rc = require_sm100()) return rc;
  const int warps = 4;
  const size_t smem = (size_t)warps * W * k * 12;
  int grid = (Q + warps - 1) / warps;
  if (grid > 148 * 16) grid = 148 * 16;
  if (smem > 48 * 1024)
    if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(&topk_merge_kernel), smem)) return rc;
  topk_merge_kernel<<<grid, warps * 32, smem, (cudaStream_t)stream>>>(scores, reinterpret_cast<const long long*>(idx), W, Q, k,
                                                                      out_scores, reinterpret_cast<long long*>(out_idx));
  CVPCE_LAUNCH_OK("topk_merge_kernel");
  return CVPCE_OK;
}
