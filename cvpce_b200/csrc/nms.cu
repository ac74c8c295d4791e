// Batched greedy IoU-NMS for sm_100a: one CTA per image, everything after the first read lives in shared memory.
//
// Follows the semantics of torch.ops.torchvision.nms (CPU op) as reached from
// tv:models/detection/retinanet.py:559-561 / tv:ops/boxes.py:20-48 (see include/cvpce_b200.h).
//
// Per image (segment of the CSR batch):
//   1. stable (score desc, index asc) sort: 64-bit keys, bitonic network in shared memory;
//   2. the boxes are gathered once, in sorted order, into shared memory (float4 loads);
//   3. the sorted list is walked in panels of 64: a 64x64 intra-panel suppression bitmask is built with warp ballots
//      (two 32-bit ballots per row), one warp resolves the panel's keep bits serially over the 64-bit words, and
//      then ALL threads test the not-yet-suppressed tail of the list against the (<= 64) boxes the panel kept.
//      Only kept boxes ever suppress, so the work is O(N*K) IoUs instead of the O(N^2) full bitmask.
// IoU arithmetic uses __f*_rn intrinsics so that nvcc cannot contract mul+add into FMA: the keep decision has to be
// bit-identical to the x86 fp32 evaluation.
#include "common.cuh"
#include <math.h>

namespace cvpce {

constexpr int kNmsPanel = 64;

__device__ __forceinline__ uint32_t desc_score_key(float s) {
  // ascending order of this key == descending score; NaN first (torch.sort(descending=True) puts NaN first)
  if (s != s) return 0u;
  if (s == 0.f) s = 0.f;  // -0 and +0 compare equal in the reference sort
  uint32_t u = __float_as_uint(s);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // larger float -> larger uint
  return ~u;
}

__device__ __forceinline__ bool iou_exceeds(const float4 a, const float4 b, const float thr) {
  const float xx1 = fmaxf(a.x, b.x), yy1 = fmaxf(a.y, b.y);
  const float xx2 = fminf(a.z, b.z), yy2 = fminf(a.w, b.w);
  const float w = fmaxf(0.f, __fsub_rn(xx2, xx1));
  const float h = fmaxf(0.f, __fsub_rn(yy2, yy1));
  const float inter = __fmul_rn(w, h);
  // inter == 0 gives 0, -0 or NaN: none is > thr when thr >= 0 (skips the IEEE division for disjoint pairs)
  if (inter == 0.f && thr >= 0.f) return false;
  const float area_a = __fmul_rn(__fsub_rn(a.z, a.x), __fsub_rn(a.w, a.y));
  const float area_b = __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
  const float uni = __fsub_rn(__fadd_rn(area_a, area_b), inter);
  return __fdiv_rn(inter, uni) > thr;
}

struct NmsParams {
  const float4* boxes;
  const float* scores;
  const int32_t* seg_offsets;  // CSR form: segment b = [seg_offsets[b], seg_offsets[b+1])
  const int32_t* seg_counts;   // strided form (seg_offsets == nullptr): segment b = [b*seg_stride, b*seg_stride + seg_counts[b])
  int seg_stride;
  int max_n;      // capacity of the shared arrays
  int sort_cap;   // power of two >= max_n
  float iou_thr;
  float min_score;
  int max_keep;
  int64_t* keep;
  int32_t* keep_counts;
  const float2* box_scale;
  float4* kept_boxes;
  float* kept_scores;
};

__global__ void __launch_bounds__(1024, 1) nms_batched_kernel(const NmsParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // layout: keys u64[sort_cap] | sbox float4[max_n] | removed u8[max_n] (padded) | static parts below
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem_raw);
  float4* sbox = reinterpret_cast<float4*>(smem_raw + (size_t)p.sort_cap * 8);
  unsigned char* removed = reinterpret_cast<unsigned char*>(sbox + p.max_n);

  __shared__ unsigned long long panel_mask[kNmsPanel];
  __shared__ float4 kbox[kNmsPanel];
  __shared__ int s_panel_kept;
  __shared__ int s_total_kept;
  __shared__ int s_stop;

  const int b = blockIdx.x;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int lane = tid & 31, warp = tid >> 5, nwarps = nthr >> 5;
  const int seg = p.seg_offsets ? p.seg_offsets[b] : b * p.seg_stride;
  const int n = p.seg_offsets ? p.seg_offsets[b + 1] - seg : p.seg_counts[b];
  if (n > p.max_n || n < 0) {
    if (tid == 0) p.keep_counts[b] = -1;
    return;
  }
  if (n == 0) {
    if (tid == 0) p.keep_counts[b] = 0;
    return;
  }
  // power of two >= n (>= 2) for this segment only: small images sort a small network
  int cap = 2;
  while (cap < n) cap <<= 1;

  // ---- 1. keys -------------------------------------------------------------------------------------------------
  for (int i = tid; i < cap; i += nthr) {
    unsigned long long k = ~0ull;
    if (i < n) k = ((unsigned long long)desc_score_key(p.scores[seg + i]) << 32) | (unsigned)i;
    keys[i] = k;
  }
  __syncthreads();
  for (int k = 2; k <= cap; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = tid; t < (cap >> 1); t += nthr) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // insert a 0 bit at position log2(j)
        const int l = i | j;
        const unsigned long long a = keys[i], c = keys[l];
        const bool up = ((i & k) == 0);
        if ((a > c) == up) {
          keys[i] = c;
          keys[l] = a;
        }
      }
      __syncthreads();
    }
  }
  // ---- 2. gather boxes in sorted order -------------------------------------------------------------------------
  for (int i = tid; i < n; i += nthr) {
    const int idx = (int)(keys[i] & 0xffffffffu);
    sbox[i] = __ldg(p.boxes + seg + idx);
    removed[i] = 0;
  }
  if (tid == 0) {
    s_total_kept = 0;
    s_stop = 0;
  }
  __syncthreads();

  const float thr = p.iou_thr;
  const int max_keep = p.max_keep > 0 ? p.max_keep : 0x7fffffff;
  const float2 scale = p.box_scale ? p.box_scale[b] : make_float2(1.f, 1.f);

  // ---- 3. panels -----------------------------------------------------------------------------------------------
  for (int p0 = 0; p0 < n; p0 += kNmsPanel) {
    const int pn = min(kNmsPanel, n - p0);
    // A: intra-panel bitmask rows (bit c of row r set iff c > r and IoU(r,c) > thr)
    for (int r = warp; r < kNmsPanel; r += nwarps) {
      unsigned lo = 0, hi = 0;
      if (r < pn) {  // warp-uniform
        const float4 br = sbox[p0 + r];
        const int c0 = lane, c1 = lane + 32;
        const bool b0 = (c0 > r && c0 < pn) ? iou_exceeds(br, sbox[p0 + c0], thr) : false;
        const bool b1 = (c1 > r && c1 < pn) ? iou_exceeds(br, sbox[p0 + c1], thr) : false;
        lo = __ballot_sync(0xffffffffu, b0);
        hi = __ballot_sync(0xffffffffu, b1);
      }
      if (lane == 0) panel_mask[r] = ((unsigned long long)hi << 32) | lo;
    }
    __syncthreads();
    // B: one warp resolves the panel serially; every lane runs the same scalar recurrence (broadcast LDS)
    if (warp == 0) {
      const int i0 = p0 + lane, i1 = p0 + lane + 32;
      bool r0 = true, r1 = true;  // "removed" (out of range, suppressed earlier, or below min_score)
      float sc0 = 0.f, sc1 = 0.f;
      int id0 = 0, id1 = 0;
      if (lane < pn) {
        id0 = (int)(keys[i0] & 0xffffffffu);
        sc0 = __ldg(p.scores + seg + id0);
        r0 = removed[i0] || !(sc0 > p.min_score || sc0 != sc0);
      }
      if (lane + 32 < pn) {
        id1 = (int)(keys[i1] & 0xffffffffu);
        sc1 = __ldg(p.scores + seg + id1);
        r1 = removed[i1] || !(sc1 > p.min_score || sc1 != sc1);
      }
      unsigned long long rem =
          ((unsigned long long)__ballot_sync(0xffffffffu, r1) << 32) | __ballot_sync(0xffffffffu, r0);
      unsigned long long keepbits = 0;
#pragma unroll 8
      for (int i = 0; i < kNmsPanel; ++i) {
        const unsigned long long m = panel_mask[i];
        if (!((rem >> i) & 1ull)) {
          keepbits |= (1ull << i);
          rem |= m;
        }
      }
      const int base = s_total_kept;
      const int room = max_keep - base;
      const int cnt = __popcll(keepbits);
      const int take = min(cnt, room);
      // sorted scores are descending: once one is <= min_score every later candidate is too
      bool below = false;
      if (pn == kNmsPanel) {
        const float last = __shfl_sync(0xffffffffu, sc1, 31);
        below = !(last > p.min_score || last != last);
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int pos = lane + 32 * h;
        if ((keepbits >> pos) & 1ull) {
          const int rank = __popcll(keepbits & ((1ull << pos) - 1ull));
          if (rank < take) {
            const float4 bx = sbox[p0 + pos];
            kbox[rank] = bx;
            const size_t o = (size_t)seg + base + rank;
            p.keep[o] = (long long)(h ? id1 : id0);
            if (p.kept_boxes)
              p.kept_boxes[o] = make_float4(__fmul_rn(bx.x, scale.x), __fmul_rn(bx.y, scale.y),
                                            __fmul_rn(bx.z, scale.x), __fmul_rn(bx.w, scale.y));
            if (p.kept_scores) p.kept_scores[o] = h ? sc1 : sc0;
          }
        }
      }
      if (lane == 0) {
        s_panel_kept = take;
        s_total_kept = base + take;
        if (base + take >= max_keep || below) s_stop = 1;
      }
    }
    __syncthreads();
    if (s_stop) break;
    // C: the panel's kept boxes suppress the rest of the list
    const int nk = s_panel_kept;
    if (nk > 0) {
      for (int j = p0 + kNmsPanel + tid; j < n; j += nthr) {
        if (removed[j]) continue;
        const float4 bj = sbox[j];
        for (int t = 0; t < nk; ++t) {
          if (iou_exceeds(kbox[t], bj, thr)) {
            removed[j] = 1;
            break;
          }
        }
      }
    }
    __syncthreads();
  }
  if (tid == 0) p.keep_counts[b] = s_total_kept;
}

}  // namespace cvpce

namespace cvpce {

static int launch_nms(NmsParams& p, const float* boxes, const float* scores, int n_images, int max_n, float iou_threshold,
                      float min_score, int max_keep, int64_t* keep, int32_t* keep_counts, const float* box_scale,
                      float* kept_boxes, float* kept_scores, void* stream) {
  CVPCE_REQUIRE(max_n <= CVPCE_NMS_MAX_N, CVPCE_ERR_UNSUPPORTED, "nms: max_n=%d exceeds CVPCE_NMS_MAX_N=%d", max_n, CVPCE_NMS_MAX_N);
  if (int rc = require_sm100()) return rc;
  if (max_n > 0) {
    CVPCE_REQUIRE(boxes && scores && keep, CVPCE_ERR_INVALID_ARG, "nms: boxes / scores / keep is null");
    CVPCE_REQUIRE(aligned16(boxes) && (!kept_boxes || aligned16(kept_boxes)), CVPCE_ERR_INVALID_ARG,
                  "nms: boxes / kept_boxes must be 16-byte aligned");
  }
  p.boxes = reinterpret_cast<const float4*>(boxes);
  p.scores = scores;
  p.max_n = max_n > 0 ? max_n : 1;
  int cap = 2;
  while (cap < p.max_n) cap <<= 1;
  p.sort_cap = cap;
  p.iou_thr = iou_threshold;
  p.min_score = min_score;
  p.max_keep = max_keep;
  p.keep = keep;
  p.keep_counts = keep_counts;
  p.box_scale = reinterpret_cast<const float2*>(box_scale);
  p.kept_boxes = reinterpret_cast<float4*>(kept_boxes);
  p.kept_scores = kept_scores;
  const size_t smem = (size_t)cap * 8 + (size_t)p.max_n * 16 + (((size_t)p.max_n + 15) & ~(size_t)15);
  int threads = cap / 2;
  if (threads < 128) threads = 128;
  if (threads > 1024) threads = 1024;
  static thread_local size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    CVPCE_CUDA_OK(cudaFuncSetAttribute(nms_batched_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  nms_batched_kernel<<<n_images, threads, smem, (cudaStream_t)stream>>>(p);
  CVPCE_LAUNCH_OK("nms_batched_kernel");
  return CVPCE_OK;
}

}  // namespace cvpce

extern "C" int cvpce_nms_batched(const float* boxes, const float* scores, const int32_t* seg_offsets, int n_images,
                                 int max_n, float iou_threshold, float min_score, int max_keep, int64_t* keep,
                                 int32_t* keep_counts, const float* box_scale, float* kept_boxes, float* kept_scores,
                                 void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(n_images >= 0 && max_n >= 0, CVPCE_ERR_INVALID_ARG, "nms: negative size (n_images=%d max_n=%d)", n_images, max_n);
  if (n_images == 0) return CVPCE_OK;
  CVPCE_REQUIRE(seg_offsets && keep_counts, CVPCE_ERR_INVALID_ARG, "nms: seg_offsets / keep_counts is null");
  NmsParams p;
  p.seg_offsets = seg_offsets;
  p.seg_counts = nullptr;
  p.seg_stride = 0;
  return launch_nms(p, boxes, scores, n_images, max_n, iou_threshold, min_score, max_keep, keep, keep_counts, box_scale,
                    kept_boxes, kept_scores, stream);
}

extern "C" int cvpce_nms_segments(const float* boxes, const float* scores, int seg_stride, const int32_t* seg_counts,
                                  int n_images, float iou_threshold, float min_score, int max_keep, int64_t* keep,
                                  int32_t* keep_counts, const float* box_scale, float* kept_boxes, float* kept_scores,
                                  void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(n_images >= 0 && seg_stride >= 0, CVPCE_ERR_INVALID_ARG, "nms: negative size (n_images=%d seg_stride=%d)",
                n_images, seg_stride);
  if (n_images == 0) return CVPCE_OK;
  CVPCE_REQUIRE(seg_counts && keep_counts, CVPCE_ERR_INVALID_ARG, "nms: seg_counts / keep_counts is null");
  CVPCE_REQUIRE((long long)seg_stride * n_images < (1ll << 31), CVPCE_ERR_UNSUPPORTED, "nms: seg_stride * n_images overflows int32");
  NmsParams p;
  p.seg_offsets = nullptr;
  p.seg_counts = seg_counts;
  p.seg_stride = seg_stride;
  return launch_nms(p, boxes, scores, n_images, seg_stride, iou_threshold, min_score, max_keep, keep, keep_counts, box_scale,
                    kept_boxes, kept_scores, stream);
}
