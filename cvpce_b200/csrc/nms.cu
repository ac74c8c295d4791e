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

// development aid (build with CVPCE_NVCC_EXTRA=-DCVPCE_PHASE_TIMERS): cycles per phase of block 0, thread 0
#ifdef CVPCE_PHASE_TIMERS
__device__ long long g_nms_phase[8];
#define PHASE_INIT long long _t_last = clock64();
#define PHASE_COUNT g_nms_phase[5] += 1;
#define PHASE_MARK(k)                                         \
  if (blockIdx.x == 0 && threadIdx.x == 0) {                  \
    const long long _now = clock64();                         \
    g_nms_phase[k] += _now - _t_last;                         \
    _t_last = _now;                                           \
  }
#else
#define PHASE_INIT
#define PHASE_MARK(k)
#define PHASE_COUNT ;
#endif

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
  int region_a;   // bytes of the shared region that holds the sort keys, then the grid
  float size_ratio; // a kept box b can only suppress a if w_b > size_ratio * w_a (and likewise in y)
  float win_scale;  // half-extent of a candidate's search window in units of its own width / height
  int brute;      // threshold < 0 (or NaN): disjoint boxes suppress each other, every cell is visited
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

constexpr int kNmsGrid = 32;      // finest level: kNmsGrid x kNmsGrid cells over the range of box centres
constexpr int kNmsLevels = 6;     // level l: (32 >> l)^2 cells, holds kept boxes of (normalised) size [2^l, 2^(l+1)) fine cells
constexpr int kNmsCells = 1024 + 256 + 64 + 16 + 4 + 1;
constexpr int kNmsCellCap = 16;   // kept boxes per cell; the rest go to the overflow list
__device__ __forceinline__ int nms_level_base(int l) { return (4096 - (4096 >> (2 * l))) / 3; }  // 0,1024,1280,1344,1360,1364
// size class of a box whose larger side measures v fine cells (monotone in v)
__device__ __forceinline__ int nms_level(float v) {
  if (!(v >= 2.f)) return 0;
  if (v >= 32.f) return kNmsLevels - 1;
  return ilogbf(v);
}
// bytes of the region that holds the sort keys first and the grid afterwards
__host__ __device__ constexpr size_t nms_grid_bytes(int max_n) {
  return ((size_t)kNmsCells * 4 + (size_t)kNmsCells * kNmsCellCap * 2 + (size_t)max_n * 2 + 15) & ~(size_t)15;
}

__device__ __forceinline__ bool finite4(const float4 b) {
  return (fabsf(b.x) <= 3.0e38f) && (fabsf(b.y) <= 3.0e38f) && (fabsf(b.z) <= 3.0e38f) && (fabsf(b.w) <= 3.0e38f);
}

__global__ void __launch_bounds__(1024, 1) nms_batched_kernel(const NmsParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // layout: region A = keys u64[sort_cap] during the sort, then {cell_count i32[cells] | cell_entry u16[cells*cap] (6 levels, 32x32 .. 1x1) |
  // overflow u16[max_n]}; sbox float4[max_n]; orig u16[max_n] (index of sorted position i in the segment)
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem_raw);
  int* cell_count = reinterpret_cast<int*>(smem_raw);
  unsigned short* cell_entry = reinterpret_cast<unsigned short*>(smem_raw + (size_t)kNmsCells * 4);
  unsigned short* overflow = cell_entry + (size_t)kNmsCells * kNmsCellCap;
  float4* sbox = reinterpret_cast<float4*>(smem_raw + p.region_a);
  unsigned short* orig = reinterpret_cast<unsigned short*>(sbox + p.max_n);

  __shared__ unsigned long long panel_mask[kNmsPanel];
  __shared__ float4 kbox[kNmsPanel];
  __shared__ int suppressed[kNmsPanel];
  __shared__ float s_red[4][32];
  __shared__ int s_total_kept, s_stop, s_overflow;

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

  PHASE_INIT
  // ---- 1. keys -------------------------------------------------------------------------------------------------
  for (int i = tid; i < cap; i += nthr) {
    unsigned long long k = ~0ull;
    if (i < n) k = ((unsigned long long)desc_score_key(p.scores[seg + i]) << 32) | (unsigned)i;
    keys[i] = k;
  }
  __syncthreads();
  block_bitonic_sort_u64(keys, cap);  // common.cuh: strides <= 32 in registers, larger ones fused three at a time
  PHASE_MARK(0)
  // ---- 2. gather boxes in sorted order; range of the box centres -----------------------------------------------------
  float cx_lo = 3.0e38f, cx_hi = -3.0e38f, cy_lo = 3.0e38f, cy_hi = -3.0e38f;
  for (int i = tid; i < n; i += nthr) {
    const int idx = (int)(keys[i] & 0xffffffffu);
    const float4 bx = __ldg(p.boxes + seg + idx);
    sbox[i] = bx;
    orig[i] = (unsigned short)idx;
    if (finite4(bx)) {
      const float cx = 0.5f * (bx.x + bx.z), cy = 0.5f * (bx.y + bx.w);
      cx_lo = fminf(cx_lo, cx);
      cx_hi = fmaxf(cx_hi, cx);
      cy_lo = fminf(cy_lo, cy);
      cy_hi = fmaxf(cy_hi, cy);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    cx_lo = fminf(cx_lo, __shfl_xor_sync(0xffffffffu, cx_lo, o));
    cx_hi = fmaxf(cx_hi, __shfl_xor_sync(0xffffffffu, cx_hi, o));
    cy_lo = fminf(cy_lo, __shfl_xor_sync(0xffffffffu, cy_lo, o));
    cy_hi = fmaxf(cy_hi, __shfl_xor_sync(0xffffffffu, cy_hi, o));
  }
  if (lane == 0) {
    s_red[0][warp] = cx_lo;
    s_red[1][warp] = cx_hi;
    s_red[2][warp] = cy_lo;
    s_red[3][warp] = cy_hi;
  }
  __syncthreads();  // also: every key has been read, region A can become the grid
  for (int w = 0; w < nwarps; ++w) {
    cx_lo = fminf(cx_lo, s_red[0][w]);
    cx_hi = fmaxf(cx_hi, s_red[1][w]);
    cy_lo = fminf(cy_lo, s_red[2][w]);
    cy_hi = fmaxf(cy_hi, s_red[3][w]);
  }
  // cell of a coordinate: monotone in the coordinate, so a centre inside a window lies in a cell inside the window's range
  const float inv_x = (float)kNmsGrid / fmaxf(cx_hi - cx_lo, 1e-30f), inv_y = (float)kNmsGrid / fmaxf(cy_hi - cy_lo, 1e-30f);
  auto cell_x = [&](float x) { return min(kNmsGrid - 1, max(0, (int)fminf(fmaxf((x - cx_lo) * inv_x, 0.f), (float)kNmsGrid))); };
  auto cell_y = [&](float y) { return min(kNmsGrid - 1, max(0, (int)fminf(fmaxf((y - cy_lo) * inv_y, 0.f), (float)kNmsGrid))); };
  for (int i = tid; i < kNmsCells; i += nthr) cell_count[i] = 0;
  if (tid < kNmsPanel) suppressed[tid] = 0;
  if (tid == 0) {
    s_total_kept = 0;
    s_stop = 0;
    s_overflow = 0;
  }
  __syncthreads();

  PHASE_MARK(1)
  const float thr = p.iou_thr;
  const int max_keep = p.max_keep > 0 ? p.max_keep : 0x7fffffff;
  const float2 scale = p.box_scale ? p.box_scale[b] : make_float2(1.f, 1.f);
  const int T = nthr / kNmsPanel;  // threads that share one candidate of the panel (2..16)
  const float abs_margin_x = 1e-6f * (cx_hi - cx_lo), abs_margin_y = 1e-6f * (cy_hi - cy_lo);

  // ---- 3. panels of 64 candidates in sorted order -------------------------------------------------------------------
  for (int p0 = 0; p0 < n; p0 += kNmsPanel) {
    const int pn = min(kNmsPanel, n - p0);
    // the resolving warp needs the panel's scores in step C: issue the (L2-latency) loads now, they land during A and B
    float sc0 = 0.f, sc1 = 0.f;
    int id0 = 0, id1 = 0;
    if (warp == 0) {
      if (lane < pn) {
        id0 = orig[p0 + lane];
        sc0 = __ldg(p.scores + seg + id0);
      }
      if (lane + 32 < pn) {
        id1 = orig[p0 + lane + 32];
        sc1 = __ldg(p.scores + seg + id1);
      }
    }
    // A: is the candidate suppressed by a box kept in an EARLIER panel?  Only kept boxes whose centre lies in the
    //    candidate's window can exceed the threshold (header comment), so only those grid cells are visited.
    {
      const int c = tid / T, sub = tid - c * T;
      if (c < pn) {
        const float4 a = sbox[p0 + c];
        const float w = a.z - a.x, h = a.w - a.y;
        const bool full = p.brute || !finite4(a);
        if (full || (w > 0.f && h > 0.f)) {  // an empty box has no intersection: it is never suppressed
          // Kept boxes are filed by size class (level) and centre cell.  b can only suppress a if every side of b is
          // within (t, 1/t) of a's, so a visits the levels of sizes t*m_a .. m_a/t only, and in each of them the
          // cells its window covers: a bounded number of cells whatever the mix of box sizes.
          int x0 = 0, x1 = kNmsGrid - 1, y0 = 0, y1 = kNmsGrid - 1, l0 = 0, l1 = kNmsLevels - 1;
          if (!full) {
            const float cxa = 0.5f * (a.x + a.z), cya = 0.5f * (a.y + a.w);
            const float hx = w * p.win_scale + abs_margin_x + 1e-6f * fabsf(cxa);
            const float hy = h * p.win_scale + abs_margin_y + 1e-6f * fabsf(cya);
            x0 = cell_x(cxa - hx);
            x1 = cell_x(cxa + hx);
            y0 = cell_y(cya - hy);
            y1 = cell_y(cya + hy);
            const float m = fmaxf(w * inv_x, h * inv_y);
            l0 = nms_level(m * p.size_ratio);
            l1 = nms_level(m / p.size_ratio);
          }
          bool hit = false;
          for (int l = l0; l <= l1 && !hit; ++l) {
            const int g = kNmsGrid >> l, X0 = x0 >> l, Y0 = y0 >> l;
            const int ncx = (x1 >> l) - X0 + 1, ncells = ncx * ((y1 >> l) - Y0 + 1);
            const int base = nms_level_base(l) + Y0 * g + X0;
            for (int q = sub; q < ncells && !hit; q += T) {
              const int qy = q / ncx;
              const int cell = base + qy * g + (q - qy * ncx);
              const int cnt = min(cell_count[cell], kNmsCellCap);
              for (int e = 0; e < cnt; ++e) {
                if (iou_exceeds(sbox[cell_entry[cell * kNmsCellCap + e]], a, thr)) {
                  hit = true;
                  break;
                }
              }
            }
          }
          const int nover = s_overflow;
          for (int e = sub; e < nover && !hit; e += T)
            if (iou_exceeds(sbox[overflow[e]], a, thr)) hit = true;
          if (hit) suppressed[c] = 1;
        }
      }
    }
    __syncthreads();
    PHASE_MARK(2)
    // B: intra-panel bitmask rows (bit c of row r set iff c > r and IoU(r,c) > thr)
    for (int r = warp; r < kNmsPanel; r += nwarps) {
      unsigned lo = 0, hi = 0;
      if (r < pn && !suppressed[r]) {  // warp-uniform
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
    PHASE_MARK(3)
    // C: one warp resolves the panel serially (every lane runs the same scalar recurrence), writes the kept rows and
    //    files the newly kept boxes in the grid
    if (warp == 0) {
      bool r0 = true, r1 = true;  // "removed" (out of range, suppressed earlier, or below min_score)
      if (lane < pn) r0 = suppressed[lane] || !(sc0 > p.min_score || sc0 != sc0);
      if (lane + 32 < pn) r1 = suppressed[lane + 32] || !(sc1 > p.min_score || sc1 != sc1);
      // rem: bit i set = candidate i is out (out of range, suppressed by an earlier panel or an earlier kept candidate of
      // this panel, below min_score).  Row masks only have bits c > r, so bit i is final once step i has run and
      // kept = ~rem at the end.  32-bit halves with compile-time bit positions: two dependent instructions per step.
      unsigned rem_lo = __ballot_sync(0xffffffffu, r0), rem_hi = __ballot_sync(0xffffffffu, r1);
      const uint2* mask2 = reinterpret_cast<const uint2*>(panel_mask);
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const uint2 m = mask2[i];
        if (!(rem_lo & (1u << i))) {
          rem_lo |= m.x;
          rem_hi |= m.y;
        }
      }
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const uint2 m = mask2[32 + i];
        if (!(rem_hi & (1u << i))) rem_hi |= m.y;
      }
      const unsigned long long keepbits = ~(((unsigned long long)rem_hi << 32) | rem_lo);
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
            const size_t o = (size_t)seg + base + rank;
            p.keep[o] = (long long)(h ? id1 : id0);
            if (p.kept_boxes)
              p.kept_boxes[o] = make_float4(__fmul_rn(bx.x, scale.x), __fmul_rn(bx.y, scale.y),
                                            __fmul_rn(bx.z, scale.x), __fmul_rn(bx.w, scale.y));
            if (p.kept_scores) p.kept_scores[o] = h ? sc1 : sc0;
            // grid insertion: by centre cell; boxes that cannot intersect anything are left out
            const float w = bx.z - bx.x, hgt = bx.w - bx.y;
            if (p.brute || !(w <= 0.f || hgt <= 0.f)) {
              int slot = kNmsCellCap;
              int cell = 0;
              if (finite4(bx)) {
                const int l = nms_level(fmaxf(w * inv_x, hgt * inv_y));
                cell = nms_level_base(l) + (cell_y(0.5f * (bx.y + bx.w)) >> l) * (kNmsGrid >> l) + (cell_x(0.5f * (bx.x + bx.z)) >> l);
                slot = atomicAdd(&cell_count[cell], 1);
              }
              if (slot < kNmsCellCap)
                cell_entry[cell * kNmsCellCap + slot] = (unsigned short)(p0 + pos);
              else
                overflow[atomicAdd(&s_overflow, 1)] = (unsigned short)(p0 + pos);
            }
          }
        }
      }
      suppressed[lane] = 0;
      suppressed[lane + 32] = 0;
      if (lane == 0) {
        s_total_kept = base + take;
        if (base + take >= max_keep || below) s_stop = 1;
      }
    }
    __syncthreads();
    PHASE_MARK(4)
    if (blockIdx.x == 0 && tid == 0) PHASE_COUNT
    if (s_stop) break;
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
  size_t region_a = (size_t)cap * 8;
  if (region_a < nms_grid_bytes(p.max_n)) region_a = nms_grid_bytes(p.max_n);
  p.region_a = (int)region_a;
  // IoU > t needs the overlap to exceed t * max(w_a, w_b) in x (and likewise in y), which bounds the distance of the two
  // centres by w_a * max(0.5, 0.5/t - 0.5); 1/64 of slack covers the fp32 rounding of the IoU and of the centres
  p.brute = !(iou_threshold >= 0.f);
  p.win_scale = p.brute ? 0.f : (iou_threshold >= 0.5f ? 0.5f : (iou_threshold > 1e-3f ? 0.5f / iou_threshold - 0.5f : INFINITY)) * (1.f + 1.f / 64);
  if (!p.brute && !(iou_threshold > 1e-3f)) p.brute = 1;
  p.size_ratio = p.brute ? 0.f : fminf(iou_threshold, 1.f) * (1.f - 1.f / 64);
  const size_t smem = region_a + (size_t)p.max_n * 16 + (((size_t)p.max_n * 2 + 15) & ~(size_t)15);
  int threads = cap / 2;
  if (threads < 128) threads = 128;
  if (threads > 1024) threads = 1024;
  if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(&nms_batched_kernel), smem)) return rc;
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

#ifdef CVPCE_PHASE_TIMERS
extern "C" __attribute__((visibility("default"))) int cvpce_debug_nms_phases(long long* out8, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out8, cvpce::g_nms_phase, sizeof(long long) * 8);
  if (reset) {
    long long z[8] = {0};
    cudaMemcpyToSymbol(cvpce::g_nms_phase, z, sizeof(z));
  }
  return 0;
}
#endif
