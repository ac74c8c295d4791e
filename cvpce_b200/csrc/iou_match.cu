// Pairwise box IoU and the greedy IoU claim pass that follows it in the reference, for sm_100a.
// Replaces: tv:ops/boxes.py box_iou as called by cvpce/metrics.py:11-13 (iou_matrices) and cvpce/planograms.py:224, and
//           the Python double loops that consume the sorted matrix: cvpce/metrics.py:15-34 (check_matches, detection
//           evaluation) and cvpce/planograms.py:214-232 (expected-vs-detected matching inside finalize_via_ransac, the
//           step that decides which planogram positions go to the second-round re-classification of production.py:101-115).
//
// What those loops compute (neither stops after its first hit): rows are visited in index order; row i walks its columns by
// descending IoU until the first  iou < thr  and marks EVERY column not yet marked; the row counts as matched iff it marked
// at least one.  The walk order inside a row therefore does not matter and the result is
//     claim_i = { j : !(iou_ij < thr) } \ used ;  used |= claim_i ;  row_matched_i = claim_i != {}
// (NaN IoU -- 0/0 of two empty boxes -- is not "< thr" and sorts first in torch's descending sort, so it claims.)  With
// labels (planograms.py:214-224) a column is eligible only when its label equals the row's.  `thr` is compared in fp32
// like torch's 0-dim-tensor-vs-Python-float comparison.
//
// IoU is torchvision's: area = (x2-x1)*(y2-y1); wh = max(min(rb)-max(lt), 0); inter = w*h; union = (area_a+area_b)-inter;
// inter/union in IEEE fp32, no contraction => the matrix is bit-equal to torchvision's.
//
// box_iou_kernel: one thread = 4 consecutive columns x 16 rows (column boxes and areas in registers, one 16-byte streaming
// store per row when the row pitch allows), HBM-bound on the n_a*n_b*4-byte matrix.
// iou_match_kernel: one CTA per problem (image); a warp computes the claim bit-row of one row with ballots (8 rows per
// panel), then one warp resolves the panel against the `used` bit set -- the matrix is never written to memory.
#include "common.cuh"

namespace cvpce {

__device__ __forceinline__ float box_area_tv(const float4 b) { return __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y)); }

// Disjoint pairs (inter == +0) are the common case and a zero dividend sends the IEEE division down its slow path: the
// divider is fed 1 instead and the quotient of 0/union is written out by its IEEE rule (+0, -0 or NaN by the sign of union).
__device__ __forceinline__ float iou_tv(const float4 a, const float area_a, const float4 b, const float area_b) {
  const float w = fmaxf(__fsub_rn(fminf(a.z, b.z), fmaxf(a.x, b.x)), 0.f);
  const float h = fmaxf(__fsub_rn(fminf(a.w, b.w), fmaxf(a.y, b.y)), 0.f);
  const float inter = __fmul_rn(w, h);
  const float uni = __fsub_rn(__fadd_rn(area_a, area_b), inter);
  const bool zero = inter == 0.f;
  const float q = __fdiv_rn(zero ? 1.f : inter, uni);
  const float z = uni > 0.f ? 0.f : (uni < 0.f ? -0.f : __int_as_float(0x7fffffff));
  return zero ? z : q;
}

constexpr int kIouRows = 16;  // rows per CTA: the four column boxes and their areas stay in registers across them

__global__ void __launch_bounds__(256) box_iou_kernel(const float4* __restrict__ a, int n_a, const float4* __restrict__ b, int n_b,
                                                       float* __restrict__ out, int vec_ok) {
  const int j0 = 4 * (blockIdx.x * blockDim.x + threadIdx.x);
  if (j0 >= n_b) return;
  float4 bb[4];
  float ab[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    bb[t] = b[j0 + t < n_b ? j0 + t : n_b - 1];
    ab[t] = box_area_tv(bb[t]);
  }
  const int i0 = blockIdx.y * kIouRows;
  const int i1 = i0 + kIouRows < n_a ? i0 + kIouRows : n_a;
  const bool full = vec_ok && j0 + 3 < n_b;
  float* row = out + (size_t)i0 * n_b + j0;
  for (int i = i0; i < i1; ++i, row += n_b) {
    const float4 ba = a[i];  // one address per warp: a broadcast load
    const float area_a = box_area_tv(ba);
    float v[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) v[t] = iou_tv(ba, area_a, bb[t], ab[t]);
    if (full) {
      __stcs(reinterpret_cast<float4*>(row), make_float4(v[0], v[1], v[2], v[3]));  // written once, not re-read here
    } else {
#pragma unroll
      for (int t = 0; t < 4; ++t)
        if (j0 + t < n_b) row[t] = v[t];
    }
  }
}

constexpr int kMatchThreads = 256;  // small CTAs: up to 8 images per SM, so one image's serial resolve overlaps the others' IoU work
constexpr int kMatchPanel = kMatchThreads / 32;  // rows per panel = warps per CTA

struct MatchParams {
  const float4* row_boxes;
  const long long* row_labels;
  const int* row_offsets;
  const float4* col_boxes;
  const long long* col_labels;
  const int* col_offsets;
  int words_cap;  // 32-bit words of one bit-row (host: ceil(max_cols / 32))
  float thr;
  unsigned char* row_matched;
  unsigned char* col_matched;
};

__global__ void __launch_bounds__(kMatchThreads) iou_match_kernel(const MatchParams p) {
  extern __shared__ uint32_t match_smem[];
  uint32_t* used = match_smem;                // [words_cap]
  uint32_t* mask = match_smem + p.words_cap;  // [kMatchPanel][words_cap]
  const int prob = blockIdx.x;
  const int r0 = p.row_offsets[prob], n_rows = p.row_offsets[prob + 1] - r0;
  const int c0 = p.col_offsets[prob];
  int n_cols = p.col_offsets[prob + 1] - c0;
  if (n_cols > p.words_cap * 32) n_cols = p.words_cap * 32;  // caller contract: max_cols bounds every problem
  const int words = (n_cols + 31) >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int w = threadIdx.x; w < words; w += kMatchThreads) used[w] = 0u;
  __syncthreads();
  uint32_t used_reg = 0u;
  for (int base = 0; base < n_rows; base += kMatchPanel) {
    const int r = base + warp;
    if (r < n_rows) {
      const float4 rb = p.row_boxes[r0 + r];
      const float area_r = box_area_tv(rb);
      const long long rl = p.row_labels ? p.row_labels[r0 + r] : 0;
      for (int w = 0; w < words; ++w) {
        const int j = (w << 5) + lane;
        bool claim = false;
        if (j < n_cols) {
          const float4 cb = p.col_boxes[c0 + j];
          const float iou = iou_tv(rb, area_r, cb, box_area_tv(cb));
          claim = !(iou < p.thr);
          if (p.row_labels && p.col_labels) claim = claim && p.col_labels[c0 + j] == rl;
        }
        const uint32_t bits = __ballot_sync(0xffffffffu, claim);
        if (lane == 0) mask[warp * p.words_cap + w] = bits;
      }
    }
    __syncthreads();
    if (warp == 0) {
      const int rows_here = min(kMatchPanel, n_rows - base);
      if (words <= 32) {
        // <= 1024 columns: lane w keeps word w of `used` in a register; the mask loads do not depend on it and pipeline
        uint32_t m[kMatchPanel];
#pragma unroll
        for (int q = 0; q < kMatchPanel; ++q) m[q] = (q < rows_here && lane < words) ? mask[q * p.words_cap + lane] : 0u;
#pragma unroll
        for (int q = 0; q < kMatchPanel; ++q) {
          const uint32_t claim = m[q] & ~used_reg;
          used_reg |= claim;
          const uint32_t hit = __ballot_sync(0xffffffffu, claim != 0u);
          if (lane == q && q < rows_here) p.row_matched[r0 + base + q] = hit != 0u;
        }
      } else {
        for (int q = 0; q < rows_here; ++q) {
          uint32_t any = 0u;
          for (int w = lane; w < words; w += 32) {
            const uint32_t claim = mask[q * p.words_cap + w] & ~used[w];
            used[w] |= claim;
            any |= claim;
          }
          const uint32_t hit = __ballot_sync(0xffffffffu, any != 0u);
          if (lane == 0) p.row_matched[r0 + base + q] = hit != 0u;
        }
      }
    }
    __syncthreads();
  }
  if (warp == 0 && words <= 32 && lane < words) used[lane] = used_reg;
  __syncthreads();
  for (int j = threadIdx.x; j < n_cols; j += kMatchThreads) p.col_matched[c0 + j] = (used[j >> 5] >> (j & 31)) & 1u;
}

}  // namespace cvpce

extern "C" int cvpce_box_iou(const float* boxes_a, int n_a, const float* boxes_b, int n_b, float* iou, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(n_a >= 0 && n_b >= 0, CVPCE_ERR_INVALID_ARG, "box_iou: negative size n_a=%d n_b=%d", n_a, n_b);
  if (n_a == 0 || n_b == 0) return CVPCE_OK;
  CVPCE_REQUIRE(boxes_a && boxes_b && iou, CVPCE_ERR_INVALID_ARG, "box_iou: null pointer");
  CVPCE_REQUIRE(aligned16(boxes_a) && aligned16(boxes_b), CVPCE_ERR_INVALID_ARG, "box_iou: boxes must be 16-byte aligned");
  CVPCE_REQUIRE(n_b <= (1 << 30), CVPCE_ERR_UNSUPPORTED, "box_iou: n_b=%d too large", n_b);
  if (int rc = require_sm100()) return rc;
  const int vec_ok = aligned16(iou) && n_b % 4 == 0;
  const int gx = (n_b + 4 * 256 - 1) / (4 * 256);
  // blockIdx.y is limited to 65535 row groups: walk taller matrices in slabs
  const int slab = 65535 * kIouRows;
  for (int i0 = 0; i0 < n_a; i0 += slab) {
    const int rows = n_a - i0 < slab ? n_a - i0 : slab;
    box_iou_kernel<<<dim3((unsigned)gx, (unsigned)((rows + kIouRows - 1) / kIouRows), 1), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4*>(boxes_a) + i0, rows, reinterpret_cast<const float4*>(boxes_b), n_b, iou + (size_t)i0 * n_b,
        vec_ok);
    CVPCE_LAUNCH_OK("box_iou_kernel");
  }
  return CVPCE_OK;
}

extern "C" int cvpce_iou_match(const float* row_boxes, const int64_t* row_labels, const int32_t* row_offsets,
                               const float* col_boxes, const int64_t* col_labels, const int32_t* col_offsets, int n_problems,
                               int max_cols, float iou_threshold, uint8_t* row_matched, uint8_t* col_matched, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(n_problems >= 0 && max_cols >= 0, CVPCE_ERR_INVALID_ARG, "iou_match: negative size");
  if (n_problems == 0) return CVPCE_OK;
  CVPCE_REQUIRE(row_offsets && col_offsets && row_matched && col_matched, CVPCE_ERR_INVALID_ARG, "iou_match: null pointer");
  CVPCE_REQUIRE(row_boxes && col_boxes, CVPCE_ERR_INVALID_ARG, "iou_match: null boxes");
  CVPCE_REQUIRE(aligned16(row_boxes) && aligned16(col_boxes), CVPCE_ERR_INVALID_ARG, "iou_match: boxes must be 16-byte aligned");
  CVPCE_REQUIRE((row_labels == nullptr) == (col_labels == nullptr), CVPCE_ERR_INVALID_ARG,
                "iou_match: give both label arrays or neither");
  CVPCE_REQUIRE(max_cols <= CVPCE_MATCH_MAX_COLS, CVPCE_ERR_UNSUPPORTED, "iou_match: max_cols=%d exceeds %d", max_cols,
                CVPCE_MATCH_MAX_COLS);
  if (int rc = require_sm100()) return rc;
  MatchParams p;
  p.row_boxes = reinterpret_cast<const float4*>(row_boxes);
  p.row_labels = reinterpret_cast<const long long*>(row_labels);
  p.row_offsets = row_offsets;
  p.col_boxes = reinterpret_cast<const float4*>(col_boxes);
  p.col_labels = reinterpret_cast<const long long*>(col_labels);
  p.col_offsets = col_offsets;
  p.words_cap = max_cols > 0 ? (max_cols + 31) / 32 : 1;
  p.thr = iou_threshold;
  p.row_matched = row_matched;
  p.col_matched = col_matched;
  const size_t smem = (size_t)p.words_cap * (1 + kMatchPanel) * sizeof(uint32_t);
  if (smem > 40 * 1024)
    if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(&iou_match_kernel), 96 * 1024)) return rc;
  iou_match_kernel<<<(unsigned)n_problems, kMatchThreads, smem, (cudaStream_t)stream>>>(p);
  CVPCE_LAUNCH_OK("iou_match_kernel");
  return CVPCE_OK;
}
