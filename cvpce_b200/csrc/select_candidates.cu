// Pre-NMS candidate selection of the RetinaNet/GLN detector head for sm_100a:
//   per FPN level: sigmoid -> score > score_thresh -> top-k -> box decode -> clip to the image -> concatenation over levels
// Follows tv:models/detection/retinanet.py:527-557 (postprocess_detections up to the batched_nms call),
// tv:models/detection/_utils.py:183-224 (BoxCoder.decode_single) and tv:ops/boxes.py clip_boxes_to_image
// (see include/cvpce_b200.h).  The output is the input of cvpce_nms_segments, with no host synchronisation in between.
//
// One thread-block CLUSTER per image, one CTA per FPN level (the cluster is padded to a power of two; padding CTAs
// only take part in the count exchange).  A CTA
//   1. streams its level's logits (coalesced, each pass reads them once) and radix-selects the top-k scores with a
//      shared-memory histogram: one histogram pass localises the k-th score to a bin of <= a few dozen candidates,
//      further passes only run when a bin is over-full (many equal scores);
//   2. a final pass appends the candidates above the bin (certain) and inside it (boundary) to a shared-memory list,
//      which a bitonic network orders by (score descending, flat index ascending);
//   3. the level counts are exchanged through distributed shared memory, so every CTA knows where its rows start in
//      the image's output segment and writes decoded, clipped boxes straight to their final position.
// Scores are computed with det_sigmoid below: IEEE fp32 +,*,fma and one division in a fixed order, so the CPU oracle
// (oracle/oracle.c: oracle_det_expf) reproduces them bit for bit; they differ from ATen's vectorised sigmoid by a few ulp.
#include "common.cuh"
#include <cooperative_groups.h>
#include <math.h>
#include <string.h>

namespace cg = cooperative_groups;

namespace cvpce {

constexpr int kSelThreads = 1024;
constexpr int kSelBinsLog2 = 11;
constexpr int kSelBins = 1 << kSelBinsLog2;
constexpr int kSelListCap = 4096;  // certain + boundary candidates that are sorted together

// exp(x) from IEEE-exact building blocks only (Cephes expf polynomial, two-step Cody-Waite reduction, exact 2^n scaling)
__device__ __forceinline__ float det_expf(float x) {
  if (x != x) return x;
  if (x > 88.7228394f) return __int_as_float(0x7f800000);
  if (x < -103.972084f) return 0.f;
  const float n = rintf(__fmul_rn(x, 1.44269502f));
  float r = __fmaf_rn(n, -0.693359375f, x);
  r = __fmaf_rn(n, 2.12194440e-4f, r);
  float p = 1.9875691500e-4f;
  p = __fmaf_rn(p, r, 1.3981999507e-3f);
  p = __fmaf_rn(p, r, 8.3334519073e-3f);
  p = __fmaf_rn(p, r, 4.1665795894e-2f);
  p = __fmaf_rn(p, r, 1.6666665459e-1f);
  p = __fmaf_rn(p, r, 5.0000001201e-1f);
  float y = __fmaf_rn(p, __fmul_rn(r, r), r);
  y = __fadd_rn(y, 1.f);
  const int ni = (int)n;
  const int n1 = ni / 2, n2 = ni - n1;  // |n1|,|n2| <= 76: both scale factors are normal numbers
  y = __fmul_rn(y, __int_as_float((127 + n1) << 23));
  return __fmul_rn(y, __int_as_float((127 + n2) << 23));
}

__device__ __forceinline__ float det_sigmoid(float x) { return __fdiv_rn(1.f, __fadd_rn(1.f, det_expf(-x))); }

struct SelLevel {
  const float* logits;
  const float* regression;
  const float* anchors;
  long long logits_stride, regression_stride, anchors_stride;
  int n_anchors;
};

struct SelParams {
  SelLevel lv[CVPCE_SELECT_MAX_LEVELS];
  int n_levels;
  int num_classes;
  int topk;
  float thresh;
  float prefilter;    // logits below this cannot pass the threshold (conservative), skips the sigmoid
  uint32_t key_base;  // key = bits(score) - key_base for a passing score
  int shift0;         // first histogram pass: bin = key >> shift0
  float wx, wy, ww, wh, xform_clip;
  const int32_t* image_sizes;
  int seg_stride;
  float4* boxes;
  float* scores;
  long long* labels;
  int32_t* seg_counts;
  int32_t* level_counts;
};

// score of a candidate with logit x, as a selection key; false if it does not pass the threshold
__device__ __forceinline__ bool sel_key(const float x, const SelParams& p, uint32_t& key) {
  if (!(x >= p.prefilter)) return false;  // also drops NaN logits (sigmoid(NaN) > thr is false in the reference)
  const float s = det_sigmoid(x);
  if (!(s > p.thresh)) return false;
  key = __float_as_uint(s) - p.key_base;
  return true;
}

// f(flat index, logit) for every candidate of the level whose logit passes the conservative pre-filter, with DENSE lanes:
// a trained detector leaves ~1 logit in 10 above the pre-filter, so evaluating the sigmoid where the logit is loaded
// would run ~70 instructions per element at ~10% lane utilisation.  Instead every warp compacts the survivors of its
// 128-bit loads (ballot + popc) into a private shared-memory queue and, whenever 32 are waiting, evaluates one per lane.
// Per-level views of the concatenated head output start at arbitrary multiples of 4 bytes: scalar head/tail, 2 x 16 B
// in flight per thread in between.
constexpr int kSelQueue = 32 + 8 * 32;  // a full batch left over + one iteration's survivors (2 loads x 4 x 32 lanes)

constexpr size_t kSelDynSmem = (size_t)(kSelThreads / 32) * kSelQueue * 8;  // 73,728 B

struct WarpQueue {
  float* x;  // [kSelQueue] of this warp
  int* i;
  int n;     // warp-uniform
};

__device__ __forceinline__ void queue_push(WarpQueue& q, const bool pass, const float x, const int i) {
  const unsigned lane = threadIdx.x & 31u;
  const unsigned m = __ballot_sync(0xffffffffu, pass);
  if (pass) {
    const int pos = q.n + __popc(m & ((1u << lane) - 1u));
    q.x[pos] = x;
    q.i[pos] = i;
  }
  q.n += __popc(m);
}

template <class F>
__device__ __forceinline__ void queue_drain(WarpQueue& q, const bool all, F& f) {
  const int lane = threadIdx.x & 31;
  __syncwarp();
  while (q.n >= 32 || (all && q.n > 0)) {
    const int take = min(q.n, 32);
    const int pos = q.n - take + lane;
    if (lane < take) f(q.i[pos], q.x[pos]);
    q.n -= take;
  }
  __syncwarp();
}

template <class F>
__device__ __forceinline__ void stream_level(const float* __restrict__ logits, const int n, const float prefilter,
                                             float* queue_x, int* queue_i, F f) {
  const int tid = threadIdx.x;
  WarpQueue q;
  q.x = queue_x + (tid >> 5) * kSelQueue;
  q.i = queue_i + (tid >> 5) * kSelQueue;
  q.n = 0;
  int head = (int)((4u - (unsigned)((reinterpret_cast<uintptr_t>(logits) >> 2) & 3u)) & 3u);
  head = min(head, n);
  const float4* v = reinterpret_cast<const float4*>(logits + head);
  const int nv = (n - head) >> 2;
  const int t0 = head + 4 * nv;
  if (tid < 32) {  // warp 0: the unaligned head and tail elements (at most 3 + 3)
    float x = 0.f;
    int i = -1;
    if (tid < head) i = tid;
    else if (tid - head < n - t0) i = t0 + tid - head;
    if (i >= 0) x = __ldg(logits + i);
    queue_push(q, i >= 0 && x >= prefilter, x, i);
  }
  for (int j0 = (tid >> 5) * 32; j0 < nv; j0 += 2 * kSelThreads) {  // warp-uniform trip count
    const int ja = j0 + (tid & 31), jb = ja + kSelThreads;
    const bool va = ja < nv, vb = jb < nv;
    float4 xa = make_float4(0.f, 0.f, 0.f, 0.f), xb = xa;
    if (va) xa = __ldg(v + ja);
    if (vb) xb = __ldg(v + jb);
    const int ia = head + 4 * ja, ib = head + 4 * jb;
    queue_push(q, va && xa.x >= prefilter, xa.x, ia);
    queue_push(q, va && xa.y >= prefilter, xa.y, ia + 1);
    queue_push(q, va && xa.z >= prefilter, xa.z, ia + 2);
    queue_push(q, va && xa.w >= prefilter, xa.w, ia + 3);
    queue_push(q, vb && xb.x >= prefilter, xb.x, ib);
    queue_push(q, vb && xb.y >= prefilter, xb.y, ib + 1);
    queue_push(q, vb && xb.z >= prefilter, xb.z, ib + 2);
    queue_push(q, vb && xb.w >= prefilter, xb.w, ib + 3);
    queue_drain(q, false, f);
  }
  queue_drain(q, true, f);
}

// exclusive prefix sum of v over the block (blockDim.x == kSelThreads); total returned through *total
__device__ __forceinline__ int block_excl_scan(int v, int* warp_tot /* [32] smem */, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const int w = warp_tot[lane];
    int winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    warp_tot[lane] = winc - w;  // exclusive warp offsets
    if (lane == 31) warp_tot[32] = winc;
  }
  __syncthreads();
  const int res = warp_tot[warp] + inc - v;
  *total = warp_tot[32];
  __syncthreads();  // warp_tot may be reused by the next call
  return res;
}

__global__ void __launch_bounds__(kSelThreads, 1) select_candidates_kernel(const __grid_constant__ SelParams p) {
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ unsigned long long list[kSelListCap];
  __shared__ int hist[kSelBins];
  extern __shared__ __align__(16) unsigned char sel_dyn_smem[];  // per-warp compaction queues
  float* queue_x = reinterpret_cast<float*>(sel_dyn_smem);
  int* queue_i = reinterpret_cast<int*>(queue_x + (kSelThreads / 32) * kSelQueue);
  __shared__ int warp_tot[33];
  __shared__ int s_bin, s_above, s_pop;
  __shared__ int s_n_certain, s_n_boundary;
  __shared__ int s_count;  // this level's selected candidates; read by the other CTAs of the cluster

  const int level = blockIdx.x;
  const int b = blockIdx.y;
  const int tid = threadIdx.x;
  int n_sel = 0;

  if (level < p.n_levels && p.lv[level].n_anchors > 0 && p.topk > 0) {
    const SelLevel& L = p.lv[level];
    const int n = L.n_anchors * p.num_classes;
    const float* logits = L.logits + (long long)b * L.logits_stride;

    // ---- 1. localise the k-th score ---------------------------------------------------------------------------
    uint32_t lo = 0;           // keys >= lo + width: certain; keys in [lo, lo+width): boundary
    uint32_t width = 0;        // 0: everything >= lo is certain (fewer passing candidates than topk)
    uint32_t range = ~0u;      // keys considered by the current histogram pass: key - lo < range
    int shift = p.shift0;
    int need = p.topk;         // candidates still to take from keys < lo + width
    int pop = 0;               // candidates inside the boundary bin
    for (int iter = 0;; ++iter) {
      for (int i = tid; i < kSelBins; i += kSelThreads) hist[i] = 0;
      __syncthreads();
      stream_level(logits, n, p.prefilter, queue_x, queue_i, [&](int, float x) {
        uint32_t key;
        if (sel_key(x, p, key)) {
          const uint32_t rel = key - lo;
          if (key >= lo && rel < range) atomicAdd(&hist[rel >> shift], 1);
        }
      });
      __syncthreads();
      // suffix scan from the top bin: thread t owns bins (top-2t, top-2t-1)
      const int top = kSelBins - 1;
      const int h0 = hist[top - 2 * tid], h1 = hist[top - 2 * tid - 1];
      int total;
      const int excl = block_excl_scan(h0 + h1, warp_tot, &total);
      if (iter == 0 && total <= need) {  // everything that passes the threshold is selected
        pop = 0;
        width = 0;
        need = total;
        break;
      }
      if (excl < need && need <= excl + h0 + h1) {
        if (need <= excl + h0) {
          s_bin = top - 2 * tid;
          s_above = excl;
          s_pop = h0;
        } else {
          s_bin = top - 2 * tid - 1;
          s_above = excl + h0;
          s_pop = h1;
        }
      }
      __syncthreads();
      need -= s_above;
      pop = s_pop;
      lo += (uint32_t)s_bin << shift;
      width = 1u << shift;
      __syncthreads();
      if ((p.topk - need) + pop <= kSelListCap || shift == 0) break;
      range = width;
      shift = shift > kSelBinsLog2 ? shift - kSelBinsLog2 : 0;
    }
    const int n_certain = (width == 0) ? need : p.topk - need;
    const bool ordered_ties = (width != 0) && (n_certain + pop > kSelListCap);  // only possible with width == 1

    // ---- 2. collect ---------------------------------------------------------------------------------------------
    if (tid == 0) {
      s_n_certain = 0;
      s_n_boundary = 0;
    }
    __syncthreads();
    if (!ordered_ties) {
      // only scores >= the boundary bin's lower edge are collected: pre-filter on the logit of that edge (conservative
      // margin >> the few-ulp error of det_sigmoid), so a dense head does not pay a second sigmoid per anchor
      float collect_filter = p.prefilter;
      if (width != 0) {
        const float s_lo = __uint_as_float(lo + p.key_base);
        if (s_lo > 0.f && s_lo < 1.f) {
          const float x_lo = logf(s_lo / (1.f - s_lo));
          collect_filter = fmaxf(collect_filter, x_lo - 1e-3f * fmaxf(1.f, fabsf(x_lo)));
        } else if (s_lo >= 1.f) {
          collect_filter = fmaxf(collect_filter, 8.f);  // sigmoid(x) rounds to 1 only for x > 16
        }
      }
      stream_level(logits, n, collect_filter, queue_x, queue_i, [&](int i, float x) {
        uint32_t key;
        if (sel_key(x, p, key) && key >= lo) {
          const unsigned long long e = ((unsigned long long)(~key) << 32) | (unsigned)i;
          if (width == 0 || key - lo >= width)
            list[atomicAdd(&s_n_certain, 1)] = e;
          else
            list[n_certain + atomicAdd(&s_n_boundary, 1)] = e;
        }
      });
    } else {
      // more equal boundary scores than the list holds (width == 1: they are all the same score): take the `need`
      // lowest flat indices.  Rare path: every chunk pays a block scan.
      int running = 0;
      for (int base = 0; base < n; base += kSelThreads) {
        const int i = base + tid;
        uint32_t key = 0;
        const bool ok = i < n && sel_key(__ldg(logits + i), p, key) && key >= lo;
        const bool tie = ok && key == lo;
        const unsigned long long e = ((unsigned long long)(~key) << 32) | (unsigned)i;
        if (ok && !tie) list[atomicAdd(&s_n_certain, 1)] = e;
        int total;
        const int rank = running + block_excl_scan(tie ? 1 : 0, warp_tot, &total);
        if (tie && rank < need) list[n_certain + rank] = e;
        running += total;
      }
      if (tid == 0) s_n_boundary = need;
    }
    __syncthreads();
    const int n_list = s_n_certain + s_n_boundary;  // s_n_certain == n_certain by construction

    // ---- 3. order by (score desc, flat index asc): ascending sort of (~key, index) ------------------------------------
    int cap = 2;
    while (cap < n_list) cap <<= 1;
    for (int i = n_list + tid; i < cap; i += kSelThreads) list[i] = ~0ull;
    __syncthreads();
    block_bitonic_sort_u64(list, cap);
    n_sel = min(n_list, p.topk);
  }

  // ---- 4. level offsets through distributed shared memory -----------------------------------------------------------
  if (tid == 0) s_count = n_sel;
  cluster.sync();
  int offset = 0, total_image = 0;
  {
    const unsigned csize = cluster.num_blocks();
    for (unsigned r = 0; r < csize; ++r) {
      const int c = *cluster.map_shared_rank(&s_count, r);
      if ((int)r < level) offset += c;
      total_image += c;
    }
  }
  if (tid == 0) {
    if (level == 0) p.seg_counts[b] = total_image;
    if (level < p.n_levels && p.level_counts) p.level_counts[b * p.n_levels + level] = n_sel;
  }

  // ---- 5. decode + clip + write (tv:_utils.py:183-224, tv:ops/boxes.py clip_boxes_to_image) -------------------------
  if (n_sel > 0) {
    const SelLevel& L = p.lv[level];
    const float4* reg = reinterpret_cast<const float4*>(L.regression + (long long)b * L.regression_stride);
    const float4* anc = reinterpret_cast<const float4*>(L.anchors + (long long)b * L.anchors_stride);
    const float img_h = (float)p.image_sizes[2 * b], img_w = (float)p.image_sizes[2 * b + 1];
    const size_t out0 = (size_t)b * p.seg_stride + offset;
    for (int r = tid; r < n_sel; r += kSelThreads) {
      const unsigned long long e = list[r];
      const uint32_t key = ~(uint32_t)(e >> 32);
      const int flat = (int)(e & 0xffffffffu);
      const int a = flat / p.num_classes;
      const float4 rc = __ldg(reg + a);
      const float4 an = __ldg(anc + a);
      const float w = __fsub_rn(an.z, an.x), h = __fsub_rn(an.w, an.y);
      const float cx = __fadd_rn(an.x, __fmul_rn(0.5f, w)), cy = __fadd_rn(an.y, __fmul_rn(0.5f, h));
      const float dx = __fdiv_rn(rc.x, p.wx), dy = __fdiv_rn(rc.y, p.wy);
      float dw = __fdiv_rn(rc.z, p.ww), dh = __fdiv_rn(rc.w, p.wh);
      dw = dw > p.xform_clip ? p.xform_clip : dw;  // torch.clamp(max=): NaN propagates
      dh = dh > p.xform_clip ? p.xform_clip : dh;
      const float pcx = __fadd_rn(__fmul_rn(dx, w), cx), pcy = __fadd_rn(__fmul_rn(dy, h), cy);
      const float hw = __fmul_rn(0.5f, __fmul_rn(det_expf(dw), w)), hh = __fmul_rn(0.5f, __fmul_rn(det_expf(dh), h));
      float x1 = __fsub_rn(pcx, hw), y1 = __fsub_rn(pcy, hh), x2 = __fadd_rn(pcx, hw), y2 = __fadd_rn(pcy, hh);
      x1 = x1 < 0.f ? 0.f : (x1 > img_w ? img_w : x1);
      x2 = x2 < 0.f ? 0.f : (x2 > img_w ? img_w : x2);
      y1 = y1 < 0.f ? 0.f : (y1 > img_h ? img_h : y1);
      y2 = y2 < 0.f ? 0.f : (y2 > img_h ? img_h : y2);
      p.boxes[out0 + r] = make_float4(x1, y1, x2, y2);
      p.scores[out0 + r] = __uint_as_float(key + p.key_base);
      if (p.labels) p.labels[out0 + r] = (long long)(flat - a * p.num_classes);
    }
  }
  cluster.sync();  // no CTA may exit while its s_count can still be read
}

}  // namespace cvpce

extern "C" int cvpce_select_candidates(const cvpce_fpn_level* levels, int n_levels, int n_images, int num_classes,
                                       float score_thresh, int topk, const int32_t* image_sizes, const float* weights,
                                       float bbox_xform_clip, float* boxes, float* scores, int64_t* labels,
                                       int32_t* seg_counts, int32_t* level_counts, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(n_images >= 0 && n_levels >= 0 && topk >= 0, CVPCE_ERR_INVALID_ARG,
                "select_candidates: negative size (n_images=%d n_levels=%d topk=%d)", n_images, n_levels, topk);
  if (n_images == 0) return CVPCE_OK;
  CVPCE_REQUIRE(n_levels >= 1 && n_levels <= CVPCE_SELECT_MAX_LEVELS, CVPCE_ERR_UNSUPPORTED,
                "select_candidates: n_levels=%d outside [1,%d]", n_levels, CVPCE_SELECT_MAX_LEVELS);
  CVPCE_REQUIRE(topk <= CVPCE_SELECT_MAX_TOPK, CVPCE_ERR_UNSUPPORTED, "select_candidates: topk=%d exceeds %d", topk,
                CVPCE_SELECT_MAX_TOPK);
  CVPCE_REQUIRE(num_classes >= 1, CVPCE_ERR_INVALID_ARG, "select_candidates: num_classes=%d", num_classes);
  CVPCE_REQUIRE(levels && image_sizes && boxes && scores && seg_counts, CVPCE_ERR_INVALID_ARG,
                "select_candidates: levels / image_sizes / boxes / scores / seg_counts is null");
  CVPCE_REQUIRE(aligned16(boxes), CVPCE_ERR_INVALID_ARG, "select_candidates: boxes must be 16-byte aligned");
  if (int rc = require_sm100()) return rc;
  SelParams p;
  for (int l = 0; l < CVPCE_SELECT_MAX_LEVELS; ++l) p.lv[l] = SelLevel{nullptr, nullptr, nullptr, 0, 0, 0, 0};
  for (int l = 0; l < n_levels; ++l) {
    const cvpce_fpn_level& s = levels[l];
    CVPCE_REQUIRE(s.n_anchors >= 0, CVPCE_ERR_INVALID_ARG, "select_candidates: level %d has n_anchors=%d", l, s.n_anchors);
    CVPCE_REQUIRE((long long)s.n_anchors * num_classes < (1ll << 31), CVPCE_ERR_UNSUPPORTED,
                  "select_candidates: level %d has more than 2^31 candidates", l);
    if (s.n_anchors > 0) {
      CVPCE_REQUIRE(s.logits && s.regression && s.anchors, CVPCE_ERR_INVALID_ARG, "select_candidates: level %d has a null pointer", l);
      CVPCE_REQUIRE(aligned16(s.regression) && aligned16(s.anchors) && s.regression_stride % 4 == 0 && s.anchors_stride % 4 == 0,
                    CVPCE_ERR_INVALID_ARG, "select_candidates: level %d regression / anchors rows must be 16-byte aligned", l);
    }
    p.lv[l] = SelLevel{s.logits, s.regression, s.anchors, (long long)s.logits_stride, (long long)s.regression_stride,
                       (long long)s.anchors_stride, s.n_anchors};
  }
  p.n_levels = n_levels;
  p.num_classes = num_classes;
  p.topk = topk;
  p.thresh = score_thresh;
  // scores are in [0,1]: keys are the distance in fp32 bit patterns above the threshold
  uint32_t tb;
  memcpy(&tb, &score_thresh, 4);
  p.key_base = (score_thresh >= 0.f) ? tb + 1u : 0u;   // thr = +/-0 or negative: every score >= +0 ... see sel_key
  if (score_thresh == 0.f) p.key_base = 1u;            // -0.0 as threshold: bits would be 0x80000000
  if (!(score_thresh == score_thresh)) p.key_base = 0u;  // NaN threshold: nothing passes (s > NaN is false)
  const uint32_t one_bits = 0x3f800000u;
  const uint32_t range = (p.key_base <= one_bits) ? one_bits - p.key_base + 1u : 1u;
  int bits = 0;
  while (bits < 32 && (range - 1u) >> bits) ++bits;
  p.shift0 = bits > kSelBinsLog2 ? bits - kSelBinsLog2 : 0;
  // conservative logit pre-filter: sigmoid(x) <= thr certainly when x < logit(thr) - margin
  if (score_thresh > 0.f && score_thresh < 1.f) {
    const double t = log((double)score_thresh / (1.0 - (double)score_thresh));
    p.prefilter = (float)(t - 1e-3 * (fabs(t) > 1.0 ? fabs(t) : 1.0));
  } else {
    p.prefilter = -INFINITY;
  }
  p.wx = weights ? weights[0] : 1.f;
  p.wy = weights ? weights[1] : 1.f;
  p.ww = weights ? weights[2] : 1.f;
  p.wh = weights ? weights[3] : 1.f;
  p.xform_clip = bbox_xform_clip;
  p.image_sizes = image_sizes;
  p.seg_stride = n_levels * topk;
  p.boxes = reinterpret_cast<float4*>(boxes);
  p.scores = scores;
  p.labels = reinterpret_cast<long long*>(labels);
  p.seg_counts = seg_counts;
  p.level_counts = level_counts;

  if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(&select_candidates_kernel), kSelDynSmem)) return rc;
  // one CTA per level; if this device refuses a cluster of exactly n_levels CTAs, pad to the next power of two (the
  // padding CTAs only take part in the count exchange)
  // cached per (thread, device): 0 unknown, 1 yes, -1 no
  static thread_local int exact_dev = -1;
  static thread_local int exact_ok[CVPCE_SELECT_MAX_LEVELS + 1] = {0};
  int cur_dev = 0;
  CVPCE_CUDA_OK(cudaGetDevice(&cur_dev));
  if (cur_dev != exact_dev) {
    for (int i = 0; i <= CVPCE_SELECT_MAX_LEVELS; ++i) exact_ok[i] = 0;
    exact_dev = cur_dev;
  }
  int csize = n_levels;
  if (exact_ok[n_levels] == 0) {
    cudaLaunchConfig_t q = {};
    q.gridDim = dim3(n_levels, 1, 1);
    q.blockDim = dim3(kSelThreads, 1, 1);
    q.dynamicSmemBytes = kSelDynSmem;
    cudaLaunchAttribute qa[1];
    qa[0].id = cudaLaunchAttributeClusterDimension;
    qa[0].val.clusterDim.x = n_levels;
    qa[0].val.clusterDim.y = 1;
    qa[0].val.clusterDim.z = 1;
    q.attrs = qa;
    q.numAttrs = 1;
    int n_clusters = 0;
    const cudaError_t e = cudaOccupancyMaxActiveClusters(&n_clusters, select_candidates_kernel, &q);
    if (e != cudaSuccess) (void)cudaGetLastError();
    exact_ok[n_levels] = (e == cudaSuccess && n_clusters > 0) ? 1 : -1;
  }
  if (exact_ok[n_levels] < 0) {
    csize = 1;
    while (csize < n_levels) csize <<= 1;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(csize, n_images, 1);
  cfg.blockDim = dim3(kSelThreads, 1, 1);
  cfg.dynamicSmemBytes = kSelDynSmem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  CVPCE_CUDA_OK(cudaLaunchKernelEx(&cfg, select_candidates_kernel, p));
  CVPCE_LAUNCH_OK("select_candidates_kernel");
  return CVPCE_OK;
}
