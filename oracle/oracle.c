/*
 * oracle.c -- CPU restatement of the reference's proposal -> product classification path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this; the product
 * (cvpce_b200/) never does.  Plain C, fp32 arithmetic where the reference computes in fp32, compiled WITHOUT
 * -ffast-math and with -ffp-contract=off so that the NMS decisions are the x86 IEEE ones.
 *
 * Each function cites the reference lines it follows (paths relative to the laitalaj/cvpce root; "tv:" = torchvision
 * 0.26, "aten:" = PyTorch 2.11 ATen -- third-party dependencies whose compiled ops the reference calls; their source
 * is not in the reference tree, so their published algorithms are restated here and PINNED against the compiled
 * ops themselves by oracle/make_golden.py -> tests/golden/ and tests/test_oracle_golden.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define API __attribute__((visibility("default")))

/* ---------------------------------------------------------------------------------------------------------------
 * NMS.  tv:ops/boxes.py:20-48 -> torch.ops.torchvision.nms (CPU kernel nms_kernel_impl): stable descending sort of
 * the scores, areas = (x2-x1)*(y2-y1), greedy loop; j suppressed iff ovr = inter/(iarea + area_j - inter) > thr,
 * with ovr in fp32 and thr a double.  Reached from tv:models/detection/retinanet.py:559-561 (batched_nms + [:N]).
 * Returns the number of kept indices written to keep (score-descending order).
 * --------------------------------------------------------------------------------------------------------------- */
typedef struct { float s; int32_t i; } sortkey;

static int cmp_desc_stable(const void* a, const void* b) {
  const sortkey* x = (const sortkey*)a;
  const sortkey* y = (const sortkey*)b;
  const int xn = x->s != x->s, yn = y->s != y->s; /* NaN sorts first in torch's descending sort */
  if (xn != yn) return xn ? -1 : 1;
  if (!xn) {
    if (x->s > y->s) return -1;
    if (x->s < y->s) return 1;
  }
  return (x->i > y->i) - (x->i < y->i);
}

API int64_t oracle_nms(const float* boxes, const float* scores, int64_t n, double iou_threshold, int64_t* keep) {
  if (n <= 0) return 0;
  sortkey* order = (sortkey*)malloc(sizeof(sortkey) * (size_t)n);
  float* areas = (float*)malloc(sizeof(float) * (size_t)n);
  unsigned char* suppressed = (unsigned char*)calloc((size_t)n, 1);
  for (int64_t i = 0; i < n; ++i) {
    order[i].s = scores[i];
    order[i].i = (int32_t)i;
    areas[i] = (boxes[4 * i + 2] - boxes[4 * i + 0]) * (boxes[4 * i + 3] - boxes[4 * i + 1]);
  }
  qsort(order, (size_t)n, sizeof(sortkey), cmp_desc_stable);
  int64_t nk = 0;
  for (int64_t _i = 0; _i < n; ++_i) {
    const int64_t i = order[_i].i;
    if (suppressed[i]) continue;
    keep[nk++] = i;
    const float ix1 = boxes[4 * i], iy1 = boxes[4 * i + 1], ix2 = boxes[4 * i + 2], iy2 = boxes[4 * i + 3];
    const float iarea = areas[i];
    for (int64_t _j = _i + 1; _j < n; ++_j) {
      const int64_t j = order[_j].i;
      if (suppressed[j]) continue;
      const float xx1 = ix1 > boxes[4 * j] ? ix1 : boxes[4 * j];
      const float yy1 = iy1 > boxes[4 * j + 1] ? iy1 : boxes[4 * j + 1];
      const float xx2 = ix2 < boxes[4 * j + 2] ? ix2 : boxes[4 * j + 2];
      const float yy2 = iy2 < boxes[4 * j + 3] ? iy2 : boxes[4 * j + 3];
      const float dw = xx2 - xx1, dh = yy2 - yy1;
      const float w = dw > 0.f ? dw : 0.f;
      const float h = dh > 0.f ? dh : 0.f;
      const float inter = w * h;
      const float uni = iarea + areas[j] - inter;
      const float ovr = inter / uni;
      if ((double)ovr > iou_threshold) suppressed[j] = 1;
    }
  }
  free(order);
  free(areas);
  free(suppressed);
  return nk;
}

/* ---------------------------------------------------------------------------------------------------------------
 * Crop + pad + resize.  cvpce/production.py:20 (boxes.to(long); python slice image[:, y1:y2, x1:x2]) and
 * cvpce/datautils.py:234-239 resize_for_classification: canvas [3,L,L] = 0.5, canvas[:, :h, :w] = crop,
 * ttf.resize(canvas, (S,S)) -> F.interpolate(bilinear, align_corners=False, antialias) (tv:transforms/_functional_tensor.py:445-467).
 * antialias=1: aten:native/cpu/UpSampleKernel.cpp separable_upsample_generic_Nd_kernel_impl with
 *   HelperInterpBase::_compute_indices_min_size_weights_aa (bilinear filter): horizontal pass, then vertical pass.
 * antialias=0: aten upsample_bilinear2d (area_pixel_compute_source_index, two taps per axis).
 * Returns 0, or 1 when the crop is empty in both dimensions (the reference raises there); out is then filled with pad.
 * --------------------------------------------------------------------------------------------------------------- */
static int64_t slice_index(int64_t i, int64_t n) {
  if (i < 0) { i += n; if (i < 0) i = 0; }
  return i > n ? n : i;
}

static float aa_filter(float x) {
  if (x < 0.f) x = -x;
  return x < 1.f ? 1.f - x : 0.f;
}

/* weights for one output index; returns xsize, sets *xmin */
static int aa_weights(int64_t i, int64_t input_size, float scale, float support, float* w, int max_size, int64_t* xmin_out) {
  const float center = (float)((double)scale * ((double)i + 0.5));
  const float invscale = (scale >= 1.0f) ? 1.0f / scale : 1.0f;
  int64_t xmin = (int64_t)((double)(center - support) + 0.5);
  if (xmin < 0) xmin = 0;
  int64_t hi = (int64_t)((double)(center + support) + 0.5);
  if (hi > input_size) hi = input_size;
  int64_t xsize = hi - xmin;
  if (xsize < 0) xsize = 0;
  if (xsize > max_size) xsize = max_size;
  float total = 0.f;
  for (int64_t j = 0; j < xsize; ++j) {
    const float x = (float)(((double)((float)(j + xmin) - center) + 0.5) * (double)invscale);
    w[j] = aa_filter(x);
    total += w[j];
  }
  const float norm = total != 0.f ? 1.0f / total : 0.f;
  for (int64_t j = 0; j < xsize; ++j) w[j] *= norm;
  *xmin_out = xmin;
  return (int)xsize;
}

API int oracle_crop_resize(const float* image, int64_t H, int64_t W, int64_t stride_c, int64_t stride_h, const float* box,
                           int S, int antialias, float pad, float* out /* [3,S,S] */) {
  int64_t b[4];
  for (int t = 0; t < 4; ++t) b[t] = (int64_t)box[t]; /* Tensor.to(torch.long) truncates toward zero */
  const int64_t xs = slice_index(b[0], W), xe = slice_index(b[2], W);
  const int64_t ys = slice_index(b[1], H), ye = slice_index(b[3], H);
  const int64_t w = xe > xs ? xe - xs : 0, h = ye > ys ? ye - ys : 0;
  const int64_t L = h > w ? h : w;
  if (L == 0) {
    for (int64_t i = 0; i < 3 * (int64_t)S * S; ++i) out[i] = pad;
    return 1;
  }
  float* canvas = (float*)malloc(sizeof(float) * 3 * (size_t)L * (size_t)L);
  for (int64_t i = 0; i < 3 * L * L; ++i) canvas[i] = pad;
  for (int c = 0; c < 3; ++c)
    for (int64_t y = 0; y < h; ++y)
      memcpy(canvas + (c * L + y) * L, image + c * stride_c + (ys + y) * stride_h + xs, sizeof(float) * (size_t)w);
  const float scale = (float)L / (float)S;
  if (antialias) {
    const float support = scale >= 1.0f ? scale : 1.0f;
    const int max_size = (int)ceilf(support) * 2 + 1;
    float* wts = (float*)malloc(sizeof(float) * (size_t)S * (size_t)max_size);
    int64_t* mins = (int64_t*)malloc(sizeof(int64_t) * (size_t)S);
    int* sizes = (int*)malloc(sizeof(int) * (size_t)S);
    for (int o = 0; o < S; ++o) sizes[o] = aa_weights(o, L, scale, support, wts + (size_t)o * max_size, max_size, &mins[o]);
    float* tmp = (float*)malloc(sizeof(float) * 3 * (size_t)L * (size_t)S); /* [3,L,S] after the horizontal pass */
    for (int64_t cy = 0; cy < 3 * L; ++cy)
      for (int o = 0; o < S; ++o) {
        const float* src = canvas + cy * L + mins[o];
        const float* wv = wts + (size_t)o * max_size;
        float t = 0.f;
        for (int j = 0; j < sizes[o]; ++j) t += src[j] * wv[j];
        tmp[cy * S + o] = t;
      }
    for (int c = 0; c < 3; ++c)
      for (int oy = 0; oy < S; ++oy) {
        const float* wv = wts + (size_t)oy * max_size;
        for (int ox = 0; ox < S; ++ox) {
          float t = 0.f;
          for (int i = 0; i < sizes[oy]; ++i) t += tmp[((int64_t)c * L + mins[oy] + i) * S + ox] * wv[i];
          out[((int64_t)c * S + oy) * S + ox] = t;
        }
      }
    free(tmp); free(sizes); free(mins); free(wts);
  } else {
    for (int oy = 0; oy < S; ++oy) {
      float sy = scale * ((float)oy + 0.5f) - 0.5f;
      if (sy < 0.f) sy = 0.f;
      int64_t y0 = (int64_t)sy;
      if (y0 > L - 1) y0 = L - 1;
      const int64_t y1 = y0 + (y0 < L - 1 ? 1 : 0);
      const float ly1 = sy - (float)y0, ly0 = 1.f - ly1;
      for (int ox = 0; ox < S; ++ox) {
        float sx = scale * ((float)ox + 0.5f) - 0.5f;
        if (sx < 0.f) sx = 0.f;
        int64_t x0 = (int64_t)sx;
        if (x0 > L - 1) x0 = L - 1;
        const int64_t x1 = x0 + (x0 < L - 1 ? 1 : 0);
        const float lx1 = sx - (float)x0, lx0 = 1.f - lx1;
        for (int c = 0; c < 3; ++c) {
          const float* p = canvas + (int64_t)c * L * L;
          out[((int64_t)c * S + oy) * S + ox] =
              ly0 * (lx0 * p[y0 * L + x0] + lx1 * p[y0 * L + x1]) + ly1 * (lx0 * p[y1 * L + x0] + lx1 * p[y1 * L + x1]);
        }
      }
    }
  }
  free(canvas);
  return 0;
}

/* ---------------------------------------------------------------------------------------------------------------
 * Region max-pool embedding.  cvpce/models/classification.py:46-51 (MACVGG: x.amax(dim=(-2,-1)) of two maps, cat,
 * desc / norm(desc).clamp(min=eps)) and :79-85 (MACResNet).  With regions + per_region_l2 it is R-MAC
 * (Tolias, Sicre, Jegou 2016: per region channel-wise max -> L2 -> sum -> L2), the north-star extension
 * ("parity unpinned" by the reference: it has no R-MAC).
 * feat: NCHW-contiguous fp32 maps.  regions: (y0,x0,y1,x1) per map, concatenated; n_regions NULL = whole map.
 * --------------------------------------------------------------------------------------------------------------- */
API void oracle_region_pool_embed(const float* const* feats, const int32_t* Cs, const int32_t* Hs, const int32_t* Ws, int n_maps,
                                  int64_t B, const int32_t* regions, const int32_t* n_regions, int per_region_l2, int final_l2,
                                  float eps, float* out /* [B, sum C] */) {
  int64_t D = 0;
  for (int m = 0; m < n_maps; ++m) D += Cs[m];
  for (int64_t n = 0; n < B; ++n) {
    float* orow = out + n * D;
    int64_t col = 0;
    const int32_t* rg = regions;
    for (int m = 0; m < n_maps; ++m) {
      const int C = Cs[m], H = Hs[m], W = Ws[m];
      const int R = n_regions ? n_regions[m] : 1;
      float* v = (float*)malloc(sizeof(float) * (size_t)C);
      for (int c = 0; c < C; ++c) orow[col + c] = 0.f;
      for (int r = 0; r < R; ++r) {
        const int y0 = n_regions ? rg[4 * r] : 0, x0 = n_regions ? rg[4 * r + 1] : 0;
        const int y1 = n_regions ? rg[4 * r + 2] : H, x1 = n_regions ? rg[4 * r + 3] : W;
        double ss = 0.0;
        for (int c = 0; c < C; ++c) {
          const float* plane = feats[m] + ((n * C + c) * (int64_t)H) * W;
          float mx = -INFINITY;
          for (int y = y0; y < y1; ++y)
            for (int x = x0; x < x1; ++x)
              if (plane[y * W + x] > mx) mx = plane[y * W + x];
          v[c] = mx;
          ss += (double)mx * (double)mx;
        }
        float dn = 1.f;
        if (per_region_l2) {
          dn = (float)sqrt(ss);
          if (dn < eps) dn = eps;
        }
        for (int c = 0; c < C; ++c) {
          const float t = per_region_l2 ? v[c] / dn : v[c];
          orow[col + c] = (r == 0) ? t : orow[col + c] + t;
        }
      }
      free(v);
      if (n_regions) rg += 4 * R;
      col += C;
    }
    if (final_l2) {
      double ss = 0.0;
      for (int64_t c = 0; c < D; ++c) ss += (double)orow[c] * (double)orow[c];
      float dn = (float)sqrt(ss);
      if (dn < eps) dn = eps;
      for (int64_t c = 0; c < D; ++c) orow[c] = orow[c] / dn;
    }
  }
}

/* R-MAC region grid (Tolias et al. 2016): scale l = 1..levels, region side floor(2*min(H,W)/(l+1)), regions spread
 * uniformly along each axis.  Mirrors the host helper the library exports so both can be compared. */
API int oracle_rmac_regions(int H, int W, int levels, int32_t* regions, int max_regions) {
  int count = 0;
  const int w = H < W ? H : W;
  for (int l = 1; l <= levels; ++l) {
    int side = (2 * w) / (l + 1);
    if (side < 1) side = 1;
    const int ny = (H == w) ? l : l + (int)((double)(H - w) / (double)w * (l + 1) / 2.0 + 0.5);
    const int nx = (W == w) ? l : l + (int)((double)(W - w) / (double)w * (l + 1) / 2.0 + 0.5);
    for (int iy = 0; iy < ny; ++iy) {
      const int y0 = ny > 1 ? (int)((double)(H - side) * iy / (ny - 1)) : (H - side) / 2;
      for (int ix = 0; ix < nx; ++ix) {
        const int x0 = nx > 1 ? (int)((double)(W - side) * ix / (nx - 1)) : (W - side) / 2;
        if (count < max_regions && regions) {
          regions[4 * count] = y0; regions[4 * count + 1] = x0;
          regions[4 * count + 2] = y0 + side; regions[4 * count + 3] = x0 + side;
        }
        ++count;
      }
    }
  }
  return count;
}

/* ---------------------------------------------------------------------------------------------------------------
 * Cosine similarity + top-k.  cvpce/models/classification.py:87-95: distance = 1 - cosine_similarity(a, b),
 * nearest_neighbors = distances.argsort(-1)[:, :k].  cosine_similarity (aten:native/Distance.cpp) =
 * sum(a*b) / (max(||a||,eps) * max(||b||,eps)), eps = 1e-8.  Dot products are accumulated in DOUBLE here so the
 * ranking is the mathematically exact one; the reference's argsort is not stable, so ties are resolved
 * (similarity descending, gallery index ascending) -- the contract the CUDA path states.  normalize=0 skips the norms
 * (rows already unit length, as MACVGG outputs are).  out_sim holds cosine similarities (= 1 - distance).
 * --------------------------------------------------------------------------------------------------------------- */
API void oracle_sim_topk(const float* q, const float* g, int64_t Q, int64_t G, int64_t D, int k, int normalize, double eps,
                         double* out_sim /* [Q,k] */, int64_t* out_idx /* [Q,k] */) {
  double* gn = (double*)malloc(sizeof(double) * (size_t)(G > 0 ? G : 1));
  for (int64_t j = 0; j < G; ++j) {
    double s = 0.0;
    if (normalize) {
      for (int64_t d = 0; d < D; ++d) s += (double)g[j * D + d] * (double)g[j * D + d];
      s = sqrt(s);
      if (s < eps) s = eps;
    } else {
      s = 1.0;
    }
    gn[j] = s;
  }
  for (int64_t i = 0; i < Q; ++i) {
    double* ts = out_sim + i * k;
    int64_t* ti = out_idx + i * k;
    for (int t = 0; t < k; ++t) { ts[t] = -INFINITY; ti[t] = -1; }
    double qn = 1.0;
    if (normalize) {
      double s = 0.0;
      for (int64_t d = 0; d < D; ++d) s += (double)q[i * D + d] * (double)q[i * D + d];
      qn = sqrt(s);
      if (qn < eps) qn = eps;
    }
    const float* qi = q + i * D;
    for (int64_t j = 0; j < G; ++j) {
      const float* gj = g + j * D;
      double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
      int64_t d = 0;
      for (; d + 3 < D; d += 4) {
        a0 += (double)qi[d] * (double)gj[d];
        a1 += (double)qi[d + 1] * (double)gj[d + 1];
        a2 += (double)qi[d + 2] * (double)gj[d + 2];
        a3 += (double)qi[d + 3] * (double)gj[d + 3];
      }
      for (; d < D; ++d) a0 += (double)qi[d] * (double)gj[d];
      const double s = ((a0 + a1) + (a2 + a3)) / (qn * gn[j]);
      if (ti[k - 1] < 0 || s > ts[k - 1]) {
        int pos = k - 1;
        while (pos > 0 && (ti[pos - 1] < 0 || ts[pos - 1] < s)) { ts[pos] = ts[pos - 1]; ti[pos] = ti[pos - 1]; --pos; }
        ts[pos] = s;
        ti[pos] = j;
      }
    }
  }
  free(gn);
}

/* fp32 variant used only as a timed CPU baseline (reference arithmetic type): Q x G^T in fp32 + top-k. */
API void oracle_sim_topk_f32(const float* q, const float* g, int64_t Q, int64_t G, int64_t D, int k, float* out_sim,
                             int64_t* out_idx) {
  for (int64_t i = 0; i < Q; ++i) {
    float* ts = out_sim + i * k;
    int64_t* ti = out_idx + i * k;
    for (int t = 0; t < k; ++t) { ts[t] = -INFINITY; ti[t] = -1; }
    const float* qi = q + i * D;
    for (int64_t j = 0; j < G; ++j) {
      const float* gj = g + j * D;
      float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      int64_t d = 0;
      for (; d + 7 < D; d += 8)
        for (int u = 0; u < 8; ++u) acc[u] += qi[d + u] * gj[d + u];
      for (; d < D; ++d) acc[0] += qi[d] * gj[d];
      const float s = ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((acc[4] + acc[5]) + (acc[6] + acc[7]));
      if (ti[k - 1] < 0 || s > ts[k - 1]) {
        int pos = k - 1;
        while (pos > 0 && (ti[pos - 1] < 0 || ts[pos - 1] < s)) { ts[pos] = ts[pos - 1]; ti[pos] = ti[pos - 1]; --pos; }
        ts[pos] = s;
        ti[pos] = j;
      }
    }
  }
}

/* ---------------------------------------------------------------------------------------------------------------
 * Pre-NMS candidate selection.  tv:models/detection/retinanet.py:527-557, one (image, level):
 *   scores = sigmoid(logits).flatten(); keep = scores > score_thresh; topk(min(topk_candidates, #keep));
 *   anchor = idx // num_classes, label = idx % num_classes; BoxCoder.decode_single (tv:models/detection/_utils.py:183-224);
 *   clip_boxes_to_image (tv:ops/boxes.py).
 * exp / sigmoid: ATen's vectorised kernels (Sleef) are third-party compiled code; their results are reproduced to a few
 * ulp by oracle_det_expf -- a fixed sequence of IEEE fp32 operations (Cephes expf polynomial) that the CUDA kernel
 * repeats bit for bit.  Equal scores are ordered by ascending flat index (torch.topk: unspecified).
 * Returns the number of candidates written (score-descending).
 * --------------------------------------------------------------------------------------------------------------- */
static float bits_to_float(uint32_t u) {
  float f;
  memcpy(&f, &u, 4);
  return f;
}

API float oracle_det_expf(float x) {
  if (x != x) return x;
  if (x > 88.7228394f) return INFINITY;
  if (x < -103.972084f) return 0.f;
  const float n = rintf(x * 1.44269502f);
  float r = fmaf(n, -0.693359375f, x);
  r = fmaf(n, 2.12194440e-4f, r);
  float p = 1.9875691500e-4f;
  p = fmaf(p, r, 1.3981999507e-3f);
  p = fmaf(p, r, 8.3334519073e-3f);
  p = fmaf(p, r, 4.1665795894e-2f);
  p = fmaf(p, r, 1.6666665459e-1f);
  p = fmaf(p, r, 5.0000001201e-1f);
  float y = fmaf(p, r * r, r);
  y = y + 1.f;
  const int ni = (int)n;
  const int n1 = ni / 2, n2 = ni - n1;
  y = y * bits_to_float((uint32_t)(127 + n1) << 23);
  return y * bits_to_float((uint32_t)(127 + n2) << 23);
}

API float oracle_det_sigmoid(float x) { return 1.f / (1.f + oracle_det_expf(-x)); }

static float clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); } /* NaN propagates */

API int64_t oracle_select_level(const float* logits, const float* regression, const float* anchors, int64_t n_anchors,
                                int64_t num_classes, float score_thresh, int64_t topk, int64_t img_h, int64_t img_w,
                                const float* weights, float xform_clip, float* out_boxes, float* out_scores,
                                int64_t* out_labels) {
  const int64_t n = n_anchors * num_classes;
  if (n <= 0 || topk <= 0) return 0;
  sortkey* cand = (sortkey*)malloc(sizeof(sortkey) * (size_t)n);
  int64_t m = 0;
  for (int64_t i = 0; i < n; ++i) {
    const float s = oracle_det_sigmoid(logits[i]);
    if (s > score_thresh) { /* retinanet.py:533-536 */
      cand[m].s = s;
      cand[m].i = (int32_t)i;
      ++m;
    }
  }
  qsort(cand, (size_t)m, sizeof(sortkey), cmp_desc_stable);
  const int64_t k = m < topk ? m : topk; /* det_utils._topk_min, retinanet.py:539-541 */
  const float wx = weights ? weights[0] : 1.f, wy = weights ? weights[1] : 1.f;
  const float ww = weights ? weights[2] : 1.f, wh = weights ? weights[3] : 1.f;
  for (int64_t r = 0; r < k; ++r) {
    const int64_t a = cand[r].i / num_classes; /* retinanet.py:543-544 */
    const float* rc = regression + 4 * a;
    const float* an = anchors + 4 * a;
    const float w = an[2] - an[0], h = an[3] - an[1];       /* _utils.py:195-198 */
    const float hw0 = 0.5f * w, hh0 = 0.5f * h;
    const float cx = an[0] + hw0, cy = an[1] + hh0;
    const float dx = rc[0] / wx, dy = rc[1] / wy;            /* _utils.py:200-204 */
    float dw = rc[2] / ww, dh = rc[3] / wh;
    dw = dw > xform_clip ? xform_clip : dw;                  /* _utils.py:207-208 */
    dh = dh > xform_clip ? xform_clip : dh;
    const float mx = dx * w, my = dy * h;
    const float pcx = mx + cx, pcy = my + cy;                /* _utils.py:210-211 */
    const float pw = oracle_det_expf(dw) * w, ph = oracle_det_expf(dh) * h; /* _utils.py:212-213 */
    const float c2w = 0.5f * pw, c2h = 0.5f * ph;            /* _utils.py:216-217 */
    out_boxes[4 * r + 0] = clampf(pcx - c2w, 0.f, (float)img_w); /* _utils.py:219-222, boxes.py clip_boxes_to_image */
    out_boxes[4 * r + 1] = clampf(pcy - c2h, 0.f, (float)img_h);
    out_boxes[4 * r + 2] = clampf(pcx + c2w, 0.f, (float)img_w);
    out_boxes[4 * r + 3] = clampf(pcy + c2h, 0.f, (float)img_h);
    out_scores[r] = cand[r].s;
    out_labels[r] = cand[r].i % num_classes;
  }
  free(cand);
  return k;
}

/* ---------------------------------------------------------------------------------------------------------------
 * Box IoU and the greedy claim walk over its sorted rows.
 * tv:ops/boxes.py box_iou / _box_inter_union: area = (x2-x1)*(y2-y1); lt = max, rb = min, wh = (rb-lt).clamp(min=0);
 * inter = w*h; union = area_a[:,None] + area_b - inter; iou = inter/union (fp32).  Called from cvpce/metrics.py:11-13
 * and cvpce/planograms.py:224.
 * --------------------------------------------------------------------------------------------------------------- */
static float tv_iou(const float* a, const float* b) {
  const float area_a = (a[2] - a[0]) * (a[3] - a[1]);
  const float area_b = (b[2] - b[0]) * (b[3] - b[1]);
  const float ltx = a[0] > b[0] ? a[0] : b[0], lty = a[1] > b[1] ? a[1] : b[1];
  const float rbx = a[2] < b[2] ? a[2] : b[2], rby = a[3] < b[3] ? a[3] : b[3];
  float w = rbx - ltx, h = rby - lty;
  if (w < 0.f) w = 0.f;
  if (h < 0.f) h = 0.f;
  const float inter = w * h;
  const float uni = (area_a + area_b) - inter;
  return inter / uni;
}

API void oracle_box_iou(const float* a, int n_a, const float* b, int n_b, float* out) {
  for (int i = 0; i < n_a; ++i)
    for (int j = 0; j < n_b; ++j) out[(size_t)i * n_b + j] = tv_iou(a + 4 * i, b + 4 * j);
}

typedef struct { float v; int32_t j; } ioukey;

/* torch.sort(descending=True): NaN counts as the largest value; order among equal values is irrelevant to the walk */
static int cmp_iou_desc(const void* x, const void* y) {
  const ioukey* p = (const ioukey*)x;
  const ioukey* q = (const ioukey*)y;
  const int pn = p->v != p->v, qn = q->v != q->v;
  if (pn || qn) return qn - pn;
  if (p->v > q->v) return -1;
  if (p->v < q->v) return 1;
  return (p->j > q->j) - (p->j < q->j);
}

/* cvpce/metrics.py:15-34 (check_matches) and cvpce/planograms.py:214-232, restated literally: for each row, the IoUs
 * against its eligible columns are sorted descending, the walk stops at the first iou < thr (fp32 comparison: a 0-dim
 * fp32 tensor against a Python float), skips used columns and marks every other one -- it does NOT stop at the first
 * hit.  Labels (nullable): planograms.py:214-223 restricts row i to the columns with its label; visiting label by label
 * equals visiting the rows in index order because columns of different labels are disjoint. */
API void oracle_iou_match(const float* rows, const int64_t* row_labels, int n_rows, const float* cols, const int64_t* col_labels,
                          int n_cols, float thr, uint8_t* row_matched, uint8_t* col_matched) {
  ioukey* keys = (ioukey*)malloc(sizeof(ioukey) * (size_t)(n_cols > 0 ? n_cols : 1));
  memset(row_matched, 0, (size_t)n_rows);
  memset(col_matched, 0, (size_t)n_cols);
  for (int i = 0; i < n_rows; ++i) {
    int m = 0;
    for (int j = 0; j < n_cols; ++j) {
      if (row_labels && col_labels && row_labels[i] != col_labels[j]) continue;
      keys[m].v = tv_iou(rows + 4 * i, cols + 4 * j);
      keys[m].j = j;
      ++m;
    }
    qsort(keys, (size_t)m, sizeof(ioukey), cmp_iou_desc);
    for (int t = 0; t < m; ++t) {
      if (keys[t].v < thr) break;
      if (col_matched[keys[t].j]) continue;
      col_matched[keys[t].j] = 1;
      row_matched[i] = 1;
    }
  }
  free(keys);
}
