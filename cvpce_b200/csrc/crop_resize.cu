// Batched ROI crop -> pad-to-square(0.5) -> (antialiased) bilinear resize to S x S, for sm_100a.
//
// Follows cvpce/production.py:20 (boxes.to(long), python slice image[:, y1:y2, x1:x2]) and
// cvpce/datautils.py:234-239 (resize_for_classification), with the resize arithmetic of ATen's
// _upsample_bilinear2d_aa / upsample_bilinear2d CPU kernels (torchvision ttf.resize -> F.interpolate).
//
// The padded canvas is never materialised: a tap at (y,x) reads the image when y<h and x<w and is the pad value
// otherwise.  One thread produces one output pixel for all 3 channels (the taps and weights are shared by the
// channels); consecutive threads produce consecutive output x, so stores are full 128-byte lines per warp and the
// source taps of a warp fall in 1-3 neighbouring lines that stay in L1 for the next tap/row.
#include "common.cuh"

namespace cvpce {

struct CropParams {
  const int64_t* image_table;  // [n_images,5]
  int n_images;
  const float4* boxes;
  const int32_t* box_image_idx;
  int K, S, tiles;
  float pad;
  float a[3], b[3];
  void* out;
  int32_t* valid;
};

// python slice index normalisation for a dimension of size n
__device__ __forceinline__ long long slice_index(long long i, long long n) {
  if (i < 0) {
    i += n;
    if (i < 0) i = 0;
  }
  return i > n ? n : i;
}

__device__ __forceinline__ long long trunc_to_long(float v) {
  // Tensor.to(torch.long): C truncation toward zero; saturate instead of UB for huge / non-finite values
  if (!(v == v)) return 0;
  if (v >= 9.0e18f) return 9000000000000000000ll;
  if (v <= -9.0e18f) return -9000000000000000000ll;
  return (long long)v;
}

struct TapSet {
  int mn;        // first source index
  int sz;        // number of taps
  float center;  // AA: filter centre;  non-AA: lambda1
  float inv_total;
};

// ATen HelperInterpBase::_compute_indices_min_size_weights_aa restated (bilinear filter, interp_size 2)
__device__ __forceinline__ float aa_filter(float x) {
  x = fabsf(x);
  return x < 1.f ? 1.f - x : 0.f;
}

template <bool AA>
__device__ __forceinline__ TapSet make_taps(int o, int L, float scale, float support, float invscale) {
  TapSet t;
  if (AA) {
    const float center = (float)((double)scale * ((double)o + 0.5));
    long long mn = (long long)((double)(center - support) + 0.5);
    if (mn < 0) mn = 0;
    long long hi = (long long)((double)(center + support) + 0.5);
    if (hi > L) hi = L;
    long long sz = hi - mn;
    const long long max_sz = (long long)ceilf(support) * 2 + 1;
    if (sz < 0) sz = 0;
    if (sz > max_sz) sz = max_sz;
    float total = 0.f;
    for (int j = 0; j < (int)sz; ++j) total += aa_filter(((float)(j + mn) - center + 0.5f) * invscale);
    t.mn = (int)mn;
    t.sz = (int)sz;
    t.center = center;
    t.inv_total = total != 0.f ? 1.f / total : 0.f;
  } else {
    float src = scale * ((float)o + 0.5f) - 0.5f;
    if (src < 0.f) src = 0.f;
    int i0 = (int)src;
    if (i0 > L - 1) i0 = L - 1;
    t.mn = i0;
    t.sz = 2;
    t.center = src - (float)i0;  // lambda1
    t.inv_total = 1.f;
  }
  return t;
}

template <bool AA>
__device__ __forceinline__ float tap_weight(const TapSet& t, int j, float invscale) {
  if (AA) return aa_filter(((float)(j + t.mn) - t.center + 0.5f) * invscale) * t.inv_total;
  return j == 0 ? 1.f - t.center : t.center;
}

template <bool AA>
__device__ __forceinline__ int tap_index(const TapSet& t, int j, int L) {
  if (AA) return t.mn + j;
  const int i = t.mn + j;
  return i < L ? i : L - 1;  // i1 = i0 + (i0 < L-1)
}

constexpr int kCropThreads = 256;
constexpr int kCropPixPerThread = 4;

template <bool AA, int DTYPE, int LAYOUT>
__global__ void __launch_bounds__(kCropThreads) crop_resize_kernel(const CropParams p) {
  const int k = blockIdx.x / p.tiles;
  const int tile = blockIdx.x - k * p.tiles;
  const int S = p.S;
  const int npix = S * S;
  const float4 bx = __ldg(p.boxes + k);
  int img_i = p.box_image_idx ? __ldg(p.box_image_idx + k) : 0;
  if (img_i < 0 || img_i >= p.n_images) img_i = 0;
  const int64_t* row = p.image_table + (size_t)img_i * 5;
  const float* img = reinterpret_cast<const float*>(static_cast<uintptr_t>(__ldg(row + 0)));
  const long long H = __ldg(row + 1), W = __ldg(row + 2);
  const long long sc = __ldg(row + 3), sh = __ldg(row + 4);

  const long long xs = slice_index(trunc_to_long(bx.x), W), xe = slice_index(trunc_to_long(bx.z), W);
  const long long ys = slice_index(trunc_to_long(bx.y), H), ye = slice_index(trunc_to_long(bx.w), H);
  const int w = (int)(xe > xs ? xe - xs : 0), h = (int)(ye > ys ? ye - ys : 0);
  const int L = max(h, w);
  if (tile == 0 && threadIdx.x == 0 && p.valid) p.valid[k] = L > 0 ? 1 : 0;

  const float scale = (float)L / (float)S;
  const float support = (AA && scale >= 1.f) ? scale : 1.f;
  const float invscale = (AA && scale >= 1.f) ? 1.f / scale : 1.f;
  const float pad = p.pad;
  const float* base = img + ys * sh + xs;

  const int pix0 = tile * (kCropThreads * kCropPixPerThread) + threadIdx.x;
#pragma unroll 1
  for (int r = 0; r < kCropPixPerThread; ++r) {
    const int pix = pix0 + r * kCropThreads;
    if (pix >= npix) break;
    const int oy = pix / S, ox = pix - oy * S;
    float acc0 = pad, acc1 = pad, acc2 = pad;
    if (L > 0) {
      const TapSet tx = make_taps<AA>(ox, L, scale, support, invscale);
      const TapSet ty = make_taps<AA>(oy, L, scale, support, invscale);
      acc0 = acc1 = acc2 = 0.f;
      for (int i = 0; i < ty.sz; ++i) {
        const float wy = tap_weight<AA>(ty, i, invscale);
        const int y = tap_index<AA>(ty, i, L);
        const bool rowin = y < h;
        const float* rp = base + (long long)y * sh;
        float r0 = 0.f, r1 = 0.f, r2 = 0.f;
        for (int j = 0; j < tx.sz; ++j) {
          const float wx = tap_weight<AA>(tx, j, invscale);
          const int x = tap_index<AA>(tx, j, L);
          float v0 = pad, v1 = pad, v2 = pad;
          if (rowin && x < w) {
            v0 = __ldg(rp + x);
            v1 = __ldg(rp + sc + x);
            v2 = __ldg(rp + 2 * sc + x);
          }
          r0 = fmaf(wx, v0, r0);
          r1 = fmaf(wx, v1, r1);
          r2 = fmaf(wx, v2, r2);
        }
        acc0 = fmaf(wy, r0, acc0);
        acc1 = fmaf(wy, r1, acc1);
        acc2 = fmaf(wy, r2, acc2);
      }
    }
    acc0 = fmaf(acc0, p.a[0], p.b[0]);
    acc1 = fmaf(acc1, p.a[1], p.b[1]);
    acc2 = fmaf(acc2, p.a[2], p.b[2]);
    if (LAYOUT == CVPCE_NCHW) {
      const size_t o = (size_t)k * 3 * npix + pix;
      if (DTYPE == CVPCE_F32) {
        float* out = reinterpret_cast<float*>(p.out);
        __stcs(out + o, acc0);
        __stcs(out + o + npix, acc1);
        __stcs(out + o + 2 * (size_t)npix, acc2);
      } else {
        __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.out);
        out[o] = __float2bfloat16_rn(acc0);
        out[o + npix] = __float2bfloat16_rn(acc1);
        out[o + 2 * (size_t)npix] = __float2bfloat16_rn(acc2);
      }
    } else {
      const size_t o = ((size_t)k * npix + pix) * 3;
      if (DTYPE == CVPCE_F32) {
        float* out = reinterpret_cast<float*>(p.out);
        out[o] = acc0;
        out[o + 1] = acc1;
        out[o + 2] = acc2;
      } else {
        __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.out);
        out[o] = __float2bfloat16_rn(acc0);
        out[o + 1] = __float2bfloat16_rn(acc1);
        out[o + 2] = __float2bfloat16_rn(acc2);
      }
    }
  }
}

template <bool AA>
static int launch_crop(const CropParams& p, int out_dtype, int out_layout, cudaStream_t st) {
  dim3 grid((unsigned)((long long)p.tiles * p.K));
  if (out_dtype == CVPCE_F32 && out_layout == CVPCE_NCHW)
    crop_resize_kernel<AA, CVPCE_F32, CVPCE_NCHW><<<grid, kCropThreads, 0, st>>>(p);
  else if (out_dtype == CVPCE_F32)
    crop_resize_kernel<AA, CVPCE_F32, CVPCE_NHWC><<<grid, kCropThreads, 0, st>>>(p);
  else if (out_layout == CVPCE_NCHW)
    crop_resize_kernel<AA, CVPCE_BF16, CVPCE_NCHW><<<grid, kCropThreads, 0, st>>>(p);
  else
    crop_resize_kernel<AA, CVPCE_BF16, CVPCE_NHWC><<<grid, kCropThreads, 0, st>>>(p);
  CVPCE_LAUNCH_OK("crop_resize_kernel");
  return CVPCE_OK;
}

}  // namespace cvpce

extern "C" int cvpce_crop_resize(const int64_t* image_table, int n_images, const float* boxes,
                                 const int32_t* box_image_idx, int K, int S, int antialias, float pad_value,
                                 const float* affine, int out_dtype, int out_layout, void* out, int32_t* valid,
                                 void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(K >= 0 && S > 0 && S <= 4096, CVPCE_ERR_INVALID_ARG, "crop_resize: bad K=%d or S=%d", K, S);
  CVPCE_REQUIRE(out_dtype == CVPCE_F32 || out_dtype == CVPCE_BF16, CVPCE_ERR_INVALID_ARG, "crop_resize: bad out_dtype %d", out_dtype);
  CVPCE_REQUIRE(out_layout == CVPCE_NCHW || out_layout == CVPCE_NHWC, CVPCE_ERR_INVALID_ARG, "crop_resize: bad out_layout %d", out_layout);
  if (K == 0) return CVPCE_OK;
  CVPCE_REQUIRE(image_table && boxes && out && n_images > 0, CVPCE_ERR_INVALID_ARG, "crop_resize: null pointer or n_images<=0");
  CVPCE_REQUIRE(aligned16(boxes), CVPCE_ERR_INVALID_ARG, "crop_resize: boxes must be 16-byte aligned");
  if (int rc = require_sm100()) return rc;
  CropParams p;
  p.image_table = image_table;
  p.n_images = n_images;
  p.boxes = reinterpret_cast<const float4*>(boxes);
  p.box_image_idx = box_image_idx;
  p.K = K;
  p.S = S;
  p.tiles = (S * S + kCropThreads * kCropPixPerThread - 1) / (kCropThreads * kCropPixPerThread);
  CVPCE_REQUIRE((long long)p.tiles * K < 2147483647ll, CVPCE_ERR_UNSUPPORTED, "crop_resize: K*tiles too large");
  p.pad = pad_value;
  for (int c = 0; c < 3; ++c) {
    p.a[c] = affine ? affine[2 * c] : 1.f;
    p.b[c] = affine ? affine[2 * c + 1] : 0.f;
  }
  p.out = out;
  p.valid = valid;
  return antialias ? launch_crop<true>(p, out_dtype, out_layout, (cudaStream_t)stream)
                   : launch_crop<false>(p, out_dtype, out_layout, (cudaStream_t)stream);
}
