// Batched ROI crop -> pad-to-square(0.5) -> (antialiased) bilinear resize to S x S, for sm_100a.
//
// Follows cvpce/production.py:20 (boxes.to(long), python slice image[:, y1:y2, x1:x2]) and
// cvpce/datautils.py:234-239 (resize_for_classification), with the resize arithmetic of ATen's
// _upsample_bilinear2d_aa / upsample_bilinear2d CPU kernels (torchvision ttf.resize -> F.interpolate).
//
// The padded canvas is never materialised: a tap at (y,x) reads the image when y<h and x<w and is the pad value
// otherwise.
//
// Two kernels:
//  * crop_resize_col_kernel (antialias, S <= 1024: the production case).  The triangle filter is separable and, down a
//    column of the output, the vertical taps slide over consecutive source rows.  A thread owns ONE output column of a
//    band of output rows and marches down it: each source row is filtered horizontally once (the column's taps and
//    weights live in registers, read through the read-only path; a warp reads a contiguous span of the row) into a
//    register ring of the last NT filtered rows, and each output row is the ring weighted by that row's vertical taps
//    (broadcast reads of a small shared-memory table) -> affine -> one store per pixel (a warp writes 512 contiguous
//    bytes in the 4-channel layout).  No shared-memory traffic for pixel data, no barrier inside the walk.  The band's
//    source rows are read exactly once, so they are prefetched into L1 six rows ahead and the next row's taps are
//    loaded while the current row is filtered.  CTA-uniform specialisation on the crop's widest filter (2 / 4 / 6 taps:
//    up-sampling, down-sampling by <= 2x, <= 3x); wider filters (boxes more than 3 S pixels long) take direct evaluation
//    inside the kernel.  Round 1's shared-memory band kernel (2.56 ms for 6400 4-channel crops, long_scoreboard-bound,
//    ~20 thread instructions per output value) was replaced by this one (1.32 ms, 0.93 of HBM on the bytes it moves).
//  * crop_resize_kernel (direct form: one thread per output pixel, all taps): plain bilinear (antialias=0, 4 taps) and
//    S > 1024.
#include "common.cuh"
#include <cuda_fp16.h>

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

// direct evaluation of one output pixel (shared by the direct kernel and the band kernel's fallback)
template <bool AA>
__device__ __forceinline__ void direct_pixel(const float* __restrict__ base, long long sh, long long sc, int h, int w, int L, int ox,
                                             int oy, float scale, float support, float invscale, float pad, float& acc0,
                                             float& acc1, float& acc2) {
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

// two fp32 values -> one 32-bit word of 16-bit outputs (bf16 or IEEE half, round to nearest even)
template <int DTYPE>
__device__ __forceinline__ uint32_t pack16x2(float lo, float hi) {
  if (DTYPE == CVPCE_F16) {
    const __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
  }
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}
template <int DTYPE>
__device__ __forceinline__ unsigned short to16(float x) {
  if (DTYPE == CVPCE_F16) return __half_as_ushort(__float2half_rn(x));
  return __bfloat16_as_ushort(__float2bfloat16_rn(x));
}

template <int DTYPE, int LAYOUT>
__device__ __forceinline__ void store_pixel(void* outp, size_t k, int npix, int pix, float a0, float a1, float a2) {
  if (LAYOUT == CVPCE_NCHW) {
    const size_t o = k * 3 * (size_t)npix + pix;
    if (DTYPE == CVPCE_F32) {
      float* out = reinterpret_cast<float*>(outp);
      __stcs(out + o, a0);
      __stcs(out + o + npix, a1);
      __stcs(out + o + 2 * (size_t)npix, a2);
    } else {
      unsigned short* out = reinterpret_cast<unsigned short*>(outp);
      out[o] = to16<DTYPE>(a0);
      out[o + npix] = to16<DTYPE>(a1);
      out[o + 2 * (size_t)npix] = to16<DTYPE>(a2);
    }
  } else if (LAYOUT == CVPCE_NHWC4) {  // 4 channels per pixel, the 4th is zero: one aligned vector store
    const size_t o = (k * (size_t)npix + pix) * 4;
    if (DTYPE == CVPCE_F32) {
      __stcs(reinterpret_cast<float4*>(reinterpret_cast<float*>(outp) + o), make_float4(a0, a1, a2, 0.f));
    } else {
      __stcs(reinterpret_cast<uint2*>(reinterpret_cast<unsigned short*>(outp) + o),
             make_uint2(pack16x2<DTYPE>(a0, a1), pack16x2<DTYPE>(a2, 0.f)));
    }
  } else {
    const size_t o = (k * (size_t)npix + pix) * 3;
    if (DTYPE == CVPCE_F32) {
      float* out = reinterpret_cast<float*>(outp);
      out[o] = a0;
      out[o + 1] = a1;
      out[o + 2] = a2;
    } else {
      unsigned short* out = reinterpret_cast<unsigned short*>(outp);
      out[o] = to16<DTYPE>(a0);
      out[o + 1] = to16<DTYPE>(a1);
      out[o + 2] = to16<DTYPE>(a2);
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------------
// column-marching kernel (antialias)
// ------------------------------------------------------------------------------------------------------------------------
constexpr int kColMaxTaps = 6;    // widest filter the register ring holds (down-sampling by up to 3x)
constexpr int kColMaxBand = 256;  // output rows per CTA (shared-memory tables of the vertical taps)

struct ColLayout {
  int band_rows;  // output rows per CTA
  int bands;      // CTAs per crop
};

// one output pixel (3 channels) at a precomputed byte address; NCHW: channel planes plane_bytes apart
template <int DTYPE, int LAYOUT>
__device__ __forceinline__ void store_pixel_at(unsigned char* at, size_t plane_bytes, float a0, float a1, float a2) {
  if (LAYOUT == CVPCE_NCHW) {
    if (DTYPE == CVPCE_F32) {
      __stcs(reinterpret_cast<float*>(at), a0);
      __stcs(reinterpret_cast<float*>(at + plane_bytes), a1);
      __stcs(reinterpret_cast<float*>(at + 2 * plane_bytes), a2);
    } else {
      *reinterpret_cast<unsigned short*>(at) = to16<DTYPE>(a0);
      *reinterpret_cast<unsigned short*>(at + plane_bytes) = to16<DTYPE>(a1);
      *reinterpret_cast<unsigned short*>(at + 2 * plane_bytes) = to16<DTYPE>(a2);
    }
  } else if (LAYOUT == CVPCE_NHWC4) {
    if (DTYPE == CVPCE_F32) __stcs(reinterpret_cast<float4*>(at), make_float4(a0, a1, a2, 0.f));
    else __stcs(reinterpret_cast<uint2*>(at), make_uint2(pack16x2<DTYPE>(a0, a1), pack16x2<DTYPE>(a2, 0.f)));
  } else {
    if (DTYPE == CVPCE_F32) {
      float* o = reinterpret_cast<float*>(at);
      o[0] = a0;
      o[1] = a1;
      o[2] = a2;
    } else {
      unsigned short* o = reinterpret_cast<unsigned short*>(at);
      o[0] = to16<DTYPE>(a0);
      o[1] = to16<DTYPE>(a1);
      o[2] = to16<DTYPE>(a2);
    }
  }
}

// One band of one crop.  NT: compile-time upper bound of the tap count of every column and row of this crop (CTA-uniform).
template <int NT, int ST, int DTYPE, int LAYOUT>
__device__ __forceinline__ void col_walk(const CropParams& p, size_t k, const float* __restrict__ base, long long sc, long long sh,
                                         int h, int w, float invscale, int b0, int nb, const TapSet tx, float* s_yw, int* s_need,
                                         const int* s_ymn, const int* s_ysz, const float* s_yc, const float* s_yinv) {
  constexpr bool PF = NT <= 4;  // software prefetch of the next source row's taps (the wide ring has no registers to spare)
  const int S = ST ? ST : p.S;
  const int x = threadIdx.x;
  const float pad = p.pad;
  // ---- vertical taps of the band, right-aligned to the ring: ring slot j of output row r holds source row
  //      ymn + ysz - NT + j, so slots below NT - ysz get weight 0
  for (int o = x; o < nb; o += blockDim.x) {
    TapSet t;
    t.mn = s_ymn[o];
    t.sz = s_ysz[o];
    t.center = s_yc[o];
    t.inv_total = s_yinv[o];
    s_need[o] = t.mn + t.sz;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int i = j - (NT - t.sz);
      s_yw[o * NT + j] = i >= 0 ? tap_weight<true>(t, i, invscale) : 0.f;
    }
  }
  // ---- this thread's column.  Taps [0, hin) of the column read the image, taps [hin, hsz) the pad canvas (a constant
  //      contribution, wpad).  Every thread loads a window of NT CONSECUTIVE image columns [c0, c0 + NT) that lies inside
  //      the crop and contains its image taps (the window is shifted left at the right edge; weights of window positions
  //      that are not taps are 0), so the loads have constant offsets and no lane diverges.  Crops narrower than NT
  //      pixels (CTA-uniform) clamp the window positions instead.
  const bool wide_enough = w >= NT;
  const int hin = (x < S && h > 0) ? max(0, min(tx.sz, w - tx.mn)) : 0;
  const int c0 = wide_enough ? max(0, min(tx.mn, w - NT)) : 0;
  const int shift = hin > 0 ? tx.mn - c0 : 0;  // window position of tap 0
  float wx[NT];
  float wpad = 0.f, padsum = 0.f;
#pragma unroll
  for (int j = 0; j < NT; ++j) wx[j] = 0.f;
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const float wj = (x < S && j < tx.sz) ? tap_weight<true>(tx, j, invscale) : 0.f;
    padsum = fmaf(wj, pad, padsum);
    if (j >= hin) wpad = fmaf(wj, pad, wpad);
    else {
      const int pos = wide_enough ? j + shift : min(tx.mn + j, max(w - 1, 0));  // narrow crops: weight goes to the clamped column
#pragma unroll
      for (int q = 0; q < NT; ++q)
        if (q == pos) wx[q] += wj;
    }
  }
  const int heff = w > 0 ? h : 0;  // rows [0, heff) of the crop hold image pixels
  const int top0 = s_ymn[0];       // first source row the band needs
  __syncthreads();
  if (x >= S) return;
  const int last_need = min(s_need[nb - 1], heff);  // rows >= last_need are never loaded

  // row pointers of the three channels at this thread's window, advanced by one source row per push
  const float* q0 = base + (long long)top0 * sh + c0;
  const float* q1 = q0 + sc;
  const float* q2 = q1 + sc;
  // L1 prefetch kPrefetchRows source rows ahead of the row being loaded: the band's source rows are read exactly once, so
  // without it every row costs a full DRAM round trip per warp (ncu: 52 % of the stall samples on the first use of a load)
  constexpr int kPrefetchRows = 6;
  const long long pf_step = (long long)kPrefetchRows * sh;
  auto prefetch_row = [&](const float* r0) {
    asm volatile("prefetch.global.L1 [%0];" ::"l"(r0));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(r0 + sc));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(r0 + 2 * sc));
  };
  for (int i = 1; i < kPrefetchRows; ++i)
    if (top0 + i < last_need) prefetch_row(q0 + (long long)i * sh);
  float v[PF ? NT : 1][3];  // the window of source row `top`, loaded one row ahead (PF only)
#pragma unroll
  for (int j = 0; j < (PF ? NT : 1); ++j) v[j][0] = v[j][1] = v[j][2] = 0.f;
  auto load_row = [&]() {
    if (!PF) return;
    if (wide_enough) {
#pragma unroll
      for (int j = 0; j < (PF ? NT : 1); ++j) {
        v[j][0] = __ldg(q0 + j);
        v[j][1] = __ldg(q1 + j);
        v[j][2] = __ldg(q2 + j);
      }
    } else {
#pragma unroll
      for (int j = 0; j < (PF ? NT : 1); ++j) {
        const int o = min(j, w - 1);
        v[j][0] = __ldg(q0 + o);
        v[j][1] = __ldg(q1 + o);
        v[j][2] = __ldg(q2 + o);
      }
    }
  };
  int top = top0;  // next source row to filter; rows < top are in the ring or no longer needed
  if (PF && top < last_need) load_row();

  float ring[NT][3];
#pragma unroll
  for (int j = 0; j < NT; ++j) ring[j][0] = ring[j][1] = ring[j][2] = 0.f;
  const float a0 = p.a[0], a1 = p.a[1], a2 = p.a[2], o0 = p.b[0], o1 = p.b[1], o2 = p.b[2];
  const size_t npix = (size_t)S * S;
  const size_t esz = DTYPE == CVPCE_F32 ? 4 : 2;
  const size_t px = LAYOUT == CVPCE_NCHW ? 1 : (LAYOUT == CVPCE_NHWC4 ? 4 : 3);
  unsigned char* outp = reinterpret_cast<unsigned char*>(p.out) +
                        ((LAYOUT == CVPCE_NCHW ? k * 3 * npix : k * npix * px) + ((size_t)b0 * S + x) * px) * esz;
  const size_t out_row_bytes = (size_t)S * px * esz;
  const size_t plane_bytes = npix * esz;
  const int* need_p = s_need;
  const float* wr = s_yw;
#pragma unroll 1
  for (int r = 0; r < nb; ++r, ++need_p, wr += NT) {
    const int need = *need_p;  // tap ranges are monotone in the output row
#pragma unroll 1
    while (top < need) {  // CTA-uniform
#pragma unroll
      for (int j = 0; j + 1 < NT; ++j) {
        ring[j][0] = ring[j + 1][0];
        ring[j][1] = ring[j + 1][1];
        ring[j][2] = ring[j + 1][2];
      }
      float t0 = padsum, t1 = padsum, t2 = padsum;
      if (top < heff) {
        t0 = t1 = t2 = wpad;
        if (PF) {
#pragma unroll
          for (int j = 0; j < NT; ++j) {
            t0 = fmaf(wx[j], v[j][0], t0);
            t1 = fmaf(wx[j], v[j][1], t1);
            t2 = fmaf(wx[j], v[j][2], t2);
          }
        } else {  // wide filters: two window positions in flight at a time keep the register ring the only large array
#pragma unroll
          for (int j = 0; j < NT; j += 2) {
            const int oa = wide_enough ? j : min(j, w - 1), ob = wide_enough ? j + 1 : min(j + 1, w - 1);
            const float xa0 = __ldg(q0 + oa), xa1 = __ldg(q1 + oa), xa2 = __ldg(q2 + oa);
            const float xb0 = __ldg(q0 + ob), xb1 = __ldg(q1 + ob), xb2 = __ldg(q2 + ob);
            t0 = fmaf(wx[j + 1], xb0, fmaf(wx[j], xa0, t0));
            t1 = fmaf(wx[j + 1], xb1, fmaf(wx[j], xa1, t1));
            t2 = fmaf(wx[j + 1], xb2, fmaf(wx[j], xa2, t2));
          }
        }
      }
      ring[NT - 1][0] = t0;
      ring[NT - 1][1] = t1;
      ring[NT - 1][2] = t2;
      ++top;
      if (top + kPrefetchRows - 1 < last_need) prefetch_row(q0 + pf_step);
      q0 += sh;
      q1 += sh;
      q2 += sh;
      if (PF && top < last_need) load_row();
    }
    float c0v = 0.f, c1v = 0.f, c2v = 0.f;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const float wj = wr[j];
      c0v = fmaf(wj, ring[j][0], c0v);
      c1v = fmaf(wj, ring[j][1], c1v);
      c2v = fmaf(wj, ring[j][2], c2v);
    }
    store_pixel_at<DTYPE, LAYOUT>(outp, plane_bytes, fmaf(c0v, a0, o0), fmaf(c1v, a1, o1), fmaf(c2v, a2, o2));
    outp += out_row_bytes;
  }
}

template <int ST, int DTYPE, int LAYOUT>
__global__ void __launch_bounds__(ST ? ((ST + 31) / 32 * 32) : 1024, ST ? 4 : 1) crop_resize_col_kernel(const CropParams p, const ColLayout lay) {
  __shared__ __align__(16) float s_yw[kColMaxBand * kColMaxTaps];
  __shared__ int s_ymn[kColMaxBand], s_ysz[kColMaxBand], s_need[kColMaxBand];
  __shared__ float s_yc[kColMaxBand], s_yinv[kColMaxBand];
  __shared__ int s_max_taps;
  const int S = ST ? ST : p.S;
  const int k = blockIdx.x / lay.bands;
  const int band = blockIdx.x - k * lay.bands;
  const int b0 = band * lay.band_rows;
  const int nb = min(S, b0 + lay.band_rows) - b0;
  const int tid = threadIdx.x;

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
  if (band == 0 && tid == 0 && p.valid) p.valid[k] = L > 0 ? 1 : 0;
  const float pad = p.pad;
  const int npix = S * S;

  if (L == 0) {  // empty crop: pad value through the affine
    const float e0 = fmaf(pad, p.a[0], p.b[0]), e1 = fmaf(pad, p.a[1], p.b[1]), e2 = fmaf(pad, p.a[2], p.b[2]);
    if (tid < S)
      for (int r = 0; r < nb; ++r) store_pixel<DTYPE, LAYOUT>(p.out, (size_t)k, npix, (b0 + r) * S + tid, e0, e1, e2);
    return;
  }
  const float scale = (float)L / (float)S;
  const float support = scale >= 1.f ? scale : 1.f;
  const float invscale = scale >= 1.f ? 1.f / scale : 1.f;
  const float* base = img + ys * sh + xs;

  // ---- tap sets: this thread's column, and the band's rows (one per thread, kept in shared memory)
  if (tid == 0) s_max_taps = 0;
  __syncthreads();
  TapSet tx;
  tx.mn = 0; tx.sz = 0; tx.center = 0.f; tx.inv_total = 0.f;
  int my_max = 0;
  if (tid < S) {
    tx = make_taps<true>(tid, L, scale, support, invscale);
    my_max = tx.sz;
  }
  for (int o = tid; o < nb; o += blockDim.x) {
    const TapSet t = make_taps<true>(b0 + o, L, scale, support, invscale);
    my_max = max(my_max, t.sz);
    s_ymn[o] = t.mn;
    s_ysz[o] = t.sz;
    s_yc[o] = t.center;
    s_yinv[o] = t.inv_total;
  }
  my_max = __reduce_max_sync(0xffffffffu, my_max);
  if ((tid & 31) == 0 && my_max > 0) atomicMax(&s_max_taps, my_max);
  __syncthreads();
  const int max_taps = s_max_taps;
  // CTA-uniform specialisation on the widest filter of this crop (2: up-sampling; 4: down-sampling by <= 2x; ...)
  if (max_taps <= 2) col_walk<2, ST, DTYPE, LAYOUT>(p, (size_t)k, base, sc, sh, h, w, invscale, b0, nb, tx, s_yw, s_need, s_ymn, s_ysz, s_yc, s_yinv);
  else if (max_taps <= 4) col_walk<4, ST, DTYPE, LAYOUT>(p, (size_t)k, base, sc, sh, h, w, invscale, b0, nb, tx, s_yw, s_need, s_ymn, s_ysz, s_yc, s_yinv);
  else if (max_taps <= kColMaxTaps) col_walk<kColMaxTaps, ST, DTYPE, LAYOUT>(p, (size_t)k, base, sc, sh, h, w, invscale, b0, nb, tx, s_yw, s_need, s_ymn, s_ysz, s_yc, s_yinv);
  else {  // filters wider than the register ring (down-sampling by more than 3x): direct evaluation of the band
    for (int it = tid; it < nb * S; it += blockDim.x) {
      const int r = it / S, ox = it - r * S;
      float v0, v1, v2;
      direct_pixel<true>(base, sh, sc, h, w, L, ox, b0 + r, scale, support, invscale, pad, v0, v1, v2);
      store_pixel<DTYPE, LAYOUT>(p.out, (size_t)k, npix, (b0 + r) * S + ox, fmaf(v0, p.a[0], p.b[0]), fmaf(v1, p.a[1], p.b[1]),
                                 fmaf(v2, p.a[2], p.b[2]));
    }
  }
}

template <int ST, int DTYPE, int LAYOUT>
static int launch_col_st(const CropParams& p, const ColLayout& lay, cudaStream_t st) {
  const int threads = (p.S + 31) / 32 * 32;
  crop_resize_col_kernel<ST, DTYPE, LAYOUT><<<(unsigned)((long long)p.K * lay.bands), threads, 0, st>>>(p, lay);
  CVPCE_LAUNCH_OK("crop_resize_col_kernel");
  return CVPCE_OK;
}

template <int DTYPE, int LAYOUT>
static int launch_col_one(const CropParams& p, const ColLayout& lay, cudaStream_t st) {
  if (p.S == 256) return launch_col_st<256, DTYPE, LAYOUT>(p, lay, st);
  if (p.S == 224) return launch_col_st<224, DTYPE, LAYOUT>(p, lay, st);
  return launch_col_st<0, DTYPE, LAYOUT>(p, lay, st);
}

// Returns -1 when the shape is outside what the column kernel handles.
static int launch_col(const CropParams& p, int out_dtype, int out_layout, cudaStream_t st) {
  const int S = p.S;
  if (S < 1 || S > 1024) return -1;
  ColLayout lay;
  // band height: 64 output rows when there are enough crops to fill the machine several times, fewer rows otherwise
  int band_rows = tuning("crop_band_rows");
  if (band_rows < 8 || band_rows > kColMaxBand || (band_rows & (band_rows - 1))) band_rows = 64;
  const long long want_ctas = 8ll * sm_count();
  while (band_rows > 8 && (long long)p.K * ((S + band_rows - 1) / band_rows) < want_ctas) band_rows >>= 1;
  lay.band_rows = band_rows;
  lay.bands = (S + band_rows - 1) / band_rows;
  if ((long long)p.K * lay.bands >= 2147483647ll) return -1;
  if (out_layout == CVPCE_NHWC4) {
    if (out_dtype == CVPCE_F32) return launch_col_one<CVPCE_F32, CVPCE_NHWC4>(p, lay, st);
    if (out_dtype == CVPCE_F16) return launch_col_one<CVPCE_F16, CVPCE_NHWC4>(p, lay, st);
    return launch_col_one<CVPCE_BF16, CVPCE_NHWC4>(p, lay, st);
  }
  if (out_layout == CVPCE_NCHW) {
    if (out_dtype == CVPCE_F32) return launch_col_one<CVPCE_F32, CVPCE_NCHW>(p, lay, st);
    if (out_dtype == CVPCE_F16) return launch_col_one<CVPCE_F16, CVPCE_NCHW>(p, lay, st);
    return launch_col_one<CVPCE_BF16, CVPCE_NCHW>(p, lay, st);
  }
  if (out_dtype == CVPCE_F32) return launch_col_one<CVPCE_F32, CVPCE_NHWC>(p, lay, st);
  if (out_dtype == CVPCE_F16) return launch_col_one<CVPCE_F16, CVPCE_NHWC>(p, lay, st);
  return launch_col_one<CVPCE_BF16, CVPCE_NHWC>(p, lay, st);
}

// ------------------------------------------------------------------------------------------------------------------------
// direct kernel: one thread per output pixel
// ------------------------------------------------------------------------------------------------------------------------
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
    if (L > 0) direct_pixel<AA>(base, sh, sc, h, w, L, ox, oy, scale, support, invscale, pad, acc0, acc1, acc2);
    store_pixel<DTYPE, LAYOUT>(p.out, (size_t)k, npix, pix, fmaf(acc0, p.a[0], p.b[0]), fmaf(acc1, p.a[1], p.b[1]),
                               fmaf(acc2, p.a[2], p.b[2]));
  }
}

template <bool AA>
static int launch_crop(const CropParams& p, int out_dtype, int out_layout, cudaStream_t st) {
  dim3 grid((unsigned)((long long)p.tiles * p.K));
  if (out_layout == CVPCE_NHWC4 && out_dtype == CVPCE_F32)
    crop_resize_kernel<AA, CVPCE_F32, CVPCE_NHWC4><<<grid, kCropThreads, 0, st>>>(p);
  else if (out_layout == CVPCE_NHWC4 && out_dtype == CVPCE_F16)
    crop_resize_kernel<AA, CVPCE_F16, CVPCE_NHWC4><<<grid, kCropThreads, 0, st>>>(p);
  else if (out_layout == CVPCE_NHWC4)
    crop_resize_kernel<AA, CVPCE_BF16, CVPCE_NHWC4><<<grid, kCropThreads, 0, st>>>(p);
  else if (out_dtype == CVPCE_F32 && out_layout == CVPCE_NCHW)
    crop_resize_kernel<AA, CVPCE_F32, CVPCE_NCHW><<<grid, kCropThreads, 0, st>>>(p);
  else if (out_dtype == CVPCE_F32)
    crop_resize_kernel<AA, CVPCE_F32, CVPCE_NHWC><<<grid, kCropThreads, 0, st>>>(p);
  else if (out_dtype == CVPCE_F16 && out_layout == CVPCE_NCHW)
    crop_resize_kernel<AA, CVPCE_F16, CVPCE_NCHW><<<grid, kCropThreads, 0, st>>>(p);
  else if (out_dtype == CVPCE_F16)
    crop_resize_kernel<AA, CVPCE_F16, CVPCE_NHWC><<<grid, kCropThreads, 0, st>>>(p);
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
  CVPCE_REQUIRE(out_dtype == CVPCE_F32 || out_dtype == CVPCE_BF16 || out_dtype == CVPCE_F16, CVPCE_ERR_INVALID_ARG,
                "crop_resize: bad out_dtype %d", out_dtype);
  CVPCE_REQUIRE(out_layout == CVPCE_NCHW || out_layout == CVPCE_NHWC || out_layout == CVPCE_NHWC4, CVPCE_ERR_INVALID_ARG,
                "crop_resize: bad out_layout %d", out_layout);
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
  // kernel choice: the column-marching kernel for every antialias shape it covers (S <= 1024); cvpce_tuning_set("crop_kernel", 3)
  // forces the direct kernel (cross-check of the two kernels)
  if (antialias && tuning("crop_kernel") != 3) {
    const int rc = launch_col(p, out_dtype, out_layout, (cudaStream_t)stream);
    if (rc >= 0) return rc;
  }
  return antialias ? launch_crop<true>(p, out_dtype, out_layout, (cudaStream_t)stream)
                   : launch_crop<false>(p, out_dtype, out_layout, (cudaStream_t)stream);
}
