// 2x2 / stride-2 max pooling of a channels-last map for sm_100a: the pooling layers BETWEEN the convolutions of the
// (not rebuilt) torch VGG16 trunk.  at::max_pool_forward_nhwc, which torch runs for nn.MaxPool2d on channels-last
// input, took 28 % of the GPU time of the whole classification step (profiles/r1b_c3x4shelves_from_head_launches.txt);
// it is a pure HBM streaming op: every input value is read once, every output written once.
// Replaces: nn.MaxPool2d(kernel_size=2, stride=2) of tv:models/vgg.py make_layers ('M' entries of cfg 'D'), reached from
// cvpce/models/classification.py:31-37 (MACVGG.block1 / block2).  Same result bit for bit (max; NaN propagates like
// at::max_pool2d: a NaN in the window wins).
//
// One thread = 16 bytes of channels of one output pixel: four 128-bit loads (2 rows x 2 pixels; consecutive threads read
// consecutive addresses of both input rows), one 128-bit store.  blockIdx.x = output row (b, yo): no integer division
// except one per thread to split (xo, channel group).
#include "common.cuh"
#include <cuda_fp16.h>

namespace cvpce {

__device__ __forceinline__ float max_nan(float a, float b) { return (b > a || b != b) ? b : a; }  // at::max_pool2d's rule

__device__ __forceinline__ float4 ld_stream(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

__device__ __forceinline__ float4 max4_f32(float4 a, float4 b, float4 c, float4 d) {
  float4 r;
  r.x = max_nan(max_nan(a.x, b.x), max_nan(c.x, d.x));
  r.y = max_nan(max_nan(a.y, b.y), max_nan(c.y, d.y));
  r.z = max_nan(max_nan(a.z, b.z), max_nan(c.z, d.z));
  r.w = max_nan(max_nan(a.w, b.w), max_nan(c.w, d.w));
  return r;
}

__device__ __forceinline__ uint32_t max2_bf16(uint32_t a, uint32_t b) {
  const __nv_bfloat162 x = *reinterpret_cast<const __nv_bfloat162*>(&a), y = *reinterpret_cast<const __nv_bfloat162*>(&b);
  const float xl = __low2float(x), xh = __high2float(x), yl = __low2float(y), yh = __high2float(y);
  const __nv_bfloat162 r = __floats2bfloat162_rn(max_nan(xl, yl), max_nan(xh, yh));  // exact: both are bf16 values
  return *reinterpret_cast<const uint32_t*>(&r);
}

__device__ __forceinline__ uint32_t max2_f16(uint32_t a, uint32_t b) {
  const __half2 x = *reinterpret_cast<const __half2*>(&a), y = *reinterpret_cast<const __half2*>(&b);
  const float2 xf = __half22float2(x), yf = __half22float2(y);
  const __half2 r = __floats2half2_rn(max_nan(xf.x, yf.x), max_nan(xf.y, yf.y));  // exact: both are half values
  return *reinterpret_cast<const uint32_t*>(&r);
}

__device__ __forceinline__ float4 max4_f16(float4 a, float4 b, float4 c, float4 d) {
  float4 r;
  uint32_t* ro = reinterpret_cast<uint32_t*>(&r);
  const uint32_t *pa = reinterpret_cast<const uint32_t*>(&a), *pb = reinterpret_cast<const uint32_t*>(&b),
                 *pc = reinterpret_cast<const uint32_t*>(&c), *pd = reinterpret_cast<const uint32_t*>(&d);
#pragma unroll
  for (int i = 0; i < 4; ++i) ro[i] = max2_f16(max2_f16(pa[i], pb[i]), max2_f16(pc[i], pd[i]));
  return r;
}

__device__ __forceinline__ float4 max4_bf16(float4 a, float4 b, float4 c, float4 d) {
  float4 r;
  uint32_t* ro = reinterpret_cast<uint32_t*>(&r);
  const uint32_t *pa = reinterpret_cast<const uint32_t*>(&a), *pb = reinterpret_cast<const uint32_t*>(&b),
                 *pc = reinterpret_cast<const uint32_t*>(&c), *pd = reinterpret_cast<const uint32_t*>(&d);
#pragma unroll
  for (int i = 0; i < 4; ++i) ro[i] = max2_bf16(max2_bf16(pa[i], pb[i]), max2_bf16(pc[i], pd[i]));
  return r;
}

// vec = 16-byte groups per pixel (C*esize/16); row_vecs = Wo * vec
template <int DT>
__global__ void __launch_bounds__(256) maxpool2x2_nhwc_kernel(const float4* __restrict__ in, float4* __restrict__ out, int H, int W,
                                                               int Ho, int Wo, int vec, int row_vecs) {
  const int row = blockIdx.x;  // b * Ho + yo
  const int b = row / Ho, yo = row - b * Ho;
  const float4* r0 = in + ((size_t)b * H + 2 * yo) * W * vec;
  const float4* r1 = r0 + (size_t)W * vec;
  float4* o = out + (size_t)row * row_vecs;
  for (int j = blockIdx.y * blockDim.x + threadIdx.x; j < row_vecs; j += gridDim.y * blockDim.x) {
    const int xo = j / vec, c = j - xo * vec;
    const int i0 = 2 * xo * vec + c;
    const float4 a = ld_stream(r0 + i0), bq = ld_stream(r0 + i0 + vec), cq = ld_stream(r1 + i0), d = ld_stream(r1 + i0 + vec);
    o[j] = DT == CVPCE_BF16 ? max4_bf16(a, bq, cq, d) : (DT == CVPCE_F16 ? max4_f16(a, bq, cq, d) : max4_f32(a, bq, cq, d));
  }
}

}  // namespace cvpce

extern "C" int cvpce_maxpool2x2_nhwc(const void* in, int dtype, int B, int H, int W, int C, void* out, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(B >= 0 && H >= 0 && W >= 0 && C >= 1, CVPCE_ERR_INVALID_ARG, "maxpool2x2: bad shape B=%d H=%d W=%d C=%d", B, H, W, C);
  CVPCE_REQUIRE(dtype == CVPCE_F32 || dtype == CVPCE_BF16 || dtype == CVPCE_F16, CVPCE_ERR_INVALID_ARG, "maxpool2x2: dtype %d", dtype);
  const int Ho = H / 2, Wo = W / 2;  // floor mode, no padding (nn.MaxPool2d defaults)
  if (B == 0 || Ho == 0 || Wo == 0) return CVPCE_OK;
  const int per16 = dtype == CVPCE_F32 ? 4 : 8;
  CVPCE_REQUIRE(C % per16 == 0, CVPCE_ERR_UNSUPPORTED, "maxpool2x2: C=%d must be a multiple of %d (16-byte channel groups)", C, per16);
  CVPCE_REQUIRE(in && out && aligned16(in) && aligned16(out), CVPCE_ERR_INVALID_ARG, "maxpool2x2: null or misaligned pointer");
  CVPCE_REQUIRE((long long)B * Ho < (1ll << 31), CVPCE_ERR_UNSUPPORTED, "maxpool2x2: too many output rows");
  if (int rc = require_sm100()) return rc;
  const int vec = C / per16;
  const long long row_vecs = (long long)Wo * vec;
  CVPCE_REQUIRE(row_vecs < (1ll << 30), CVPCE_ERR_UNSUPPORTED, "maxpool2x2: output row too long");
  int gy = (int)((row_vecs + 255) / 256);
  if (gy > 64) gy = 64;
  dim3 grid((unsigned)(B * Ho), (unsigned)gy, 1);
  const float4* i4 = reinterpret_cast<const float4*>(in);
  float4* o4 = reinterpret_cast<float4*>(out);
  if (dtype == CVPCE_F32)
    maxpool2x2_nhwc_kernel<CVPCE_F32><<<grid, 256, 0, (cudaStream_t)stream>>>(i4, o4, H, W, Ho, Wo, vec, (int)row_vecs);
  else if (dtype == CVPCE_BF16)
    maxpool2x2_nhwc_kernel<CVPCE_BF16><<<grid, 256, 0, (cudaStream_t)stream>>>(i4, o4, H, W, Ho, Wo, vec, (int)row_vecs);
  else
    maxpool2x2_nhwc_kernel<CVPCE_F16><<<grid, 256, 0, (cudaStream_t)stream>>>(i4, o4, H, W, Ho, Wo, vec, (int)row_vecs);
  CVPCE_LAUNCH_OK("maxpool2x2_nhwc_kernel");
  return CVPCE_OK;
}
