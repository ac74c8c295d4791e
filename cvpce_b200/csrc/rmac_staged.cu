// R-MAC region pooling (regions + per-region L2 + sum + final L2) as a persistent, TMA-staged streaming kernel for sm_100a.
//
// Follows the same semantics as region_pool.cu (cvpce/models/classification.py:46-51 generalised to region lists; the
// north-star's "R-MAC kernel staging VGG16 conv5 feature tiles in shared memory, warp-shuffle max and norm reductions").
//
// Why a second kernel: the cluster kernel of region_pool.cu gives every CTA 1/8 of one image (128 KB at the VGG shapes)
// between barriers -- 0.48-0.51 of HBM channels-last, 0.17 NCHW (4-byte loads).  Here ONE persistent CTA per SM owns whole
// images, and the loads are taken off the threads entirely:
//  * a producer thread streams the image's maps, 16 KB at a time, into a shared-memory ring with cp.async.bulk (the TMA
//    engine, mbarrier complete_tx) and runs ahead across map and image boundaries, so the finishing phase of image n
//    (norms, descriptor store) overlaps the loads of image n + 1;
//  * channels-last maps: within a ROW SEGMENT (the rows between two region boundaries) the rows are folded into one row
//    of running maxima element by element -- no boundary inside a row matters yet, so the streaming part is branch-free:
//    each of the 512 consumer threads owns up to two 16-byte vectors of every 16 KB chunk of a row, keeps their maxima in
//    registers (LDS.128 + 4 FMNMX / HMNMX2 per vector) and the whole CTA writes that row of column maxima to shared
//    memory when the row segment ends; only then -- 6 times per map for three R-MAC levels -- are the regions that
//    contain the row segment updated from it (a group of threads per region, a thread per 16-byte channel group).
//    The first versions walked the pixels with one boundary test per pixel, then per run of pixels: with one warp per
//    scheduler that bookkeeping cost 280 cycles per run and 520 per chunk (cycle counters below, tools/rmac_prof.py),
//    0.13 ms per image whatever the bytes;
//  * NCHW maps: a consumer warp owns a group of staged planes, a lane one 16-byte strip of columns of one plane: it walks
//    down the rows (LDS.128 per row), the strip's maxima of a row segment are written back in place over the plane,
//    then one lane per (plane, cell) and one per (plane, region) finish the plane;
//  * all channels of an image are in one CTA, so the per-region L2 norms and the final L2 norm are block reductions (no
//    cluster, no workspace, no second launch).
// Every feature value is read from HBM exactly once.
// Two kernels share the code: rmac_staged_kernel (16 consumer warps do everything; any mix of layouts) and
// rmac_staged_ws_kernel (channels-last only: 16 streaming warps that never stop + 8 region warps; the automatic choice from
// 1.2 MB per image).  DESIGN.md section 4.3 has the measurements, including the variants that were slower.
#include "common.cuh"
#include <cuda_bf16.h>
#include <math.h>
#include <string.h>
#include <algorithm>

namespace cvpce {

constexpr int kRmacConsWarps = 16;                        // consumer warps
constexpr int kRmacCons = kRmacConsWarps * 32;
constexpr int kRmacThreads = kRmacCons + 32;              // + the producer warp
constexpr int kRmacMaxSegs = 24;                          // elementary segments per axis
constexpr int kRmacMaxStages = 8;
#ifndef CVPCE_RMAC_STAGE_BYTES
#define CVPCE_RMAC_STAGE_BYTES 16384
#endif
constexpr int kRmacStageBytes = CVPCE_RMAC_STAGE_BYTES;   // one ring stage = one bulk copy (a multiple of 8 KB: whole vectors per consumer thread)
constexpr int kRmacStageVecs = kRmacStageBytes / 16;
constexpr int kRmacVecsPerThread = kRmacStageVecs / kRmacCons;  // 16-byte vectors of a chunk per consumer thread (channels-last maps)
static_assert(kRmacVecsPerThread * kRmacCons == kRmacStageVecs, "a stage must be a whole number of vectors per consumer thread");
constexpr int kRmacFullBars = 16;                         // "full" barriers are tied to the chunk index mod 16 (see the NCHW consumer)
constexpr int kRmacMaxRegions = CVPCE_POOL_MAX_REGIONS;   // per map
constexpr int kRmacMaxRegionsTotal = 64;
constexpr int kRmacMaxRowParts = 4;                       // channels-last rows of up to 4 chunks (64 KB)

// development aid (build with -DCVPCE_RMAC_PROF, tools/rmac_prof.py): cycle counts of block 0 (producer thread; thread 0 of the consumers)
#ifdef CVPCE_RMAC_PROF
__device__ long long g_rmac_prof[16];
#define RM_PROF_ADD(k, v) do { if (blockIdx.x == 0) g_rmac_prof[k] += (v); } while (0)
#define RM_CLOCK() clock64()
#else
#define RM_PROF_ADD(k, v) do { (void)(v); } while (0)
#define RM_CLOCK() 0ll
#endif

struct RmacMap {
  const void* data;
  long long image_bytes;
  int C, H, W;
  int nhwc;          // 1: dense channels-last; 0: dense NCHW
  int n_regions, region_begin, out_offset;
  int nys;           // row segments (both layouts)
  int n_chunks;      // per image
  // channels-last: a chunk is `rows_per_chunk` whole rows (row_parts == 1) or one of the `row_parts` 16 KB parts of a row
  int row_bytes, row_parts, rows_per_chunk;
  int pix_vecs;      // 16-byte vectors per pixel
  // NCHW: a chunk is `upc` whole planes; a warp pass handles `ppp` planes, `lpr` lanes per row (16 bytes per lane)
  int plane_bytes, upc, lpr, ppp, nxs;
  unsigned char xb[kRmacMaxSegs + 2], yb[kRmacMaxSegs + 2];  // segment s = [b[s], b[s+1]), b[0] = 0, b[n] = W / H (one spare entry: the walkers read b[n+1])
};

struct RmacParams {
  RmacMap maps[CVPCE_POOL_MAX_MAPS];
  int4 regions[kRmacMaxRegionsTotal]; // (y0, x0, y1, x1)
  uchar4 segs[kRmacMaxRegionsTotal];  // (first row segment, end row segment, first column segment, end column segment)
  int n_maps, B;
  int per_region_l2, final_l2;
  float eps;
  float* out;
  long long out_stride;
  int nstages;
  int off_cm, off_seg, off_vals, off_desc, off_wscr, off_bar;  // byte offsets into dynamic shared memory (the ring is at 0)
  int wscr_floats;                                    // per consumer warp (NCHW maps: cell maxima)
};

__device__ __forceinline__ uint32_t rm_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void rm_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(rm_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void rm_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rm_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void rm_mbar_arrive_n(uint64_t* bar, uint32_t n) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(rm_smem_u32(bar)), "r"(n) : "memory");
}
__device__ __forceinline__ bool rm_mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(rm_smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void rm_mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!rm_mbar_try_wait(bar, parity)) __nanosleep(20);
}
// the producer is ahead by the depth of the ring: it can afford to sleep between polls
__device__ __forceinline__ void rm_mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
  while (!rm_mbar_try_wait(bar, parity)) __nanosleep(100);
}
// global -> shared bulk copy by the TMA engine; completion is counted in bytes on `bar`.  The maps are read once: evict-first.
__device__ __forceinline__ void rm_bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(rm_smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(rm_smem_u32(bar)), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void rm_cons_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kRmacCons) : "memory"); }

// Position in the ring: stage `st` (0 .. nstages-1, any count) and the running chunk counter `it`; the chunk's "full" barrier
// is full[it % 16] in phase it / 16.
struct RmacRing {
  uint32_t it;
  int st;
  __device__ __forceinline__ void advance(int nst) {
    ++it;
    st = (st + 1 == nst) ? 0 : st + 1;
  }
  __device__ __forceinline__ int bar() const { return (int)(it & (kRmacFullBars - 1)); }
  __device__ __forceinline__ uint32_t parity() const { return (it >> 4) & 1u; }
};

// element-wise maximum of two 16-byte vectors of the map's element type
template <int DTYPE>
__device__ __forceinline__ uint4 rm_vmax(const uint4 a, const uint4 b) {
  uint4 r;
  if (DTYPE == CVPCE_F32) {
    r.x = __float_as_uint(fmaxf(__uint_as_float(a.x), __uint_as_float(b.x)));
    r.y = __float_as_uint(fmaxf(__uint_as_float(a.y), __uint_as_float(b.y)));
    r.z = __float_as_uint(fmaxf(__uint_as_float(a.z), __uint_as_float(b.z)));
    r.w = __float_as_uint(fmaxf(__uint_as_float(a.w), __uint_as_float(b.w)));
  } else {
    __nv_bfloat162 t;
    t = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a.x), *reinterpret_cast<const __nv_bfloat162*>(&b.x)); r.x = *reinterpret_cast<uint32_t*>(&t);
    t = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a.y), *reinterpret_cast<const __nv_bfloat162*>(&b.y)); r.y = *reinterpret_cast<uint32_t*>(&t);
    t = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a.z), *reinterpret_cast<const __nv_bfloat162*>(&b.z)); r.z = *reinterpret_cast<uint32_t*>(&t);
    t = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a.w), *reinterpret_cast<const __nv_bfloat162*>(&b.w)); r.w = *reinterpret_cast<uint32_t*>(&t);
  }
  return r;
}
template <int DTYPE>
__device__ __forceinline__ uint4 rm_vneg_inf() {
  const uint32_t w = (DTYPE == CVPCE_F32) ? 0xff800000u : 0xff80ff80u;
  return make_uint4(w, w, w, w);
}

// warp-specialised form (rmac_staged_ws_kernel): 16 streaming warps, 8 region warps, the producer warp
constexpr int kWsStream = kRmacCons;           // streaming threads (= the consumers of the plain kernel)
constexpr int kWsReduceWarps = 8;
constexpr int kWsReduce = kWsReduceWarps * 32;  // region / norm threads
constexpr int kWsThreads = kWsStream + kWsReduce + 32;
constexpr int kWsBarReady = 2, kWsBarFree = 3, kWsBarReduce = 4;  // named barriers: cm ready (768), cm free (768), region team (256)

template <int NT, int BAR>
__device__ __forceinline__ void rm_team_sync() { asm volatile("bar.sync %0, %1;" ::"n"(BAR), "n"(NT) : "memory"); }

// ---------------------------------------------------------------------------------------------------------------------
// The region stage of one row segment by a team of NT threads (t = 0 .. NT-1, named barrier BAR): cm = the row of column
// maxima of row segment `ys`, complete and visible to the team.  It is shared-memory-bandwidth bound (cycle counters,
// profiles/r4n): letting every region re-read its columns of cm moved ~320 KB per row-segment end at the VGG shapes.  Two
// steps instead: the maximum of every COLUMN SEGMENT once (each column of cm is read once), then a region = the maximum of
// its few column segments.  Thread (g, q) = (group, 16-byte channel vector).  With RELEASE_CM the team hands cm back to the
// streaming warps as soon as the first step is done.
// ---------------------------------------------------------------------------------------------------------------------
template <int DTYPE, int NT, int BAR, bool RELEASE_CM>
__device__ __forceinline__ void rmac_region_stage(const RmacParams& p, const RmacMap& mp, const uint4* cm, uint4* segm, float* vals, int ys, int t,
                                                  bool release) {
  constexpr int E = (DTYPE == CVPCE_F32) ? 4 : 8;  // channels per 16-byte vector
  const int C = mp.C, nq = mp.pix_vecs, nxs = mp.nxs, n_regions = mp.n_regions, region_begin = mp.region_begin;
  const int n_groups = NT / nq;
  const int g = t / nq, q = t - g * nq;
  if (g < n_groups) {
    for (int sgi = g; sgi < nxs; sgi += n_groups) {
      int x = mp.xb[sgi];
      const int xe = mp.xb[sgi + 1];
      const uint4* src = cm + x * nq + q;
      uint4 m0 = rm_vneg_inf<DTYPE>(), m1 = m0;
      for (; x + 2 <= xe; x += 2, src += 2 * nq) {
        m0 = rm_vmax<DTYPE>(m0, src[0]);
        m1 = rm_vmax<DTYPE>(m1, src[nq]);
      }
      if (x < xe) m0 = rm_vmax<DTYPE>(m0, src[0]);
      segm[sgi * nq + q] = rm_vmax<DTYPE>(m0, m1);
    }
  }
  rm_team_sync<NT, BAR>();
  if (RELEASE_CM && release) {
    __threadfence_block();
    asm volatile("bar.arrive %0, %1;" ::"n"(kWsBarFree), "n"(kWsStream + kWsReduce) : "memory");
  }
  if (g < n_groups) {
    for (int r = g; r < n_regions; r += n_groups) {
      const uchar4 sg = p.segs[region_begin + r];
      if (ys < sg.x || ys >= sg.y) continue;
      const uint4* src = segm + sg.z * nq + q;
      uint4 m0 = src[0];
      for (int sgi = sg.z + 1; sgi < sg.w; ++sgi) {
        src += nq;
        m0 = rm_vmax<DTYPE>(m0, src[0]);
      }
      float f[E];
      if (DTYPE == CVPCE_F32) {
        f[0] = __uint_as_float(m0.x); f[1] = __uint_as_float(m0.y); f[2] = __uint_as_float(m0.z); f[3] = __uint_as_float(m0.w);
      } else {
        f[0] = __uint_as_float(m0.x << 16); f[1] = __uint_as_float(m0.x & 0xffff0000u);
        f[2] = __uint_as_float(m0.y << 16); f[3] = __uint_as_float(m0.y & 0xffff0000u);
        f[4 % E] = __uint_as_float(m0.z << 16); f[5 % E] = __uint_as_float(m0.z & 0xffff0000u);
        f[6 % E] = __uint_as_float(m0.w << 16); f[7 % E] = __uint_as_float(m0.w & 0xffff0000u);
      }
      float* dst = vals + r * C + q * E;
#pragma unroll
      for (int e = 0; e < E; e += 4) {
        float4 o = make_float4(f[e], f[e + 1], f[e + 2], f[e + 3]);
        if (ys > sg.x) {
          const float4 old = *reinterpret_cast<const float4*>(dst + e);
          o.x = fmaxf(o.x, old.x); o.y = fmaxf(o.y, old.y); o.z = fmaxf(o.z, old.z); o.w = fmaxf(o.w, old.w);
        }
        *reinterpret_cast<float4*>(dst + e) = o;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// channels-last map: region maxima -> vals[r][C].  P = 16 KB parts per row (1: a chunk holds whole rows).
// ---------------------------------------------------------------------------------------------------------------------
template <int DTYPE, int P, bool WS>  // WS: the caller is a streaming warp of rmac_staged_ws_kernel (the region stage runs elsewhere)
__device__ __forceinline__ void rmac_consume_nhwc(const RmacParams& p, const RmacMap& mp, const unsigned char* ring, uint64_t* full,
                                                  uint64_t* empty, uint4* cm, uint4* segm, float* vals, RmacRing& rg, int tid, int lane) {
  const int nst = p.nstages;
  const int H = mp.H, row_bytes = mp.row_bytes, R = mp.rows_per_chunk;
  const int RV = row_bytes >> 4;  // vectors per row
  constexpr int J = kRmacVecsPerThread;
  uint4 acc[P][J];
#pragma unroll
  for (int pc = 0; pc < P; ++pc)
#pragma unroll
    for (int j = 0; j < J; ++j) acc[pc][j] = rm_vneg_inf<DTYPE>();
  int y = 0, ys = 0, ynext = mp.yb[1];

  // the row segment ends: its row of column maxima goes to shared memory, every region that contains it is updated
  auto row_segment_end = [&]() {
    const long long tr0 = RM_CLOCK();
    if (WS) asm volatile("bar.sync %0, %1;" ::"n"(kWsBarFree), "n"(kWsStream + kWsReduce) : "memory");  // the region warps are done with cm
    else rm_cons_sync();  // the previous row segment's regions are done with cm
#pragma unroll
    for (int pc = 0; pc < P; ++pc) {
#pragma unroll
      for (int j = 0; j < J; ++j) {
        const int v = pc * kRmacStageVecs + tid + j * kRmacCons;
        if (v < RV) cm[v] = acc[pc][j];
        acc[pc][j] = rm_vneg_inf<DTYPE>();
      }
    }
    if (WS) {  // hand the row of column maxima to the region warps and go on streaming
      __threadfence_block();
      asm volatile("bar.arrive %0, %1;" ::"n"(kWsBarReady), "n"(kWsStream + kWsReduce) : "memory");
    } else {
      rm_cons_sync();
      rmac_region_stage<DTYPE, kRmacCons, 1, false>(p, mp, cm, segm, vals, ys, tid, false);
    }
    ++ys;
    ynext = mp.yb[ys + 1];
    if (tid == 0) RM_PROF_ADD(3, RM_CLOCK() - tr0);
  };

  if (P == 1) {
    while (y < H) {
      const long long tw0 = RM_CLOCK();
      rm_mbar_wait(&full[rg.bar()], rg.parity());
      if (tid == 0) RM_PROF_ADD(2, RM_CLOCK() - tw0);
      const uint4* buf = reinterpret_cast<const uint4*>(ring + (size_t)rg.st * kRmacStageBytes);
      const int nrows = min(R, H - y);
      bool released = false;
      for (int rr = 0; rr < nrows; ++rr, buf += RV) {
#pragma unroll
        for (int j = 0; j < J; ++j) {
          const int v = tid + j * kRmacCons;
          if (v < RV) acc[0][j] = rm_vmax<DTYPE>(acc[0][j], buf[v]);
        }
        ++y;
        if (y == ynext) {
          if (rr + 1 == nrows) {  // the stage is not needed any more: hand it back before the region stage
            __syncwarp();
            if (lane == 0) rm_mbar_arrive_n(&empty[rg.st], 1);
            released = true;
          }
          row_segment_end();
        }
      }
      if (!released) {
        __syncwarp();
        if (lane == 0) rm_mbar_arrive_n(&empty[rg.st], 1);
      }
      rg.advance(nst);
    }
  } else {
    while (y < H) {
#pragma unroll
      for (int pc = 0; pc < P; ++pc) {
        const long long tw0 = RM_CLOCK();
        rm_mbar_wait(&full[rg.bar()], rg.parity());
        if (tid == 0) RM_PROF_ADD(2, RM_CLOCK() - tw0);
        if (tid == 0 && rg.it == 0) RM_PROF_ADD(15, RM_CLOCK() - tw0);  // the wait for the very first chunk
        const uint4* buf = reinterpret_cast<const uint4*>(ring + (size_t)rg.st * kRmacStageBytes);
        const int nvec = min(kRmacStageVecs, RV - pc * kRmacStageVecs);
#pragma unroll
        for (int j = 0; j < J; ++j) {
          const int v = tid + j * kRmacCons;
          if (v < nvec) acc[pc][j] = rm_vmax<DTYPE>(acc[pc][j], buf[v]);
        }
        __syncwarp();
        if (lane == 0) rm_mbar_arrive_n(&empty[rg.st], 1);
        rg.advance(nst);
      }
      ++y;
      if (y == ynext) row_segment_end();
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// NCHW map: a warp pass handles `ppp` staged planes; a lane owns one 16-byte strip of columns of one plane (lane = (plane,
// strip)) and walks down the rows: one LDS.128 and 4 (f32) / 8 (bf16) FMNMX per row, the strip's maxima of a row segment are
// written back IN PLACE over row `segment index` of the plane (a lane only ever touches its own strip, and that row has
// been read by then) -- no scratch for the column maxima.  Then one lane per (plane, cell) reduces the cell's columns and
// one lane per (plane, region) the region's cells.
// A chunk holds up to `gpc` plane groups; its work is dealt to `nslots` = min(gpc, warps) warps, rotating from chunk to
// chunk, and ONLY those warps touch the chunk's barriers (they share out the arrivals the stage's "empty" barrier expects).
// A warp therefore skips chunks; that is safe with one-bit phase parities because the "full" barrier of chunk k is
// full[k % 16]: a warp's slot pattern repeats every 16 chunks, so it meets consecutive phases of the barriers it uses, and the
// consumer barrier between maps puts all warps behind every chunk of the previous map.
// ---------------------------------------------------------------------------------------------------------------------
template <int DTYPE>
__device__ __forceinline__ float rm_ld_elem(const unsigned char* base, int idx) {
  if (DTYPE == CVPCE_F32) return reinterpret_cast<const float*>(base)[idx];
  return __uint_as_float((uint32_t)reinterpret_cast<const unsigned short*>(base)[idx] << 16);
}

template <int DTYPE>
__device__ __forceinline__ void rmac_consume_nchw(const RmacParams& p, const RmacMap& mp, unsigned char* ring, uint64_t* full,
                                                  uint64_t* empty, float* cells, float* vals, RmacRing& rg, int warp, int lane) {
  constexpr int VW = (DTYPE == CVPCE_F32) ? 4 : 8;  // elements per 16 bytes
  constexpr int ESZ = (DTYPE == CVPCE_F32) ? 4 : 2;
  const int nst = p.nstages;
  const int C = mp.C, H = mp.H, W = mp.W, nxs = mp.nxs, nys = mp.nys, upc = mp.upc, lpr = mp.lpr, ppp = mp.ppp;
  const int n_chunks = mp.n_chunks, n_regions = mp.n_regions, region_begin = mp.region_begin, plane_bytes = mp.plane_bytes;
  const int row_bytes = W * ESZ;
  const int pp = lane / lpr, xv = lane - pp * lpr;
  const int ncell = nys * nxs;
  const int gpc = (upc + ppp - 1) / ppp;
  const int nslots = min(gpc, kRmacConsWarps);
  for (int k = 0; k < n_chunks; ++k, rg.advance(nst)) {
    const int slot = (warp - (int)((rg.it * (uint32_t)nslots) & (kRmacConsWarps - 1)) + kRmacConsWarps) & (kRmacConsWarps - 1);
    if (slot >= nslots) continue;
    const uint32_t arrivals = kRmacConsWarps / nslots + (slot < kRmacConsWarps % nslots ? 1 : 0);
    const long long tw0 = RM_CLOCK();
    rm_mbar_wait(&full[rg.bar()], rg.parity());
    if (warp == 0 && lane == 0) RM_PROF_ADD(2, RM_CLOCK() - tw0);
    const int u0 = k * upc;
    const int nu = min(upc, C - u0);
    const int ngroups = (nu + ppp - 1) / ppp;
    unsigned char* buf = ring + (size_t)rg.st * kRmacStageBytes;
    bool wrote = false;
    for (int g = slot; g < ngroups; g += nslots) {
      wrote = true;
      const int pl = g * ppp + pp;
      const bool live = pp < ppp && pl < nu;
      {
        unsigned char* strip = buf + (size_t)pl * plane_bytes + xv * 16;
        float acc[VW];
#pragma unroll
        for (int e = 0; e < VW; ++e) acc[e] = -INFINITY;
        int ysr = 0, next = mp.yb[1];
#pragma unroll 4
        for (int y = 0; y < H; ++y) {
          if (live) {
            const uint4 w = *reinterpret_cast<const uint4*>(strip + y * row_bytes);
            if (DTYPE == CVPCE_F32) {
              acc[0] = fmaxf(acc[0], __uint_as_float(w.x)); acc[1] = fmaxf(acc[1], __uint_as_float(w.y));
              acc[2] = fmaxf(acc[2], __uint_as_float(w.z)); acc[3] = fmaxf(acc[3], __uint_as_float(w.w));
            } else {
              acc[0] = fmaxf(acc[0], __uint_as_float(w.x << 16)); acc[1] = fmaxf(acc[1], __uint_as_float(w.x & 0xffff0000u));
              acc[2] = fmaxf(acc[2], __uint_as_float(w.y << 16)); acc[3] = fmaxf(acc[3], __uint_as_float(w.y & 0xffff0000u));
              acc[4 % VW] = fmaxf(acc[4 % VW], __uint_as_float(w.z << 16)); acc[5 % VW] = fmaxf(acc[5 % VW], __uint_as_float(w.z & 0xffff0000u));
              acc[6 % VW] = fmaxf(acc[6 % VW], __uint_as_float(w.w << 16)); acc[7 % VW] = fmaxf(acc[7 % VW], __uint_as_float(w.w & 0xffff0000u));
            }
          }
          if (y + 1 == next) {  // warp-uniform: the row segment ends; its column maxima replace row `ysr` of the strip
            if (live) {
              uint4 o;
              if (DTYPE == CVPCE_F32) {
                o = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
              } else {  // maxima of bf16 values are bf16 values: the upper halves are exact
                o.x = (__float_as_uint(acc[0]) >> 16) | (__float_as_uint(acc[1]) & 0xffff0000u);
                o.y = (__float_as_uint(acc[2]) >> 16) | (__float_as_uint(acc[3]) & 0xffff0000u);
                o.z = (__float_as_uint(acc[4 % VW]) >> 16) | (__float_as_uint(acc[5 % VW]) & 0xffff0000u);
                o.w = (__float_as_uint(acc[6 % VW]) >> 16) | (__float_as_uint(acc[7 % VW]) & 0xffff0000u);
              }
              *reinterpret_cast<uint4*>(strip + ysr * row_bytes) = o;
            }
#pragma unroll
            for (int e = 0; e < VW; ++e) acc[e] = -INFINITY;
            ++ysr;
            next = mp.yb[ysr + 1];
          }
        }
      }
      __syncwarp();
      for (int c = lane; c < ncell * ppp; c += 32) {
        const int ps = c / ncell, cc = c - ps * ncell;
        const int plr = g * ppp + ps;
        if (plr >= nu) continue;
        const int ysr = cc / nxs, xsr = cc - ysr * nxs;
        const unsigned char* row = buf + (size_t)plr * plane_bytes + ysr * row_bytes;
        float v0 = -INFINITY, v1 = -INFINITY;
        int xx = mp.xb[xsr];
        const int xe = mp.xb[xsr + 1];
        for (; xx + 1 < xe; xx += 2) {
          v0 = fmaxf(v0, rm_ld_elem<DTYPE>(row, xx));
          v1 = fmaxf(v1, rm_ld_elem<DTYPE>(row, xx + 1));
        }
        if (xx < xe) v0 = fmaxf(v0, rm_ld_elem<DTYPE>(row, xx));
        cells[c] = fmaxf(v0, v1);
      }
      __syncwarp();
      for (int q = lane; q < n_regions * ppp; q += 32) {
        const int ps = q / n_regions, r = q - ps * n_regions;
        const int plr = g * ppp + ps;
        if (plr >= nu) continue;
        const uchar4 sg = p.segs[region_begin + r];
        float v0 = -INFINITY, v1 = -INFINITY;
        for (int ysr = sg.x; ysr < sg.y; ++ysr) {
          const float* crow = cells + ps * ncell + ysr * nxs;
          int xsr = sg.z;
          for (; xsr + 1 < sg.w; xsr += 2) {
            v0 = fmaxf(v0, crow[xsr]);
            v1 = fmaxf(v1, crow[xsr + 1]);
          }
          if (xsr < sg.w) v0 = fmaxf(v0, crow[xsr]);
        }
        vals[r * C + u0 + plr] = fmaxf(v0, v1);
      }
      __syncwarp();  // the cell scratch is reused by this warp's next pass
    }
    // the stage was written with ordinary stores and will next be written by the TMA engine (async proxy)
    if (wrote) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (lane == 0) rm_mbar_arrive_n(&empty[rg.st], arrivals);
  }
}

template <int DTYPE>
__global__ void __launch_bounds__(kRmacThreads, 1) rmac_staged_kernel(const __grid_constant__ RmacParams gp) {
  extern __shared__ __align__(128) unsigned char rm_smem[];
  // control data is read out of shared memory (the parameter block is 3 KB: indexed loads through the constant cache are slow)
  __shared__ __align__(16) RmacParams p;
  __shared__ float s_part[kRmacMaxRegions][kRmacConsWarps];
  __shared__ float s_denom[kRmacMaxRegions];
  __shared__ float s_wsum[kRmacConsWarps];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < (int)(sizeof(RmacParams) / 4); i += kRmacThreads) reinterpret_cast<int*>(&p)[i] = reinterpret_cast<const int*>(&gp)[i];
  unsigned char* ring = rm_smem;
  uint4* cm = reinterpret_cast<uint4*>(rm_smem + gp.off_cm);
  uint4* segm = reinterpret_cast<uint4*>(rm_smem + gp.off_seg);
  float* vals = reinterpret_cast<float*>(rm_smem + gp.off_vals);
  float* desc = reinterpret_cast<float*>(rm_smem + gp.off_desc);
  float* wscr_all = reinterpret_cast<float*>(rm_smem + gp.off_wscr);
  uint64_t* full = reinterpret_cast<uint64_t*>(rm_smem + gp.off_bar);
  uint64_t* empty = full + kRmacFullBars;
  const int nst = gp.nstages;
  if (tid == 0) {
    for (int s = 0; s < kRmacFullBars; ++s) rm_mbar_init(&full[s], 1);
    for (int s = 0; s < nst; ++s) rm_mbar_init(&empty[s], kRmacConsWarps);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kRmacConsWarps) {
    // ---- producer: one thread keeps the ring full, across maps and images -----------------------------------------------
    if (lane == 0) {
      uint64_t policy;
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
      RmacRing rg = {0u, 0};
      uint32_t rev = 0;  // revolutions of the ring (parity of the stage's "empty" phase)
      const long long tp0 = RM_CLOCK();
      for (int n = blockIdx.x; n < p.B; n += gridDim.x) {
        for (int m = 0; m < p.n_maps; ++m) {
          const RmacMap& mp = p.maps[m];
          const unsigned char* src = reinterpret_cast<const unsigned char*>(mp.data) + (long long)n * mp.image_bytes;
          const int n_chunks = mp.n_chunks;
          const int H = mp.H, C = mp.C, P = mp.row_parts, R = mp.rows_per_chunk, upc = mp.upc, row_bytes = mp.row_bytes, plane_bytes = mp.plane_bytes;
          const bool nhwc = mp.nhwc != 0;
          int row = 0, part = 0;
          for (int k = 0; k < n_chunks; ++k) {
            long long off;
            uint32_t bytes;
            if (!nhwc) {  // whole planes
              off = (long long)k * upc * plane_bytes;
              bytes = (uint32_t)min(upc, C - k * upc) * (uint32_t)plane_bytes;
            } else if (P == 1) {  // whole rows
              off = (long long)row * row_bytes;
              const int nr = min(R, H - row);
              bytes = (uint32_t)nr * (uint32_t)row_bytes;
              row += nr;
            } else {  // one 16 KB part of a row
              off = (long long)row * row_bytes + (long long)part * kRmacStageBytes;
              bytes = (uint32_t)min(kRmacStageBytes, row_bytes - part * kRmacStageBytes);
              if (++part == P) {
                part = 0;
                ++row;
              }
            }
            const long long t0 = RM_CLOCK();
            rm_mbar_wait_relaxed(&empty[rg.st], (rev & 1) ^ 1);
            RM_PROF_ADD(0, RM_CLOCK() - t0);
            rm_mbar_expect_tx(&full[rg.bar()], bytes);
            rm_bulk_load(ring + (size_t)rg.st * kRmacStageBytes, src + off, bytes, &full[rg.bar()], policy);
            rg.advance(nst);
            if (rg.st == 0) ++rev;
          }
        }
      }
      RM_PROF_ADD(1, RM_CLOCK() - tp0);
    }
    return;  // the consumers wait for every copy that was issued, so nothing is in flight when the CTA ends
  }

  // ---- consumers -----------------------------------------------------------------------------------------------------------
  RmacRing rg = {0u, 0};
  const long long tc0 = RM_CLOCK();
  for (int n = blockIdx.x; n < p.B; n += gridDim.x) {
    float sq = 0.f;
    for (int m = 0; m < p.n_maps; ++m) {
      const RmacMap& mp = p.maps[m];
      if (mp.nhwc) {
        if (mp.row_parts == 1) rmac_consume_nhwc<DTYPE, 1, false>(p, mp, ring, full, empty, cm, segm, vals, rg, tid, lane);
        else if (mp.row_parts == 2) rmac_consume_nhwc<DTYPE, 2, false>(p, mp, ring, full, empty, cm, segm, vals, rg, tid, lane);
        else if (mp.row_parts == 3) rmac_consume_nhwc<DTYPE, 3, false>(p, mp, ring, full, empty, cm, segm, vals, rg, tid, lane);
        else rmac_consume_nhwc<DTYPE, 4, false>(p, mp, ring, full, empty, cm, segm, vals, rg, tid, lane);
      } else {
        rmac_consume_nchw<DTYPE>(p, mp, ring, full, empty, wscr_all + warp * p.wscr_floats, vals, rg, warp, lane);
      }
      const long long ts0 = RM_CLOCK();
      rm_cons_sync();  // vals[r][C] of this map is complete, every chunk of the map has been consumed
      if (tid == 0) RM_PROF_ADD(6, RM_CLOCK() - ts0);
      const long long tf0 = RM_CLOCK();
      const int C = mp.C, n_regions = mp.n_regions;
      if (p.per_region_l2) {
        for (int r = 0; r < n_regions; ++r) {
          float s = 0.f;
          for (int c = tid; c < C; c += kRmacCons) {
            const float v = vals[r * C + c];
            s = fmaf(v, v, s);
          }
          s = warp_sum(s);
          if (lane == 0) s_part[r][warp] = s;
        }
        rm_cons_sync();
        if (tid < n_regions) {
          float t = 0.f;
#pragma unroll
          for (int w = 0; w < kRmacConsWarps; ++w) t += s_part[tid][w];
          s_denom[tid] = fmaxf(sqrtf(t), p.eps);
        }
        rm_cons_sync();
      }
      for (int c = tid; c < C; c += kRmacCons) {
        float d = 0.f;
        for (int r = 0; r < n_regions; ++r) {
          float v = vals[r * C + c];
          if (p.per_region_l2) v = v / s_denom[r];
          d = (r == 0) ? v : d + v;
        }
        desc[mp.out_offset + c] = d;
        sq = fmaf(d, d, sq);
      }
      rm_cons_sync();  // vals, cm and s_denom are reused by the next map
      if (tid == 0) RM_PROF_ADD(4, RM_CLOCK() - tf0);
    }
    const long long tg0 = RM_CLOCK();
    float denom = 1.f;
    if (p.final_l2) {
      sq = warp_sum(sq);
      if (lane == 0) s_wsum[warp] = sq;
      rm_cons_sync();
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < kRmacConsWarps; ++w) t += s_wsum[w];
      denom = fmaxf(sqrtf(t), p.eps);
    }
    const RmacMap& last = p.maps[p.n_maps - 1];
    const int D = last.out_offset + last.C;
    float* orow = p.out + (long long)n * p.out_stride;
    for (int c = tid; c < D; c += kRmacCons) orow[c] = p.final_l2 ? desc[c] / denom : desc[c];
    rm_cons_sync();  // desc / s_wsum are reused by the next image
    if (tid == 0) RM_PROF_ADD(7, RM_CLOCK() - tg0);
  }
  if (tid == 0) RM_PROF_ADD(5, RM_CLOCK() - tc0);
}

// ---------------------------------------------------------------------------------------------------------------------
// Warp-specialised form for channels-last maps: rmac_staged_ws_kernel.  In rmac_staged_kernel the 16 consumer warps stop
// streaming 12 times per image for the region stage and twice for the norms (~25 us of an ~80 us image, during which only the
// ring's 112 KB keep the loads going).  Here the 16 STREAMING warps never stop: at the end of a row segment they drop their row
// of column maxima into cm and go on with the next rows; 8 REGION warps (256 threads) pick cm up (named barriers: "ready" /
// "free", bar.arrive on the giving side, bar.sync on the taking side), reduce it to column-segment maxima -- cm is free again
// after that first step --, update the regions, and do the per-map and per-image norms and the store while the next image is
// already streaming.  vals, segm, desc belong to the region warps alone.
// ---------------------------------------------------------------------------------------------------------------------
template <int DTYPE>
__device__ __forceinline__ void rmac_ws_reducer(const RmacParams& p, const uint4* cm, uint4* segm, float* vals, float* desc,
                                                float (*s_part)[kWsReduceWarps], float* s_denom, float* s_wsum, int t) {
  const int lane = t & 31, warp = t >> 5;
  asm volatile("bar.arrive %0, %1;" ::"n"(kWsBarFree), "n"(kWsStream + kWsReduce) : "memory");  // cm starts out free
  for (int n = blockIdx.x; n < p.B; n += gridDim.x) {
    const bool last_image = n + (int)gridDim.x >= p.B;
    float sq = 0.f;
    for (int m = 0; m < p.n_maps; ++m) {
      const RmacMap& mp = p.maps[m];
      const int nys = mp.nys;
      for (int ys = 0; ys < nys; ++ys) {
        asm volatile("bar.sync %0, %1;" ::"n"(kWsBarReady), "n"(kWsStream + kWsReduce) : "memory");  // the row of column maxima is in cm
        // the very last hand-back has no taker: it is not given
        const bool release = !(last_image && m == p.n_maps - 1 && ys == nys - 1);
        rmac_region_stage<DTYPE, kWsReduce, kWsBarReduce, true>(p, mp, cm, segm, vals, ys, t, release);
        rm_team_sync<kWsReduce, kWsBarReduce>();  // segm is rewritten by the next row segment; vals complete after the last one
      }
      const int C = mp.C, n_regions = mp.n_regions;
      if (p.per_region_l2) {
        // this is on the critical path of the hand-off (and the tail of the launch): eight regions at a time, so that their
        // shuffle reductions overlap instead of forming 14 chains of five dependent shuffles
        for (int r0 = 0; r0 < n_regions; r0 += 8) {
          float ssum[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            ssum[j] = 0.f;
            if (r0 + j < n_regions)
              for (int c = t; c < C; c += kWsReduce) {
                const float v = vals[(r0 + j) * C + c];
                ssum[j] = fmaf(v, v, ssum[j]);
              }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int j = 0; j < 8; ++j) ssum[j] += __shfl_xor_sync(0xffffffffu, ssum[j], o);
          }
          if (lane == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (r0 + j < n_regions) s_part[r0 + j][warp] = ssum[j];
          }
        }
        rm_team_sync<kWsReduce, kWsBarReduce>();
        if (t < n_regions) {
          float tot = 0.f;
#pragma unroll
          for (int w = 0; w < kWsReduceWarps; ++w) tot += s_part[t][w];
          s_denom[t] = fmaxf(sqrtf(tot), p.eps);
        }
        rm_team_sync<kWsReduce, kWsBarReduce>();
      }
      for (int c = t; c < C; c += kWsReduce) {
        float d = 0.f;
        int r = 0;
        for (; r + 4 <= n_regions; r += 4) {  // four independent quotients, summed in region order
          float v0 = vals[r * C + c], v1 = vals[(r + 1) * C + c], v2 = vals[(r + 2) * C + c], v3 = vals[(r + 3) * C + c];
          if (p.per_region_l2) {
            v0 = v0 / s_denom[r];
            v1 = v1 / s_denom[r + 1];
            v2 = v2 / s_denom[r + 2];
            v3 = v3 / s_denom[r + 3];
          }
          d = (r == 0) ? v0 : d + v0;
          d += v1;
          d += v2;
          d += v3;
        }
        for (; r < n_regions; ++r) {
          float v = vals[r * C + c];
          if (p.per_region_l2) v = v / s_denom[r];
          d = (r == 0) ? v : d + v;
        }
        desc[mp.out_offset + c] = d;
        sq = fmaf(d, d, sq);
      }
      rm_team_sync<kWsReduce, kWsBarReduce>();  // vals and s_denom are reused by the next map
    }
    float denom = 1.f;
    if (p.final_l2) {
      sq = warp_sum(sq);
      if (lane == 0) s_wsum[warp] = sq;
      rm_team_sync<kWsReduce, kWsBarReduce>();
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < kWsReduceWarps; ++w) tot += s_wsum[w];
      denom = fmaxf(sqrtf(tot), p.eps);
    }
    const RmacMap& last = p.maps[p.n_maps - 1];
    const int D = last.out_offset + last.C;
    float* orow = p.out + (long long)n * p.out_stride;
    for (int c = t; c < D; c += kWsReduce) orow[c] = p.final_l2 ? desc[c] / denom : desc[c];
    rm_team_sync<kWsReduce, kWsBarReduce>();  // desc / s_wsum are reused by the next image
  }
}

template <int DTYPE>
__global__ void __launch_bounds__(kWsThreads, 1) rmac_staged_ws_kernel(const __grid_constant__ RmacParams gp) {
  extern __shared__ __align__(128) unsigned char rm_smem[];
  __shared__ __align__(16) RmacParams p;
  __shared__ float s_part[kRmacMaxRegions][kWsReduceWarps];
  __shared__ float s_denom[kRmacMaxRegions];
  __shared__ float s_wsum[kWsReduceWarps];

  const long long t_entry = RM_CLOCK();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < (int)(sizeof(RmacParams) / 4); i += kWsThreads) reinterpret_cast<int*>(&p)[i] = reinterpret_cast<const int*>(&gp)[i];
  unsigned char* ring = rm_smem;
  uint4* cm = reinterpret_cast<uint4*>(rm_smem + gp.off_cm);
  uint4* segm = reinterpret_cast<uint4*>(rm_smem + gp.off_seg);
  float* vals = reinterpret_cast<float*>(rm_smem + gp.off_vals);
  float* desc = reinterpret_cast<float*>(rm_smem + gp.off_desc);
  uint64_t* full = reinterpret_cast<uint64_t*>(rm_smem + gp.off_bar);
  uint64_t* empty = full + kRmacFullBars;
  const int nst = gp.nstages;
  if (tid == 0) {
    for (int s = 0; s < kRmacFullBars; ++s) rm_mbar_init(&full[s], 1);
    for (int s = 0; s < nst; ++s) rm_mbar_init(&empty[s], kRmacConsWarps);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kRmacConsWarps + kWsReduceWarps) {
    // ---- producer (channels-last maps only) ---------------------------------------------------------------------------------
    if (lane == 0) {
      uint64_t policy;
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
      RmacRing rg = {0u, 0};
      uint32_t rev = 0;
      for (int n = blockIdx.x; n < p.B; n += gridDim.x) {
        for (int m = 0; m < p.n_maps; ++m) {
          const RmacMap& mp = p.maps[m];
          const unsigned char* src = reinterpret_cast<const unsigned char*>(mp.data) + (long long)n * mp.image_bytes;
          const int n_chunks = mp.n_chunks, H = mp.H, P = mp.row_parts, R = mp.rows_per_chunk, row_bytes = mp.row_bytes;
          int row = 0, part = 0;
          for (int k = 0; k < n_chunks; ++k) {
            long long off;
            uint32_t bytes;
            if (P == 1) {  // whole rows
              off = (long long)row * row_bytes;
              const int nr = min(R, H - row);
              bytes = (uint32_t)nr * (uint32_t)row_bytes;
              row += nr;
            } else {  // one part of a row
              off = (long long)row * row_bytes + (long long)part * kRmacStageBytes;
              bytes = (uint32_t)min(kRmacStageBytes, row_bytes - part * kRmacStageBytes);
              if (++part == P) {
                part = 0;
                ++row;
              }
            }
            rm_mbar_wait_relaxed(&empty[rg.st], (rev & 1) ^ 1);
            rm_mbar_expect_tx(&full[rg.bar()], bytes);
            rm_bulk_load(ring + (size_t)rg.st * kRmacStageBytes, src + off, bytes, &full[rg.bar()], policy);
            rg.advance(nst);
            if (rg.st == 0) ++rev;
          }
        }
      }
    }
    return;
  }
  if (warp >= kRmacConsWarps) {
    // ---- region warps ----------------------------------------------------------------------------------------------------------
    rmac_ws_reducer<DTYPE>(p, cm, segm, vals, desc, s_part, s_denom, s_wsum, tid - kWsStream);
    if (tid == kWsStream) RM_PROF_ADD(14, RM_CLOCK() - t_entry);  // entry -> the region warps are done (the CTA's work is complete)
    return;
  }
  // ---- streaming warps ---------------------------------------------------------------------------------------------------------
  RmacRing rg = {0u, 0};
  if (tid == 0) RM_PROF_ADD(12, RM_CLOCK() - t_entry);  // prologue (parameter copy, barrier init, CTA barrier)
  for (int n = blockIdx.x; n < p.B; n += gridDim.x) {
    for (int m = 0; m < p.n_maps; ++m) {
      const RmacMap& mp = p.maps[m];
      if (mp.row_parts == 1) rmac_consume_nhwc<DTYPE, 1, true>(p, mp, ring, full, empty, cm, segm, vals, rg, tid, lane);
      else if (mp.row_parts == 2) rmac_consume_nhwc<DTYPE, 2, true>(p, mp, ring, full, empty, cm, segm, vals, rg, tid, lane);
      else if (mp.row_parts == 3) rmac_consume_nhwc<DTYPE, 3, true>(p, mp, ring, full, empty, cm, segm, vals, rg, tid, lane);
      else rmac_consume_nhwc<DTYPE, 4, true>(p, mp, ring, full, empty, cm, segm, vals, rg, tid, lane);
    }
  }
  if (tid == 0) RM_PROF_ADD(13, RM_CLOCK() - t_entry);  // entry -> the streaming warps are done
}

// Elementary segments of [0, dim) between the distinct boundaries of the intervals; returns the count (0: too many).
static int rmac_segments(const int* lo, const int* hi, int n, int dim, unsigned char* b) {
  int pts[2 * CVPCE_POOL_MAX_REGIONS + 2];
  int np = 0;
  pts[np++] = 0;
  pts[np++] = dim;
  for (int i = 0; i < n; ++i) {
    pts[np++] = lo[i];
    pts[np++] = hi[i];
  }
  std::sort(pts, pts + np);
  np = (int)(std::unique(pts, pts + np) - pts);
  if (np - 1 > kRmacMaxSegs || dim > 255) return 0;
  for (int i = 0; i < np; ++i) b[i] = (unsigned char)pts[i];
  return np - 1;
}
static int rmac_seg_index(const unsigned char* b, int ns, int v) {
  for (int i = 0; i <= ns; ++i)
    if (b[i] == v) return i;
  return 0;
}

// Launches the staged kernel when it covers the call.  Returns CVPCE_OK (launched), kRmacNotApplicable (the caller falls
// back to the cluster kernel) or an error code.  Arguments are already validated by region_pool_impl.
int rmac_staged_launch(const cvpce_feature_map* maps, int n_maps, int dtype, int B, const int32_t* regions, const int32_t* n_regions,
                       int per_region_l2, int final_l2, float eps, float* out, int64_t out_stride, void* stream) {
  const int mode = tuning("rmac_kernel");  // 0: automatic; 1: cluster kernel; 2: rmac_staged_kernel whenever applicable;
                                           // 4: whenever applicable, the warp-specialised kernel where it covers the maps
  if (mode == 1 || !(dtype == CVPCE_F32 || dtype == CVPCE_BF16) || B <= 0) return kRmacNotApplicable;
  // One CTA per image: below about half the SMs the cluster kernel's 8 CTAs per image fill the GPU better.
  if (mode == 0 && B < sm_count() / 2) return kRmacNotApplicable;
  const int esz = dtype == CVPCE_F32 ? 4 : 2;
  RmacParams p;
  memset(&p, 0, sizeof(p));
  p.n_maps = n_maps;
  p.B = B;
  p.per_region_l2 = per_region_l2 ? 1 : 0;
  p.final_l2 = final_l2 ? 1 : 0;
  p.eps = eps;
  p.out = out;
  p.out_stride = out_stride;
  int rbeg = 0, col = 0;
  size_t cm_bytes = 0, seg_bytes = 0, vals_floats = 0, wscr_floats = 0;
  const int32_t* rg = regions;
  for (int m = 0; m < n_maps; ++m) {
    const cvpce_feature_map& fm = maps[m];
    RmacMap& pm = p.maps[m];
    pm.data = fm.data;
    pm.C = fm.C; pm.H = fm.H; pm.W = fm.W;
    pm.image_bytes = (long long)fm.stride_n * esz;
    if (!aligned16(fm.data) || pm.image_bytes % 16 != 0 || fm.H > 255 || fm.W > 255) return kRmacNotApplicable;
    const long long pix_bytes = (long long)fm.C * esz, plane_bytes = (long long)fm.H * fm.W * esz;
    const long long cl_row_bytes = pix_bytes * fm.W, pl_row_bytes = (long long)fm.W * esz;
    // channels-last: pixels of whole 16-byte vectors, at most one region-stage thread group's worth of them, rows of at most 4 chunks
    const bool dense_nhwc = fm.stride_c == 1 && fm.stride_w == fm.C && fm.stride_h == (int64_t)fm.W * fm.C && pix_bytes % 16 == 0 &&
                            pix_bytes / 16 <= kRmacCons && cl_row_bytes <= (long long)kRmacMaxRowParts * kRmacStageBytes;
    const bool dense_nchw = fm.stride_w == 1 && fm.stride_h == fm.W && fm.stride_c == (int64_t)fm.H * fm.W && pl_row_bytes % 16 == 0 &&
                            pl_row_bytes / 16 <= 32 && plane_bytes <= kRmacStageBytes;
    if (!dense_nhwc && !dense_nchw) return kRmacNotApplicable;
    pm.nhwc = dense_nhwc ? 1 : 0;
    pm.n_regions = n_regions ? n_regions[m] : 1;
    if (pm.n_regions < 1 || pm.n_regions > kRmacMaxRegions || rbeg + pm.n_regions > kRmacMaxRegionsTotal) return kRmacNotApplicable;
    pm.region_begin = rbeg;
    int x0[CVPCE_POOL_MAX_REGIONS], x1[CVPCE_POOL_MAX_REGIONS], y0[CVPCE_POOL_MAX_REGIONS], y1[CVPCE_POOL_MAX_REGIONS];
    for (int r = 0; r < pm.n_regions; ++r) {
      if (rg) {
        y0[r] = rg[0]; x0[r] = rg[1]; y1[r] = rg[2]; x1[r] = rg[3];
        rg += 4;
      } else {
        y0[r] = 0; x0[r] = 0; y1[r] = fm.H; x1[r] = fm.W;
      }
      p.regions[rbeg + r] = make_int4(y0[r], x0[r], y1[r], x1[r]);
    }
    pm.nxs = rmac_segments(x0, x1, pm.n_regions, fm.W, pm.xb);
    pm.nys = rmac_segments(y0, y1, pm.n_regions, fm.H, pm.yb);
    if (pm.nxs == 0 || pm.nys == 0) return kRmacNotApplicable;
    for (int r = 0; r < pm.n_regions; ++r)
      p.segs[rbeg + r] = make_uchar4((unsigned char)rmac_seg_index(pm.yb, pm.nys, y0[r]), (unsigned char)rmac_seg_index(pm.yb, pm.nys, y1[r]),
                                     (unsigned char)rmac_seg_index(pm.xb, pm.nxs, x0[r]), (unsigned char)rmac_seg_index(pm.xb, pm.nxs, x1[r]));
    pm.row_parts = 1;
    pm.rows_per_chunk = 1;
    if (pm.nhwc) {
      pm.row_bytes = (int)cl_row_bytes;
      pm.pix_vecs = (int)(pix_bytes / 16);
      if (pm.row_bytes <= kRmacStageBytes) {
        pm.rows_per_chunk = kRmacStageBytes / pm.row_bytes;
        pm.n_chunks = (fm.H + pm.rows_per_chunk - 1) / pm.rows_per_chunk;
      } else {
        pm.row_parts = (pm.row_bytes + kRmacStageBytes - 1) / kRmacStageBytes;
        pm.n_chunks = fm.H * pm.row_parts;
      }
      cm_bytes = std::max(cm_bytes, (size_t)pm.row_bytes);
      seg_bytes = std::max(seg_bytes, (size_t)pm.nxs * (size_t)pix_bytes);
    } else {
      pm.plane_bytes = (int)plane_bytes;
      pm.lpr = (int)(pl_row_bytes / 16);
      pm.ppp = 32 / pm.lpr;
      pm.upc = std::max(1, kRmacStageBytes / pm.plane_bytes);
      if (pm.upc > pm.ppp) pm.upc -= pm.upc % pm.ppp;
      pm.n_chunks = (fm.C + pm.upc - 1) / pm.upc;
      wscr_floats = std::max(wscr_floats, (size_t)pm.ppp * pm.nys * pm.nxs);
    }
    rbeg += pm.n_regions;
    pm.out_offset = col;
    col += fm.C;
    vals_floats = std::max(vals_floats, (size_t)pm.n_regions * fm.C);
  }
  auto up128 = [](size_t v) { return (v + 127) & ~(size_t)127; };
  wscr_floats = (wscr_floats + 3) & ~(size_t)3;
  const size_t bar_bytes = (kRmacFullBars + kRmacMaxStages) * sizeof(uint64_t);
  const size_t rest = up128(cm_bytes) + up128(seg_bytes) + up128(vals_floats * 4) + up128((size_t)col * 4) + up128(wscr_floats * 4 * kRmacConsWarps) + up128(bar_bytes);
  const size_t budget = 222 * 1024;  // static (4.2 KB) + dynamic shared memory of a CTA may reach 227 KB
  if (rest + 3 * (size_t)kRmacStageBytes > budget) return kRmacNotApplicable;
  int nst = (int)std::min<size_t>(kRmacMaxStages, (budget - rest) / kRmacStageBytes);
  const int forced = tuning("rmac_stages");
  if (forced >= 2 && forced < nst) nst = forced;
  p.nstages = nst;
  size_t off = (size_t)nst * kRmacStageBytes;
  p.off_cm = (int)off;    off += up128(cm_bytes);
  p.off_seg = (int)off;   off += up128(seg_bytes);
  p.off_vals = (int)off;  off += up128(vals_floats * 4);
  p.off_desc = (int)off;  off += up128((size_t)col * 4);
  p.off_wscr = (int)off;  off += up128(wscr_floats * 4 * kRmacConsWarps);
  p.off_bar = (int)off;   off += up128(bar_bytes);
  p.wscr_floats = (int)wscr_floats;
  // channels-last only, pixels of at most 256 vectors (one region-team thread per vector): the warp-specialised kernel
  bool ws_ok = true;
  for (int m = 0; m < n_maps; ++m) ws_ok = ws_ok && p.maps[m].nhwc && p.maps[m].pix_vecs <= kWsReduce;
  // measured (tools/kbench.py rmac, same box): it wins from about 1.2 MB per image (VGG relu4_3 + relu5_3: 0.1475 -> 0.1413 ms at
  // B = 256 f32, 0.098 -> 0.092 bf16) and loses a little below (two 16 x 16 x 512 f32 maps: 0.092 -> 0.096)
  long long bytes_per_image = 0;
  for (int m = 0; m < n_maps; ++m) bytes_per_image += (long long)maps[m].C * maps[m].H * maps[m].W * esz;
  const bool use_ws = ws_ok && (mode == 4 || (mode == 0 && tuning("rmac_ws") != 1 && bytes_per_image >= 1200 * 1024));
  auto kern = use_ws ? (dtype == CVPCE_F32 ? rmac_staged_ws_kernel<CVPCE_F32> : rmac_staged_ws_kernel<CVPCE_BF16>)
                     : (dtype == CVPCE_F32 ? rmac_staged_kernel<CVPCE_F32> : rmac_staged_kernel<CVPCE_BF16>);
  if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), off)) return rc;
  // equal image counts per CTA: a CTA that alone carries one image more than the others would double the run time
  const int per_cta = (B + sm_count() - 1) / sm_count();
  const int grid = (B + per_cta - 1) / per_cta;
  kern<<<grid, use_ws ? kWsThreads : kRmacThreads, off, (cudaStream_t)stream>>>(p);
  CVPCE_LAUNCH_OK("rmac_staged_kernel");
  return CVPCE_OK;
}

}  // namespace cvpce

#ifdef CVPCE_RMAC_PROF
extern "C" __attribute__((visibility("default"))) int cvpce_debug_rmac_prof(long long* out16, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out16, cvpce::g_rmac_prof, sizeof(long long) * 16);
  if (reset) {
    long long z[16] = {0};
    cudaMemcpyToSymbol(cvpce::g_rmac_prof, z, sizeof(z));
  }
  return 0;
}
#endif
