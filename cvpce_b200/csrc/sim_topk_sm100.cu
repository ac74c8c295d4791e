// Cosine similarity + fused per-row top-k on the sm_100a tensor cores (tcgen05 + TMEM + TMA).
//
// Replaces cvpce/models/classification.py:87-95 (distance / nearest_neighbors: gather two [Q,G,D] tensors, reduce, full
// argsort) with S = Q x G^T computed tile by tile and reduced on the fly; S never reaches HBM.
//
// Kernel 1  sim_topk_tc_kernel   (one CTA per SM, 256 threads, warp-specialised)
//   warp 0      TMA producer: 128B-swizzled K-major bf16 tiles of Q (128 x 64) and G (256 x 64) into a shared-memory ring
//               (for D <= 512 the whole 128 x D query block is loaded once and stays resident; only G streams)
//   warp 1      one elected thread issues tcgen05.mma (M=128, N=256, K=16, bf16 x bf16 -> f32) into TMEM; the 512 TMEM
//               columns hold TWO 128x256 accumulators so the MMAs of tile n+1 overlap the epilogue of tile n
//   warp 2      TMEM allocator
//   warps 4-7   epilogue: tcgen05.ld gives every thread one complete query row of the tile (lane == row), so each thread
//               keeps its row's running top-k' (sorted, in registers) with NO cross-thread traffic; a value is compared
//               against the current k'-th score first, so the common case costs one compare per similarity.
//               Two list forms: KEY lists (k' = 16 and list lifetimes of <= 16 tiles: a candidate is one float -- the score
//               with the column in its low 12 mantissa bits -- kept sorted with min / max alone, a whole 32-column chunk
//               merged by a sorting network while the list warms up) and compare / select (score, index) lists otherwise.
//   work        grid form: CTA = (query block, gallery split); linear form (key lists, operands L2-resident): the
//               (query block, tile) items are cut into equal contiguous ranges over all SMs, a range crossing query blocks
//               restarts its lists (cheap with key lists).  Each CTA writes k' candidates per row and piece to a workspace slot.
// Kernel 2  sim_select_kernel    (one warp per query row)
//   merges the row's candidate lists (few lists: bitonic folding through shuffles; more: rank by binary search in shared
//   memory), re-scores in exact fp32, from the fp32 operands (parity mode), the prefix of the survivors that the bf16 (+ key
//   truncation) error bound cannot exclude from the top-k -- query row in registers, 16 candidates per pass, their partial sums
//   reduced together by a reduce-scatter butterfly -- and writes the final top-k ordered by (similarity desc, index asc).
//
// Measurements that shaped this file (profiles/r2*_sim_*, r3*_sim_*):
//   a single-CTA tcgen05.mma reads 96 of the SM's 128 B/clk of shared memory: 75 % tensor-pipe ceiling, reached at the
//   1M-gallery shard shape (0.95 of the sustained cuBLAS rate).  At the small C3 shape the time was list warm-up (~17 us per
//   list restart with compare / select lists), 100 of 148 SMs busy, and a selection kernel as long as half the search: key
//   lists + linear ranges + the re-score rewrite took it from 0.1506 to 0.1157 ms on the same box, results bit-identical.
//   CTA pairs (cta_group::2), TMA multicast, a persistent pair grid, 8 / 16 epilogue warps, pipelined TMEM loads and a deferred
//   insert queue were built, validated and measured without gain; they are kept under profiles/r2_experiments/, not here.
#include "common.cuh"
#include <cuda.h>
#include <math.h>
#include <mutex>
#include <unordered_map>

namespace cvpce {

// ------------------------------------------------------------------------------------------------------------------------
// tile configuration
// ------------------------------------------------------------------------------------------------------------------------
constexpr int BM = 128;            // query rows per CTA (UMMA M)
constexpr int BN = 256;            // gallery rows per tile (UMMA N)
constexpr int BK = 64;             // bf16 elements per k-block = 128 bytes = one swizzle span
constexpr int UMMA_K = 16;
constexpr int A_TILE_BYTES = BM * BK * 2;   // 16 KB
constexpr int B_TILE_BYTES = BN * BK * 2;   // 32 KB
constexpr int kStagesStream = 4;   // ring depth when A and B both stream (48 KB per stage)
constexpr int kStagesResident = 3; // ring depth when A is resident (32 KB per stage)
constexpr int kMaxResidentKB = 8;  // D <= 512
constexpr int kTcThreads = 256;
constexpr int kEpilogueWarp0 = 4;
constexpr uint32_t kTmemCols = 512;

// ------------------------------------------------------------------------------------------------------------------------
// PTX wrappers
// ------------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}

// D[tmem] (+)= A[smem] * B[smem]^T, bf16 inputs, f32 accumulate, single CTA
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// 32 lanes x 32 consecutive fp32 columns: thread t of the warp receives row (lane base + t), columns [col, col+32)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
// 32 lanes x 64 consecutive fp32 columns in one instruction
__device__ __forceinline__ void tmem_ld_32x64(uint32_t taddr, uint32_t (&r)[64]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
      "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
      "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]), "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]),
        "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]),
        "=r"(r[46]), "=r"(r[47]), "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]),
        "=r"(r[55]), "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// the same wait, tied to the registers an earlier (asynchronous) tcgen05.ld writes: no use of r[] can be scheduled above it
__device__ __forceinline__ void tmem_ld_wait_for(uint32_t (&r)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]), "+r"(r[9]),
                 "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]), "+r"(r[17]), "+r"(r[18]),
                 "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]),
                 "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
}

// Shared-memory matrix descriptor: K-major operand, 128-byte swizzle, rows at a 128-byte pitch, 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);  // start address
  d |= (uint64_t)0 << 16;                        // leading byte offset: unused for a single swizzle atom along K
  d |= (uint64_t)(1024u >> 4) << 32;             // stride byte offset between 8-row groups
  d |= (uint64_t)1 << 46;                        // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                        // layout type: SWIZZLE_128B
  return d;
}
// Instruction descriptor for kind::f16: D=f32, A=B=bf16, both K-major, N=256, M=128.
constexpr uint32_t kInstrDesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

// r[j] for a run-time j without spilling the register array to local memory: a 5-level select tree (31 SELs)
__device__ __forceinline__ uint32_t pick32(const uint32_t (&r)[32], int j) {
  uint32_t a[16], b[8], c[4], d[2];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = (j & 1) ? r[2 * i + 1] : r[2 * i];
#pragma unroll
  for (int i = 0; i < 8; ++i) b[i] = (j & 2) ? a[2 * i + 1] : a[2 * i];
#pragma unroll
  for (int i = 0; i < 4; ++i) c[i] = (j & 4) ? b[2 * i + 1] : b[2 * i];
#pragma unroll
  for (int i = 0; i < 2; ++i) d[i] = (j & 8) ? c[2 * i + 1] : c[2 * i];
  return (j & 16) ? d[1] : d[0];
}

// ------------------------------------------------------------------------------------------------------------------------
// per-thread running top-KP, sorted descending, in registers.  Equal scores keep arrival order (= ascending index).
// ------------------------------------------------------------------------------------------------------------------------
template <int KP>
struct TopList {
  float s[KP];
  int i[KP];
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int j = 0; j < KP; ++j) {
      s[j] = -INFINITY;
      i[j] = -1;
    }
  }
  __device__ __forceinline__ float threshold() const { return s[KP - 1]; }
  // requires v > s[KP-1].  Every new entry is a function of the OLD list only (ge[j] is monotone because the list is sorted),
  // so the 2*KP selects are independent: with one epilogue warp per scheduler a serial bubble-up was latency-bound
  // (15 dependent compare -> select steps per insert; measured: ~60 us of a 200 us launch at the C3 shape).
  __device__ __forceinline__ void insert(float v, int idx) {
    bool ge[KP];
#pragma unroll
    for (int j = 0; j < KP; ++j) ge[j] = s[j] >= v;  // equal scores keep arrival order (= ascending index)
#pragma unroll
    for (int j = KP - 1; j > 0; --j) {
      s[j] = ge[j] ? s[j] : (ge[j - 1] ? v : s[j - 1]);
      i[j] = ge[j] ? i[j] : (ge[j - 1] ? idx : i[j - 1]);
    }
    s[0] = ge[0] ? s[0] : v;
    i[0] = ge[0] ? i[0] : idx;
  }
};

// ------------------------------------------------------------------------------------------------------------------------
// Key lists (short pieces: at most 16 gallery tiles = 4096 columns per list lifetime).
// A candidate is ONE float: the similarity with its low 12 mantissa bits replaced by the column inside the piece, encoded so
// that among similarities of the same 2^-11-relative bucket the lower column is the larger key.  A sorted list of keys is
// then maintained with min / max alone (no index array, no compare + select pairs): inserting costs 31 FMNMX instead of
// ~100 compare / select instructions, and a whole 32-column chunk can be merged by a sorting network (~480 FMNMX, the same
// for every lane) while a list warms up, instead of one warp-wide insert per event of any lane.  The price is that
// candidates are ranked by similarities truncated to 11 mantissa bits (2^-12 relative, 1/16 of the bf16 operand rounding the
// re-score margin already covers); the selection kernel adds it to its error bound.
// ------------------------------------------------------------------------------------------------------------------------
constexpr int kKeyColBits = 12;
constexpr uint32_t kKeyColMask = (1u << kKeyColBits) - 1u;
constexpr int kKeyMaxTiles = (1 << kKeyColBits) / BN;  // 16
constexpr int kKeyNetworkMin = 6;  // survivors of some lane in a chunk from which the sorting network is cheaper than the inserts

__device__ __forceinline__ float make_key(uint32_t bits, uint32_t col) {
  // positive floats grow with their low bits, negative ones shrink: store (4095 - col) resp. col so that a lower column wins
  const uint32_t low = (bits & 0x80000000u) ? col : (kKeyColMask - col);
  return __uint_as_float((bits & ~kKeyColMask) | low);
}
// smallest float of the key's similarity bucket: every similarity that could produce a key above `key` is >= it
__device__ __forceinline__ float key_bucket_floor(float key) {
  if (key == -INFINITY) return key;
  const uint32_t b = __float_as_uint(key);
  return __uint_as_float((b & 0x80000000u) ? (b | kKeyColMask) : (b & ~kKeyColMask));
}
__device__ __forceinline__ void key_decode(float key, int col_base, float& score, int& idx) {
  if (key == -INFINITY) {
    score = -INFINITY;
    idx = -1;
    return;
  }
  const uint32_t b = __float_as_uint(key);
  const uint32_t low = b & kKeyColMask;
  idx = col_base + (int)((b & 0x80000000u) ? low : (kKeyColMask - low));
  score = __uint_as_float(b & ~kKeyColMask);  // the bucket's representative: independent of where the piece started
}

// compare-exchange, larger first
__device__ __forceinline__ void ce_desc(float& a, float& b) {
  const float hi = fmaxf(a, b), lo = fminf(a, b);
  a = hi;
  b = lo;
}
// in-register bitonic sort, descending; N a power of two, every index a compile-time constant
template <int N>
__device__ __forceinline__ void bitonic_sort_desc(float (&a)[N]) {
#pragma unroll
  for (int k = 2; k <= N; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const int l = i ^ j;
        if (l > i) {
          if ((i & k) == 0) ce_desc(a[i], a[l]);
          else ce_desc(a[l], a[i]);
        }
      }
    }
  }
}
// bitonic sequence -> sorted descending
template <int N>
__device__ __forceinline__ void bitonic_merge_desc(float (&a)[N]) {
#pragma unroll
  for (int j = N >> 1; j > 0; j >>= 1) {
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const int l = i ^ j;
      if (l > i) ce_desc(a[i], a[l]);
    }
  }
}

struct KeyList {  // 16 keys, sorted descending; empty slots are -inf
  float s[16];
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int j = 0; j < 16; ++j) s[j] = -INFINITY;
  }
  __device__ __forceinline__ float threshold() const { return s[15]; }
  // keys are unique, so the sorted insert is min / max only; a key below the last entry falls off
  __device__ __forceinline__ void insert(float key) {
    float n[16];
    n[0] = fmaxf(s[0], key);
#pragma unroll
    for (int j = 1; j < 16; ++j) n[j] = fminf(s[j - 1], fmaxf(s[j], key));
#pragma unroll
    for (int j = 0; j < 16; ++j) s[j] = n[j];
  }
  // the 16 best of (list, 32 keys): sort the halves, keep the upper half of each merge
  __device__ __forceinline__ void merge32(float (&k)[32]) {
    float a[16], b[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      a[j] = k[j];
      b[j] = k[16 + j];
    }
    bitonic_sort_desc<16>(a);
    bitonic_sort_desc<16>(b);
#pragma unroll
    for (int j = 0; j < 16; ++j) a[j] = fmaxf(a[j], b[15 - j]);  // bitonic, holds the chunk's 16 best
    bitonic_merge_desc<16>(a);
#pragma unroll
    for (int j = 0; j < 16; ++j) s[j] = fmaxf(s[j], a[15 - j]);  // bitonic, holds the 16 best of list and chunk
    bitonic_merge_desc<16>(s);
  }
};

// One 32-column chunk against a key list.  col = column of r[0] inside the piece (< 4096 - 31); n_valid < 32 on the tail tile.
__device__ __forceinline__ void scan_chunk_keys(KeyList& top, const uint32_t (&r)[32], uint32_t col, int n_valid) {
  const float floor_ = key_bucket_floor(top.threshold());
  uint32_t m = 0;
#pragma unroll
  for (int j = 0; j < 32; ++j) m |= (__uint_as_float(r[j]) >= floor_) ? (1u << j) : 0u;  // NaN never passes
  if (n_valid < 32) m &= (1u << n_valid) - 1u;
  const unsigned most = __reduce_max_sync(0xffffffffu, (unsigned)__popc(m));
  if (most == 0) return;
  if (most >= (unsigned)kKeyNetworkMin) {
    float k[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) k[j] = ((m >> j) & 1u) ? make_key(r[j], col + j) : -INFINITY;
    top.merge32(k);
    return;
  }
  while (m) {
    const int j = __ffs(m) - 1;
    m &= m - 1;
    top.insert(make_key(pick32(r, j), col + j));
  }
}

// Work distribution.  An ITEM is (query block of 128 rows, gallery tile of 256 rows), numbered m_blk * n_tiles + tile.
//   grid   (LINEAR = false): CTA b = (m_blk = b % m_blocks, split = b / m_blocks) takes tiles [split*tps, (split+1)*tps) of ONE
//          query block; the CTAs of a wave stream the same gallery range in step (a gallery larger than L2 is read from HBM
//          about once per wave).  The row's list of this CTA goes to slot `split`.
//   linear (LINEAR = true):  CTA b takes items [b*N/G, (b+1)*N/G), G = gridDim.x = the SM count: every SM gets the same number
//          of tiles (+-1) whatever m_blocks is (C3: 50 query blocks x 32 tiles = 10.8 tiles on each of 148 SMs instead of 16
//          tiles on 100 of them).  The range is cut into PIECES at the query-block boundaries, one list lifetime each; the
//          list of a piece goes to slot b - (first CTA that touches the query block).  Pays only with the key lists (a list
//          restart costs ~17 us with the compare / select lists) and when both operands stay L2-resident.
struct TcParams {
  int Q, G, D;
  int n_tiles;          // ceil(G / BN)
  int m_blocks;         // ceil(Q / BM)
  int n_split;          // grid: gallery splits (gridDim.x = m_blocks * n_split, m fastest)
  int tiles_per_split;  // grid: ceil(n_tiles / n_split)
  long long n_items;    // linear: m_blocks * n_tiles
  int max_lists;        // lists per row in the workspace (grid: n_split)
  float* cand_s;        // [Q, max_lists, KP]
  int* cand_i;
};

// first CTA of a linear launch whose range [c*N/G, (c+1)*N/G) holds item x
__host__ __device__ __forceinline__ long long tc_owner(long long x, long long G, long long N) { return ((x + 1) * G - 1) / N; }

// the pieces of a CTA's item range, in order (the three roles walk them independently and identically)
template <bool LINEAR>
struct PieceWalk {
  long long it, end;
  int n_tiles;
  bool first;
  __device__ __forceinline__ PieceWalk(const TcParams& p) : n_tiles(p.n_tiles), first(true) {
    if (LINEAR) {
      it = (long long)blockIdx.x * p.n_items / gridDim.x;
      end = ((long long)blockIdx.x + 1) * p.n_items / gridDim.x;
    } else {
      const int m_blk = blockIdx.x % p.m_blocks, split = blockIdx.x / p.m_blocks;
      it = (long long)m_blk * p.n_tiles + min(p.n_tiles, split * p.tiles_per_split);
      end = (long long)m_blk * p.n_tiles + min(p.n_tiles, (split + 1) * p.tiles_per_split);
    }
  }
  // next piece: tiles [tb, te) of query block m_blk.  The grid form yields exactly one piece, even an empty one (empty
  // gallery: the row's list must still be written).
  __device__ __forceinline__ bool next(const TcParams& p, int& m_blk, int& tb, int& te) {
    if (!LINEAR) {
      if (!first) return false;
      first = false;
      m_blk = blockIdx.x % p.m_blocks;
      tb = (int)(it - (long long)m_blk * n_tiles);
      te = (int)(end - (long long)m_blk * n_tiles);
      return true;
    }
    if (it >= end) return false;
    m_blk = (int)(it / n_tiles);
    tb = (int)(it - (long long)m_blk * n_tiles);
    te = (int)min((long long)n_tiles, (long long)tb + (end - it));
    it += te - tb;
    return true;
  }
};

// KEYS: key lists in the epilogue (KP == 16 and at most kKeyMaxTiles tiles per piece; the host guarantees both)
// LINEAR: linear item ranges (see TcParams); the query tiles stream (a range changes query blocks)
template <bool RESIDENT_A, int KP, bool KEYS, bool LINEAR>
__global__ void __launch_bounds__(kTcThreads, 1)
sim_topk_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_g, const TcParams p) {
  static_assert(!(LINEAR && RESIDENT_A), "a linear range changes query blocks: the query tiles must stream");
  constexpr int kStages = RESIDENT_A ? kStagesResident : kStagesStream;
  constexpr int kStageBytes = RESIDENT_A ? B_TILE_BYTES : (A_TILE_BYTES + B_TILE_BYTES);
  extern __shared__ __align__(1024) uint8_t smem_dyn[];
  // 1024-byte alignment is required by the 128B swizzle atoms
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
  uint8_t* a_res = smem;                                                   // RESIDENT_A: [KB][128x64] tiles
  uint8_t* ring = smem + (RESIDENT_A ? kMaxResidentKB * A_TILE_BYTES : 0);  // [kStages][kStageBytes]

  __shared__ __align__(8) uint64_t full_bar[kStagesStream];
  __shared__ __align__(8) uint64_t empty_bar[kStagesStream];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ __align__(8) uint64_t a_full_bar;
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int KB = p.D / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_g);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full_bar[a], 1);
      mbar_init(&tmem_empty_bar[a], 4);  // one arrival per epilogue warp
    }
    mbar_init(&a_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(&tmem_base_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // ===================================================== TMA producer =====================================================
    if (lane == 0) {
      if (RESIDENT_A) {
        const int m0 = (int)(blockIdx.x % p.m_blocks) * BM;
        mbar_expect_tx(&a_full_bar, (uint32_t)(KB * A_TILE_BYTES));
        for (int kb = 0; kb < KB; ++kb) tma_load_2d(a_res + kb * A_TILE_BYTES, &tmap_q, &a_full_bar, kb * BK, m0);
      }
      int stage = 0;
      uint32_t phase = 0;
      PieceWalk<LINEAR> walk(p);
      int m_blk, t0, t1;
      while (walk.next(p, m_blk, t0, t1)) {
      const int m0 = m_blk * BM;
      for (int t = t0; t < t1; ++t) {
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_expect_tx(&full_bar[stage], (uint32_t)kStageBytes);
          uint8_t* st = ring + stage * kStageBytes;
          if (!RESIDENT_A) {
            tma_load_2d(st, &tmap_q, &full_bar[stage], kb * BK, m0);
            tma_load_2d(st + A_TILE_BYTES, &tmap_g, &full_bar[stage], kb * BK, t * BN);
          } else {
            tma_load_2d(st, &tmap_g, &full_bar[stage], kb * BK, t * BN);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
      }
    }
  } else if (warp == 1) {
    // ====================================================== MMA issuer ======================================================
    if (lane == 0) {
      if (RESIDENT_A) {
        mbar_wait(&a_full_bar, 0);
        tc_fence_after();
      }
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc_phase = 0;  // bit a = phase of accumulator a
      int acc = 0;             // the accumulators alternate over the CTA's whole range, across pieces
      PieceWalk<LINEAR> walk(p);
      int m_blk, t0, t1;
      while (walk.next(p, m_blk, t0, t1)) {
      for (int t = t0; t < t1; ++t, acc ^= 1) {
        mbar_wait(&tmem_empty_bar[acc], ((acc_phase >> acc) & 1u) ^ 1u);  // epilogue has drained this accumulator
        acc_phase ^= 1u << acc;
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint8_t* st = ring + stage * kStageBytes;
          const uint32_t a_addr = RESIDENT_A ? smem_u32(a_res + kb * A_TILE_BYTES) : smem_u32(st);
          const uint32_t b_addr = RESIDENT_A ? smem_u32(st) : smem_u32(st + A_TILE_BYTES);
          const uint64_t a_desc = make_smem_desc(a_addr);
          const uint64_t b_desc = make_smem_desc(b_addr);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the swizzle atom: +2 in the (address >> 4) field
            umma_bf16(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), kInstrDesc, (uint32_t)((kb | k) != 0));
          }
          umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs have read it
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full_bar[acc]);  // accumulator complete -> epilogue
      }
      }
    }
  } else if (warp >= kEpilogueWarp0) {
    // ======================================================= epilogue =======================================================
    const int ew = warp - kEpilogueWarp0;      // == warp % 4: the TMEM lane quadrant this warp may read
    TopList<KEYS ? 1 : KP> top;
    KeyList ktop;
    uint32_t acc_phase = 0;
    int acc = 0;
    PieceWalk<LINEAR> walk(p);
    int m_blk, t0, t1;
    while (walk.next(p, m_blk, t0, t1)) {
    const int row = m_blk * BM + ew * 32 + lane;  // query row owned by this thread in this piece
    if constexpr (KEYS) ktop.init();
    else top.init();
    for (int t = t0; t < t1; ++t, acc ^= 1) {
      mbar_wait(&tmem_full_bar[acc], (acc_phase >> acc) & 1u);
      acc_phase ^= 1u << acc;
      tc_fence_after();
      const int n0 = t * BN;
      const int valid = min(BN, p.G - n0);  // columns beyond the gallery are TMA zero-fill: must not become candidates
      const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        if (c0 >= valid) break;  // warp-uniform
        uint32_t r[32];
        tmem_ld_32x32(taddr + (uint32_t)c0, r);
        tmem_ld_wait();
        if constexpr (KEYS) {
          scan_chunk_keys(ktop, r, (uint32_t)((t - t0) * BN + c0), valid - c0);
        } else {
          // common case: one compare per similarity against the row's current k'-th score; the (rare) survivors are
          // collected in a bitmask and inserted below in ascending column order, so equal scores keep the lower index
          const float thr = top.threshold();
          uint32_t m = 0;
#pragma unroll
          for (int j = 0; j < 32; ++j) m |= (__uint_as_float(r[j]) > thr) ? (1u << j) : 0u;
          if (c0 + 32 > valid) m &= (1u << (valid - c0)) - 1u;  // tail tile: drop the zero-filled columns
          while (m) {
            const int j = __ffs(m) - 1;
            m &= m - 1;
            const float v = __uint_as_float(pick32(r, j));
            if (v > top.threshold()) top.insert(v, n0 + c0 + j);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
    }
    if (row < p.Q) {
      const int slot = LINEAR ? (int)((long long)blockIdx.x - tc_owner((long long)m_blk * p.n_tiles, gridDim.x, p.n_items))
                              : (int)(blockIdx.x / p.m_blocks);
      float* cs = p.cand_s + ((size_t)row * p.max_lists + slot) * KP;
      int* ci = p.cand_i + ((size_t)row * p.max_lists + slot) * KP;
      if constexpr (KEYS) {
#pragma unroll
        for (int j = 0; j < 16; j += 4) {
          float4 sc;
          int4 ix;
          key_decode(ktop.s[j], t0 * BN, sc.x, ix.x);
          key_decode(ktop.s[j + 1], t0 * BN, sc.y, ix.y);
          key_decode(ktop.s[j + 2], t0 * BN, sc.z, ix.z);
          key_decode(ktop.s[j + 3], t0 * BN, sc.w, ix.w);
          *reinterpret_cast<float4*>(cs + j) = sc;
          *reinterpret_cast<int4*>(ci + j) = ix;
        }
      } else {
#pragma unroll
        for (int j = 0; j < KP; j += 4) {
          *reinterpret_cast<float4*>(cs + j) = make_float4(top.s[j], top.s[j + 1], top.s[j + 2], top.s[j + 3]);
          *reinterpret_cast<int4*>(ci + j) = make_int4(top.i[j], top.i[j + 1], top.i[j + 2], top.i[j + 3]);
        }
      }
    }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------------------------------
// select: merge split candidates, exact fp32 re-score, final ordering.  One warp per query row.
// ------------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool before(float sa, int ia, float sb, int ib) {
  // strict total order (score desc, index asc); invalid (idx < 0) entries last
  if ((ia < 0) || (ib < 0)) return (ib < 0) && (ia >= 0);
  return sa > sb || (sa == sb && ia < ib);
}

// Where the final [Q,k] block goes: one local buffer, or -- gallery-sharded search -- the same slot of a [W,Q,k] gather
// buffer on EVERY rank of the node (peer pointers over NVLink): the all-gather is the epilogue's store, row by row,
// overlapped with the re-scoring of the other rows, instead of a collective after the kernel.
constexpr int kMaxOutTargets = 16;
struct OutTargets {
  int n;
  float* s[kMaxOutTargets];
  long long* i[kMaxOutTargets];
};

// The workspace holds n_split candidate lists of KP entries per row (one per gallery split), each sorted by (score desc,
// index asc); indices are unique across the lists of a row.

constexpr int kSelWarps = 4;
constexpr int kSelBitonicLists = 8;  // up to this many lists per row are folded in by bitonic merges
constexpr int kSelMaxCand = 512;  // candidates of a row staged in shared memory (more take the serial merge)

// exact fp32 re-score of the first n_rescore candidates of a row (lane c holds candidate c in my_i and receives its score)
template <bool FULL>
__device__ __forceinline__ void rescore_rows(const float* __restrict__ qr, const float* __restrict__ g32, const float* __restrict__ g_inv,
                                             float qi, int D, int n_rescore, int my_i, float& my_s, int lane) {
  float4 qv[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const int d = u * 128 + lane * 4;
    qv[u] = (FULL || d < D) ? *reinterpret_cast<const float4*>(qr + d) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  // 16 candidates per pass, fully unrolled: the per-lane partial sums of all of them are then reduced together by a
  // reduce-scatter butterfly (16 shuffles for 16 candidates instead of 5 each; the same pairing order 16-8-4-2-1 as warp_sum,
  // so every score has the bits a per-candidate butterfly gives).  Candidate c of the pass ends up in lanes 2c and 2c+1.
  for (int base = 0; base < n_rescore; base += 16) {  // warp-uniform (a second pass only with 32-entry lists)
    float part[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      part[c] = 0.f;
      if (base + c < n_rescore) {
        const int gi = __shfl_sync(0xffffffffu, my_i, base + c);
        const float* gr = g32 + (size_t)gi * D + lane * 4;
        float4 gv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
          gv[u] = (FULL || u * 128 + lane * 4 < D) ? __ldg(reinterpret_cast<const float4*>(gr + u * 128)) : make_float4(0.f, 0.f, 0.f, 0.f);
        float ax = 0.f, ay = 0.f, az = 0.f, aw = 0.f;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          ax = fmaf(qv[u].x, gv[u].x, ax);
          ay = fmaf(qv[u].y, gv[u].y, ay);
          az = fmaf(qv[u].z, gv[u].z, az);
          aw = fmaf(qv[u].w, gv[u].w, aw);
        }
        part[c] = (ax + ay) + (az + aw);
      }
    }
#pragma unroll
    for (int half = 8, bit = 16; half >= 1; half >>= 1, bit >>= 1) {
#pragma unroll
      for (int j = 0; j < half; ++j) {
        const bool up = (lane & bit) != 0;
        const float send = up ? part[j] : part[j + half];
        const float keep = up ? part[j + half] : part[j];
        part[j] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
      }
    }
    part[0] += __shfl_xor_sync(0xffffffffu, part[0], 1);
    const float tot = __shfl_sync(0xffffffffu, part[0], (2 * (lane - base)) & 31);
    if (lane >= base && lane < base + 16 && lane < n_rescore) {
      const float ginv = g_inv ? __ldg(g_inv + my_i) : 1.f;
      my_s = tot * qi * ginv;
    }
  }
}

// which lists of the workspace belong to a row (see TcParams)
struct SelLists {
  int linear;       // 0: every row has n_lists0 lists; 1: one list per CTA of the linear launch that touches the row's query block
  int stride;       // lists per row in the workspace
  int n_lists0;
  int n_tiles, n_ctas;
  long long n_items;
};

template <int KP>
__global__ void __launch_bounds__(kSelWarps * 32) sim_select_kernel(const float* __restrict__ cand_s, const int* __restrict__ cand_i, int Q,
                                                                    const SelLists sl, int k, long long idx_base, int D, float approx_slack,
                                                                    const float* __restrict__ q32, const float* __restrict__ g32,
                                                                    const float* __restrict__ q_inv, const float* __restrict__ g_inv,
                                                                    const OutTargets out) {
  __shared__ float s_cs[kSelWarps][kSelMaxCand];
  __shared__ int s_ci[kSelWarps][kSelMaxCand];
  __shared__ float s_os[kSelWarps][KP];
  __shared__ int s_oi[kSelWarps][KP];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int row = blockIdx.x * kSelWarps + wib;
  if (row >= Q) return;
  int n_lists = sl.n_lists0;
  if (sl.linear) {
    const long long first = (long long)(row / BM) * sl.n_tiles;
    n_lists = (int)(tc_owner(first + sl.n_tiles - 1, sl.n_ctas, sl.n_items) - tc_owner(first, sl.n_ctas, sl.n_items) + 1);
  }
  const int C = n_lists * KP;
  const float* cs = cand_s + (size_t)row * sl.stride * KP;
  const int* ci = cand_i + (size_t)row * sl.stride * KP;
  // ---- stage 1: the KP best candidates by kernel-1 score; lane j ends up holding the j-th
  float my_s = -INFINITY;
  int my_i = -1;
  if (n_lists == 1) {
    if (lane < KP) {
      my_s = cs[lane];
      my_i = ci[lane];
    }
  } else if (n_lists <= kSelBitonicLists) {
    // few lists: fold them in one after the other.  The running best-KP sits in lanes 0..KP-1 in descending order; lane j takes
    // entry KP-1-j of the next list, the element-wise maximum of the two is a bitonic sequence holding the KP best of both, and
    // log2(KP) shuffle stages sort it again (~50 instructions per list; ranking every candidate in every other list by binary
    // search cost ~600 per row with four lists)
    float rs = -INFINITY;
    int ri = -1;
    if (lane < KP) {
      rs = cs[lane];
      ri = ci[lane];
    }
    for (int l = 1; l < n_lists; ++l) {
      float bs = -INFINITY;
      int bi = -1;
      if (lane < KP) {
        bs = cs[l * KP + (KP - 1 - lane)];
        bi = ci[l * KP + (KP - 1 - lane)];
      }
      if (before(bs, bi, rs, ri)) {
        rs = bs;
        ri = bi;
      }
#pragma unroll
      for (int j = KP / 2; j > 0; j >>= 1) {
        const float os = __shfl_xor_sync(0xffffffffu, rs, j);
        const int oi = __shfl_xor_sync(0xffffffffu, ri, j);
        const bool take = (lane & j) == 0 ? before(os, oi, rs, ri) : before(rs, ri, os, oi);
        if (take) {
          rs = os;
          ri = oi;
        }
      }
    }
    my_s = rs;
    my_i = ri;
  } else if (C <= kSelMaxCand) {
    // merge by rank: all candidates into shared memory (coalesced), then every candidate finds its overall rank = its position
    // in its own list + the number of entries of every other list that precede it (binary search: the lists are sorted in
    // the strict total order `before`); ranks < KP are the answer
    float* ws = s_cs[wib];
    int* wi = s_ci[wib];
    for (int c = lane; c < C; c += 32) {
      ws[c] = cs[c];
      wi[c] = ci[c];
    }
    __syncwarp();
    float* os = s_os[wib];
    int* oi = s_oi[wib];
    if (lane < KP) {
      os[lane] = -INFINITY;
      oi[lane] = -1;
    }
    __syncwarp();
    for (int base = 0; base < C; base += 32) {  // warp-uniform trip count
      const int c = base + lane;
      const float xs = c < C ? ws[c] : -INFINITY;
      const int xi = c < C ? wi[c] : -1;
      if (xi >= 0) {  // not the padding of a short list
        const int own = c / KP;
        int rank = c - own * KP;
        for (int l = 0; l < n_lists && rank < KP; ++l) {
          if (l == own) continue;
          const float* ls = ws + l * KP;
          const int* li = wi + l * KP;
          int lo = 0, hi = KP;  // first position of list l that does NOT precede x
#pragma unroll
          for (int it = 0; it < (KP == 16 ? 5 : 6); ++it) {
            if (lo < hi) {
              const int mid = (lo + hi) >> 1;
              if (before(ls[mid], li[mid], xs, xi)) lo = mid + 1;
              else hi = mid;
            }
          }
          rank += lo;
        }
        if (rank < KP) {  // ranks are unique (strict total order, unique indices): no two lanes write the same slot
          os[rank] = xs;
          oi[rank] = xi;
        }
      }
    }
    __syncwarp();
    if (lane < KP) {
      my_s = os[lane];
      my_i = oi[lane];
    }
  } else {
    // very many lists: KP rounds of "best head wins"; lane l owns the heads of lists l, l+32, ...
    constexpr int kMaxOwn = 8;  // up to 256 lists
    int pos[kMaxOwn];
#pragma unroll
    for (int o = 0; o < kMaxOwn; ++o) pos[o] = 0;
    for (int r = 0; r < KP; ++r) {
      float bs = -INFINITY;
      int bi = -1, bo = 0;
#pragma unroll
      for (int o = 0; o < kMaxOwn; ++o) {
        const int sp = lane + 32 * o;
        if (sp < n_lists && pos[o] < KP) {
          const float sc = cs[sp * KP + pos[o]];
          const int ix = ci[sp * KP + pos[o]];
          if (ix >= 0 && before(sc, ix, bs, bi)) {
            bs = sc;
            bi = ix;
            bo = o;
          }
        }
      }
      float ws = bs;
      int wi = bi;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float os = __shfl_xor_sync(0xffffffffu, ws, o);
        const int oi = __shfl_xor_sync(0xffffffffu, wi, o);
        if (before(os, oi, ws, wi)) {
          ws = os;
          wi = oi;
        }
      }
      if (wi < 0) break;  // warp-uniform: candidates exhausted
      if (wi == bi) {     // indices are unique across lists: exactly one lane advances
#pragma unroll
        for (int o = 0; o < kMaxOwn; ++o)
          if (o == bo) ++pos[o];
      }
      if (lane == r) {
        my_s = ws;
        my_i = wi;
      }
    }
  }
  // ---- stage 2: exact fp32 re-score (one candidate at a time, the lanes split D: a (query, gallery row) score does not depend
  //      on which other candidates the row has, so the single-GPU and the sharded search agree bit for bit).
  //      Lane j holds the candidate of approximate rank j.  Only a PREFIX of them can matter: with eps an upper bound of
  //      |approximate - exact| (unit rows in bf16: <= 2^-8 * sum|q_i g_i| <= 2^-8, plus the fp32 accumulation of D products),
  //      the k best approximate candidates all have exact scores >= a_k - eps (a_k = k-th approximate score), while a candidate
  //      -- or a gallery row that never made the list -- with approximate score < a_k - 2 eps has an exact score < a_k - eps:
  //      it cannot enter the top-k and is not re-scored.  Sparse score distributions (the 1M-row gallery) re-score ~10 of the
  //      16 candidates; dense ones (8k rows, scores within 0.01 of each other) all of them, at no extra cost in the loop
  //      (an a-posteriori test after every few candidates was measured 20 % slower there: it breaks up the load pipeline).
  if (q32 != nullptr) {
    const float* qr = q32 + (size_t)row * D;
    const float qi = q_inv ? q_inv[row] : 1.f;
    const float eps_bound = 0.00390625f * (1.f + 0.0009765625f) + (float)D * 1.1920929e-7f + approx_slack;  // slack: key lists rank truncated scores
    const float a_k = __shfl_sync(0xffffffffu, my_s, k - 1);  // -inf when the row has fewer than k candidates: everything stays
    const bool keep = lane < KP && my_i >= 0 && (lane < k || !(my_s < a_k - 2.f * eps_bound));
    const int n_rescore = __popc(__ballot_sync(0xffffffffu, keep));  // a prefix: the lanes are sorted by approximate score
    if (!keep) my_i = -1;
    if ((D & 3) == 0 && D <= 1024) {
      // The query row stays in registers (8 float4 per lane cover D <= 1024); per candidate: 8 independent 128-bit loads of
      // the gallery row and four independent FMA chains (one per float4 component; a single 32-long chain was the kernel's
      // main stall), summed (x + y) + (z + w), one butterfly.  D == 1024 (the VGG descriptor) runs without the d < D predicates.
      // A (query, gallery row) score depends only on the two rows and D.  The general loop below re-read the query slice for
      // every candidate and spent ~150 instructions per candidate: 46-66 us for 6400 rows, 40 % of the search at the C3 shape.
      if (D == 1024) rescore_rows<true>(qr, g32, g_inv, qi, D, n_rescore, my_i, my_s, lane);
      else rescore_rows<false>(qr, g32, g_inv, qi, D, n_rescore, my_i, my_s, lane);
    } else
    for (int c = 0; c < n_rescore; ++c) {
      const int gi = __shfl_sync(0xffffffffu, my_i, c);
      const float* gr = g32 + (size_t)gi * D;
      float acc = 0.f;
      if ((D & 3) == 0) {
        for (int d = lane * 4; d < D; d += 128) {
          const float4 a = *reinterpret_cast<const float4*>(qr + d);
          const float4 b = __ldg(reinterpret_cast<const float4*>(gr + d));
          acc = fmaf(a.x, b.x, acc);
          acc = fmaf(a.y, b.y, acc);
          acc = fmaf(a.z, b.z, acc);
          acc = fmaf(a.w, b.w, acc);
        }
      } else {
        for (int d = lane; d < D; d += 32) acc = fmaf(qr[d], __ldg(gr + d), acc);
      }
      acc = warp_sum(acc);
      const float sc = acc * qi * (g_inv ? g_inv[gi] : 1.f);
      if (lane == c) my_s = sc;
    }
  }
  // ---- stage 3: final rank among the KP survivors by (score desc, index asc)
  int rank = 0;
  for (int c = 0; c < KP; ++c) {
    const float os = __shfl_sync(0xffffffffu, my_s, c);
    const int oi = __shfl_sync(0xffffffffu, my_i, c);
    if (c != lane && before(os, oi, my_s, my_i)) ++rank;
  }
  const int n_valid = __popc(__ballot_sync(0xffffffffu, lane < KP && my_i >= 0));
  for (int t = 0; t < out.n; ++t) {
    float* out_s = out.s[t];
    long long* out_i = out.i[t];
    if (lane < KP && my_i >= 0 && rank < k) {
      out_s[(size_t)row * k + rank] = my_s;
      out_i[(size_t)row * k + rank] = idx_base + my_i;
    }
    // rows with fewer than k valid candidates: pad
    for (int j = n_valid + lane; j < k; j += 32) {
      out_s[(size_t)row * k + j] = -INFINITY;
      out_i[(size_t)row * k + j] = -1;
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------------
// host side: TMA descriptors (driver entry point fetched through the runtime, no -lcuda), launch plan
// ------------------------------------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

struct TmaKey {
  const void* ptr;
  long long rows;
  int D, box_rows;
  bool operator==(const TmaKey& o) const { return ptr == o.ptr && rows == o.rows && D == o.D && box_rows == o.box_rows; }
};
struct TmaKeyHash {
  size_t operator()(const TmaKey& k) const {
    return std::hash<const void*>()(k.ptr) ^ (std::hash<long long>()(k.rows) * 1315423911u) ^ (size_t)(k.D * 2654435761u) ^
           (size_t)k.box_rows;
  }
};
static std::mutex g_tma_mutex;
static std::unordered_map<TmaKey, CUtensorMap, TmaKeyHash> g_tma_cache;

void tma_cache_clear() {
  std::lock_guard<std::mutex> lock(g_tma_mutex);
  g_tma_cache.clear();
}

// 2-D bf16 tensor [rows, D] row-major; box = [box_rows, 64] with 128-byte swizzle; out-of-bounds rows read as zero.
static int make_tmap(const void* ptr, long long rows, int D, int box_rows, CUtensorMap* out) {
  TmaKey key{ptr, rows, D, box_rows};
  {
    std::lock_guard<std::mutex> lock(g_tma_mutex);
    auto it = g_tma_cache.find(key);
    if (it != g_tma_cache.end()) {
      *out = it->second;
      return CVPCE_OK;
    }
  }
  PFN_encodeTiled enc = get_encode();
  if (!enc) return set_err(CVPCE_ERR_CUDA, "sim_topk: cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {(cuuint64_t)D, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)D * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_err(CVPCE_ERR_CUDA, "sim_topk: cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  std::lock_guard<std::mutex> lock(g_tma_mutex);
  if (g_tma_cache.size() > 256) g_tma_cache.clear();
  g_tma_cache[key] = *out;
  return CVPCE_OK;
}

// k' candidates per row kept by the tcgen05 kernels: k + 8 (rounded up to 16 or 32) when they are re-scored in fp32 -- the
// margin that makes the exact top-k a subset of the bf16 top-k' -- so the re-score mode serves k <= 24.
static int kp_for(int k, bool rescore) {
  const int want = rescore ? k + 8 : k;
  return want <= 16 ? 16 : 32;
}

struct TcPlan {
  int m_blocks, n_tiles, n_split, tiles_per_split, kp;
  bool linear;   // linear item ranges over n_ctas CTAs (key lists only)
  int n_ctas, slots;
  long long n_items;
};

static TcPlan make_plan(int Q, int G, int D, int k, bool rescore) {
  TcPlan pl;
  pl.m_blocks = (Q + BM - 1) / BM;
  pl.n_tiles = (G + BN - 1) / BN;
  pl.kp = kp_for(k, rescore);
  const int sms = sm_count();
  // gallery splits: (a) keep every SM busy when there are few query blocks, (b) keep the gallery range one wave of CTAs
  // streams within ~64 MB so that it stays L2-resident while all query blocks of the wave pass over it, (c) avoid a
  // ragged last wave.  Every CTA pays a fixed cost (~20 us measured: barrier/TMEM set-up, first TMA round trips, warm-up of the
  // candidate lists) expressed in tile units so that the model stays dimensionless.
  const double tile_us = 2.0 * BM * BN * (double)D / 11.0e6;
  const double cta_overhead_tiles = 20.0 / tile_us;
  int best = 1;
  double best_cost = 1e30;
  const int max_split = pl.n_tiles < 64 ? pl.n_tiles : 64;
  for (int s = 1; s <= max_split; ++s) {
    const int tps = (pl.n_tiles + s - 1) / s;
    const int used = (pl.n_tiles + tps - 1) / tps;  // splits that actually get tiles
    if (used != s) continue;
    const long long ctas = (long long)pl.m_blocks * s;
    const long long waves = (ctas + sms - 1) / sms;
    double cost = (double)waves * (tps + cta_overhead_tiles);        // time in tile units
    const double range_mb = (double)tps * BN * D * 2 / (1024.0 * 1024.0);
    if (range_mb > 64.0 && pl.m_blocks > 1) cost *= 1.0 + 0.25 * (range_mb / 64.0 - 1.0) / 8.0;  // mild L2 penalty
    cost *= 1.0 + 0.002 * s;                                          // merge overhead per split
    if (cost < best_cost) {
      best_cost = cost;
      best = s;
    }
  }
  const int forced = tuning("sim_split");
  if (forced >= 1 && forced <= max_split) best = forced;
  pl.n_split = best;
  pl.tiles_per_split = (pl.n_tiles + best - 1) / best;
  pl.n_split = (pl.n_tiles + pl.tiles_per_split - 1) / pl.tiles_per_split;
  pl.linear = false;
  pl.n_ctas = pl.m_blocks * pl.n_split;
  pl.slots = pl.n_split;
  pl.n_items = (long long)pl.m_blocks * pl.n_tiles;
  // Linear ranges: the same number of tiles (+-1) on every SM.  Conditions: key lists (16 candidates, pieces of at most 16
  // tiles -- with them a list restart is cheap), both operands L2-resident (the CTAs no longer sweep the gallery in step),
  // no forced split, and the grid leaves SMs idle or its last wave ragged.
  const int want = tuning("sim_plan");  // 0: automatic; 1: grid; 2: linear whenever it is legal
  const long long per_cta = (pl.n_items + sms - 1) / sms;
  const double operand_mb = ((double)Q + (double)G) * D * 2 / (1024.0 * 1024.0);
  const bool legal = pl.kp == 16 && tuning("sim_keys") != 1 && forced == 0 && pl.n_items >= sms && per_cta <= kKeyMaxTiles;
  if (legal && want != 1) {
    const long long waves = ((long long)pl.n_ctas + sms - 1) / sms;
    const double grid_tiles = (double)waves * pl.tiles_per_split;  // tiles on the busiest SM with the grid
    if (want == 2 || (operand_mb <= 64.0 && grid_tiles > 1.15 * (double)per_cta)) {
      pl.linear = true;
      pl.n_ctas = sms;
      pl.slots = (int)(((long long)pl.n_tiles * sms + pl.n_items - 1) / pl.n_items) + 1;  // CTAs that can touch one query block
    }
  }
  return pl;
}

template <bool RES, int KP, bool KEYS, bool LINEAR = false>
static int launch_tc(const CUtensorMap& tq, const CUtensorMap& tg, const TcParams& p, int n_ctas, cudaStream_t st) {
  constexpr int stages = RES ? kStagesResident : kStagesStream;
  constexpr int stage_bytes = RES ? B_TILE_BYTES : (A_TILE_BYTES + B_TILE_BYTES);
  const size_t smem = 1024 + (RES ? (size_t)kMaxResidentKB * A_TILE_BYTES : 0) + (size_t)stages * stage_bytes;
  if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(&sim_topk_tc_kernel<RES, KP, KEYS, LINEAR>), smem)) return rc;
  sim_topk_tc_kernel<RES, KP, KEYS, LINEAR><<<(unsigned)n_ctas, kTcThreads, smem, st>>>(tq, tg, p);
  CVPCE_LAUNCH_OK("sim_topk_tc_kernel");
  return CVPCE_OK;
}

}  // namespace cvpce

extern "C" size_t cvpce_sim_topk_workspace_bytes(int Q, int G, int D, int k) {
  using namespace cvpce;
  if (Q <= 0 || G <= 0 || D <= 0 || k <= 0 || k > 32) return 0;
  // sized for k' = 32 and for whichever split count a tuning override may select
  const TcPlan a = make_plan(Q, G, D, k, true);
  size_t lists = (size_t)(a.slots > 64 ? a.slots : 64);  // any forced split count
  if (lists > (size_t)a.n_tiles + 1) lists = (size_t)a.n_tiles + 1;
  if (lists < (size_t)a.slots) lists = (size_t)a.slots;
  return (size_t)Q * lists * 32 * 8 + 256;
}

namespace cvpce {
static int sim_topk_impl(const void* q_bf16, const void* g_bf16, int Q, int G, int D, int k, int64_t idx_base,
                         const float* q_f32, const float* g_f32, const float* q_inv_norm, const float* g_inv_norm,
                         const OutTargets& targets, void* workspace, size_t workspace_bytes, void* stream) {
  float* out_scores = targets.s[0];
  long long* out_idx = targets.i[0];
  CVPCE_REQUIRE(Q >= 0 && G >= 0, CVPCE_ERR_INVALID_ARG, "sim_topk: negative size Q=%d G=%d", Q, G);
  CVPCE_REQUIRE(D >= 64 && D <= 4096 && D % 64 == 0, CVPCE_ERR_UNSUPPORTED, "sim_topk: D=%d must be a multiple of 64 in [64,4096] (use cvpce_sim_topk_f32)", D);
  CVPCE_REQUIRE(k >= 1 && k <= 32, CVPCE_ERR_UNSUPPORTED, "sim_topk: k=%d outside [1,32] (use cvpce_sim_topk_f32)", k);
  if (Q == 0) return CVPCE_OK;
  CVPCE_REQUIRE(q_bf16 && out_scores && out_idx && (g_bf16 || G == 0), CVPCE_ERR_INVALID_ARG, "sim_topk: null pointer");
  CVPCE_REQUIRE(aligned16(q_bf16) && aligned16(g_bf16), CVPCE_ERR_INVALID_ARG, "sim_topk: operands must be 16-byte aligned (TMA)");
  CVPCE_REQUIRE((q_f32 == nullptr) == (g_f32 == nullptr), CVPCE_ERR_INVALID_ARG, "sim_topk: q_f32 and g_f32 must be given together");
  CVPCE_REQUIRE(!q_f32 || (aligned16(q_f32) && aligned16(g_f32)), CVPCE_ERR_INVALID_ARG, "sim_topk: fp32 operands must be 16-byte aligned");
  const bool rescore = q_f32 != nullptr;
  CVPCE_REQUIRE(!rescore || k <= 24, CVPCE_ERR_UNSUPPORTED,
                "sim_topk: the exact re-score keeps k + 8 <= 32 candidates per row, so k=%d > 24 is not served (use cvpce_sim_topk_f32)", k);
  if (int rc = require_sm100()) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  uintptr_t ws = (reinterpret_cast<uintptr_t>(workspace) + 127) & ~(uintptr_t)127;
  CUtensorMap tq, tg;
  if (int rc = make_tmap(q_bf16, Q, D, BM, &tq)) return rc;
  const TcPlan pl = make_plan(Q, G > 0 ? G : 1, D, k, rescore);
  const int kp = pl.kp;
  const bool linear = pl.linear && G > 0;
  const bool resident = D <= kMaxResidentKB * BK && !linear;
  const size_t need = (size_t)Q * pl.slots * pl.kp * 8 + 256;
  CVPCE_REQUIRE(workspace && workspace_bytes >= need, CVPCE_ERR_WORKSPACE, "sim_topk: workspace of %zu bytes needed, %zu given", need, workspace_bytes);
  float* cand_s = reinterpret_cast<float*>(ws);
  int* cand_i = reinterpret_cast<int*>(ws + (size_t)Q * pl.slots * pl.kp * 4);
  TcParams p;
  p.Q = Q; p.G = G; p.D = D;
  p.n_tiles = G > 0 ? pl.n_tiles : 0;
  p.m_blocks = pl.m_blocks;
  p.n_split = pl.n_split;
  p.tiles_per_split = pl.tiles_per_split;
  p.n_items = pl.n_items;
  p.max_lists = pl.slots;
  p.cand_s = cand_s;
  p.cand_i = cand_i;
  if (int rc = make_tmap(G > 0 ? g_bf16 : q_bf16, G > 0 ? G : Q, D, BN, &tg)) return rc;
  int rc;
  const int n_ctas = linear ? pl.n_ctas : pl.m_blocks * pl.n_split;
  // key lists (cheap list warm-up) where a list lives for at most 16 tiles: the small problems whose time is set by the warm-up
  const bool keys = pl.kp == 16 && tuning("sim_keys") != 1 && (linear || pl.tiles_per_split <= kKeyMaxTiles);
  if (linear)                   rc = launch_tc<false, 16, true, true>(tq, tg, p, n_ctas, st);
  else if (pl.kp == 16 && keys) rc = resident ? launch_tc<true, 16, true>(tq, tg, p, n_ctas, st) : launch_tc<false, 16, true>(tq, tg, p, n_ctas, st);
  else if (pl.kp == 16)         rc = resident ? launch_tc<true, 16, false>(tq, tg, p, n_ctas, st) : launch_tc<false, 16, false>(tq, tg, p, n_ctas, st);
  else                          rc = resident ? launch_tc<true, 32, false>(tq, tg, p, n_ctas, st) : launch_tc<false, 32, false>(tq, tg, p, n_ctas, st);
  if (rc) return rc;
  SelLists sl;
  sl.linear = linear ? 1 : 0;
  sl.stride = pl.slots;
  sl.n_lists0 = pl.n_split;
  sl.n_tiles = p.n_tiles;
  sl.n_ctas = n_ctas;
  sl.n_items = p.n_items;
  const int grid = (Q + kSelWarps - 1) / kSelWarps;
  if (kp == 16)
    sim_select_kernel<16><<<grid, kSelWarps * 32, 0, st>>>(cand_s, cand_i, Q, sl, k, idx_base, D, keys ? 0.00049f : 0.f, q_f32, g_f32, q_inv_norm, g_inv_norm, targets);
  else
    sim_select_kernel<32><<<grid, kSelWarps * 32, 0, st>>>(cand_s, cand_i, Q, sl, k, idx_base, D, 0.f, q_f32, g_f32, q_inv_norm, g_inv_norm, targets);
  CVPCE_LAUNCH_OK("sim_select_kernel");
  return CVPCE_OK;
}
}  // namespace cvpce

extern "C" int cvpce_sim_topk(const void* q_bf16, const void* g_bf16, int Q, int G, int D, int k, int64_t idx_base,
                              const float* q_f32, const float* g_f32, const float* q_inv_norm, const float* g_inv_norm,
                              float* out_scores, int64_t* out_idx, void* workspace, size_t workspace_bytes, void* stream) {
  cvpce::OutTargets t;
  t.n = 1;
  t.s[0] = out_scores;
  t.i[0] = reinterpret_cast<long long*>(out_idx);
  return cvpce::sim_topk_impl(q_bf16, g_bf16, Q, G, D, k, idx_base, q_f32, g_f32, q_inv_norm, g_inv_norm, t, workspace,
                              workspace_bytes, stream);
}

extern "C" int cvpce_sim_topk_scatter(const void* q_bf16, const void* g_bf16, int Q, int G, int D, int k, int64_t idx_base,
                                      const float* q_f32, const float* g_f32, const float* q_inv_norm, const float* g_inv_norm,
                                      float* const* peer_scores, int64_t* const* peer_idx, int n_peers, void* workspace,
                                      size_t workspace_bytes, void* stream) {
  using namespace cvpce;
  CVPCE_REQUIRE(n_peers >= 1 && n_peers <= kMaxOutTargets && peer_scores && peer_idx, CVPCE_ERR_INVALID_ARG,
                "sim_topk_scatter: n_peers=%d outside [1,%d] or null pointer table", n_peers, kMaxOutTargets);
  OutTargets t;
  t.n = n_peers;
  for (int p = 0; p < n_peers; ++p) {
    CVPCE_REQUIRE(peer_scores[p] && peer_idx[p], CVPCE_ERR_INVALID_ARG, "sim_topk_scatter: peer %d has a null buffer", p);
    t.s[p] = peer_scores[p];
    t.i[p] = reinterpret_cast<long long*>(peer_idx[p]);
  }
  return sim_topk_impl(q_bf16, g_bf16, Q, G, D, k, idx_base, q_f32, g_f32, q_inv_norm, g_inv_norm, t, workspace, workspace_bytes,
                       stream);
}
