// exact_gemm.cu -- K3 tensor path: candidate generation for hybrid.ExactIndex.Search (hnsw-extensions/hybrid/
// exact.go:76-108) as a bf16 tcgen05 GEMM with a fused, provably safe threshold filter.
//
//   D[q][row] = sum_k Qh[q][k] * Xh[row][k]          (bf16 operands, fp32 accumulation in TMEM)
//   score     = |x|^2 - 2 D            (Euclidean: distance^2 - |q|^2, monotone in the distance)
//             = 1 - D / (|x| |q|)      (cosine distance)
// Every score is within eps_q of the same transform of the canonical fp32 distance the reference would sort by
// (bf16 rounding of both operands: |D - x.q| <= gamma |x||q|, gamma ~ 2^-7; plus fp32 rounding terms).  Per query a
// running upper bound tau_q on the k-th smallest score is kept (k slots of tile-block minima over disjoint row sets;
// shared between CTAs through an atomicMin), and every row with score <= tau_q + 2 eps_q is appended to the query's
// candidate list.  That list provably contains the k nearest rows; exact_index.cu re-evaluates it in canonical fp32.
//
// One CTA = 128 queries (UMMA M) x a range of 256-row tiles (UMMA N = 256), K in 64-element (128 B, SWIZZLE_128B)
// blocks.  Warp 0: TMA producer; warp 1: TMEM allocator + single-thread tcgen05.mma issuer; warps 2-5: epilogue
// (tcgen05.ld 32x32b, one query row per thread).  Accumulators are double-buffered in TMEM (2 x 256 columns).
// Bound: tensor pipe (2*B*N*dim flop) for B >~ 300 queries, HBM (2*N*dim bytes of bf16 rows) below.
#include <cuda.h>
#include <cuda_fp16.h>

#include <algorithm>
#include <cmath>
#include <cstring>

#include "exact_index.hpp"

namespace b200 {

constexpr int GM_M = 128;      // queries per CTA
constexpr int GM_N = 256;      // rows per tile
constexpr int GM_K = 64;       // bf16 elements per K block (128 B)
constexpr int GM_STAGES = 4;  // at most; 3 when k is large (GemmParams::stages)
constexpr int GM_CB = 8;   // candidates a thread stages in shared memory before it appends them to the query's list
constexpr int GM_A_BYTES = GM_M * GM_K * 2;   // 16 KB
constexpr int GM_B_BYTES = GM_N * GM_K * 2;   // 32 KB
// Epilogue warps: EW = 16 (four per TMEM lane quarter, each takes 64 of the tile's 256 columns) when the per-thread slot
// tables fit beside the operand stages, else 8 (128 columns each).  Four epilogue warps per scheduler hide the
// fixed-latency stalls that two cannot (ncu: issue slots 32 % busy with 8 warps).
constexpr int GM_MAX_K = 64;   // result size the slot table supports
template <int EW>
struct GemmShape {
  static constexpr int EPI = EW * 32;             // epilogue threads
  static constexpr int THREADS = 64 + EPI;        // + TMA producer warp + MMA warp
  static constexpr int COLS_W = GM_N / (EW / 4);  // columns of a tile one warp scans
  static constexpr int CHUNKS = COLS_W / 32;      // 32-column tcgen05.ld chunks per warp and tile
};

// ---- PTX wrappers ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t s_base_of(const void *p) { return smem_u32(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ void mbar_wait_backoff(uint32_t bar, uint32_t parity) {  // off the critical path
  uint32_t ok;
  for (;;) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) break;
    __nanosleep(64);
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void *tmap, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          dst),
      "l"(tmap), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(
          tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
}
// the loaded registers are operands of the wait, so no use of them can be scheduled ahead of it
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&v)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
                 "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]),
                 "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]),
                 "+r"(v[23]), "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]),
                 "+r"(v[30]), "+r"(v[31])
               :
               : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, "
      "[%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait16(uint32_t (&v)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
                 "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15])
               :
               : "memory");
}
__device__ __forceinline__ void tmem_ld4_sync(uint32_t taddr, uint32_t (&v)[4]) {  // load + wait (rare path)
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]) : : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start address >> 4 in bits
// [0,14), leading byte offset (unused for swizzled K-major) = 1 in [16,30), stride byte offset = 1024 B (eight 128 B
// rows) >> 4 in [32,46), version 1 in [46,48), layout type 2 (SWIZZLE_128B) in [61,64).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// The fold block is 16 columns (32 B) wide: K-major SWIZZLE_32B (layout type 6), eight 32 B rows = 256 B per group.
__device__ __forceinline__ uint64_t umma_desc_sw32(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(256 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)6 << 61);
}
// kind::f16 instruction descriptor: D = F32 (bits 4-5 = 1), A = B = F16 (bits 7-9, 10-12 = 0), both K-major,
// N >> 3 in bits 17-22, M >> 4 in bits 24-28.  fp16, not bf16: 11 significand bits make the provable error bound of a
// score 8x tighter, which is what decides how many candidates the fp32 re-rank has to read.
constexpr uint32_t GM_IDESC = (1u << 4) | ((uint32_t)(GM_N >> 3) << 17) | ((uint32_t)(GM_M >> 4) << 24);

__device__ __forceinline__ uint32_t ord_bits(float d) {
  uint32_t u = __float_as_uint(d);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord_float(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k);
}


struct GemmParams {
  const float *aux;        // [cap] |x|^2 (Euclidean) or 1/|x| (cosine); +inf / NaN for dead rows
  int metric, kblocks, tiles_total, tiles_per_split, warm_tiles, k, stages;
  int narrow;              // the last K block is the 16-column fold block (Euclidean): one MMA k-step, its own tensor maps
  long long B;
  const float *q_inv;      // [Bpad] Euclidean: -2 * 2^-(sx+sq); cosine: -2^-(sx+sq) / |q|  (sx, sq: operand scales)
  const float *q_margin;   // [Bpad] 2 * eps_q
  uint32_t *slots_g;       // [Bpad][k] order-preserving bits of the slot minima merged over all CTAs of a query
  int32_t *cand;           // [B][cand_cap]
  int32_t *cand_cnt;       // [B]
  int cand_cap;
};

// Operand region: `stages` ring slots of one K block each.  When the whole query tile (kblocks K blocks of A) fits in
// 64 KB it is loaded once and stays (a_res): the ring then carries row tiles only, and a third of the L2 -> SM
// traffic of a short-K problem disappears (ncu, dim 128: 30.6 GB of TMA loads per 10 k x 1 M search, 10 GB of it A).
// With the fold block (narrow) and a resident query tile, the fold block of the rows has a ring of its own (GM_NSLOTS
// slots of 256 rows x 32 B): in the main ring it would take a whole 32 KB slot, i.e. the ring would hold 1.3 instead of
// 2 tiles, and the MMA issuer ends up waiting for loads.
constexpr int GM_NSLOTS = 2;
constexpr int GM_AE_BYTES = GM_M * 32, GM_BE_BYTES = GM_N * 32;  // fold block of the queries / of a row tile
__host__ __device__ inline int gm_a_bytes(int kblocks, int narrow) {
  return narrow ? (kblocks - 1) * GM_A_BYTES + GM_AE_BYTES : kblocks * GM_A_BYTES;
}
__host__ __device__ inline bool gm_a_resident(int kblocks, int narrow) { return gm_a_bytes(kblocks, narrow) <= 64 * 1024; }
__host__ __device__ inline int gm_operand_bytes(int kblocks, int stages, int narrow) {
  return gm_a_resident(kblocks, narrow)
             ? gm_a_bytes(kblocks, narrow) + stages * GM_B_BYTES + (narrow ? GM_NSLOTS * GM_BE_BYTES : 0)
             : stages * (GM_A_BYTES + GM_B_BYTES);
}
template <int EW>
struct GemmSmem {  // S = bytes of the operand region (gm_operand_bytes)
  using Sh = GemmShape<EW>;
  static __host__ __device__ size_t slots(int S) { return (size_t)S; }                              // float[EPI][k]
  static __host__ __device__ size_t aux(int S, int k) { return slots(S) + (size_t)Sh::EPI * k * 4; }  // float[EW][2][COLS_W]
  static __host__ __device__ size_t cbuf(int S, int k) { return aux(S, k) + (size_t)EW * 2 * Sh::COLS_W * 4; }  // int[GM_CB][EPI]
  static __host__ __device__ size_t bars(int S, int k) { return cbuf(S, k) + (size_t)GM_CB * Sh::EPI * 4; }  // barriers, tmem ptr
  static __host__ __device__ size_t total(int S, int k) { return bars(S, k) + 24 * 8 + 1024; }        // + alignment slack
};

__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ float lds_f1(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_f1(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
// score of one row from its accumulator: |x|^2 - 2 q.x (Euclidean, a = |x|^2) or 1 - q.x / (|x||q|) (cosine, a = 1/|x|)
template <int METRIC>
__device__ __forceinline__ float gm_score(float acc, float a, float qf) {
  return METRIC == 0 ? fmaf(qf, acc, a) : fmaf(acc * a, qf, 1.f);
}

// four scores with two packed FMAs (FFMA2; each half a plain IEEE rn operation, so the values are those of gm_score)
template <int METRIC>
__device__ __forceinline__ void gm_score4(uint32_t v0, uint32_t v1, uint32_t v2, uint32_t v3, const float4 &a, f32x2 qf2,
                                          float &s0, float &s1, float &s2, float &s3) {
  const f32x2 acc01 = pk2(__uint_as_float(v0), __uint_as_float(v1)), acc23 = pk2(__uint_as_float(v2), __uint_as_float(v3));
  const f32x2 a01 = pk2(a.x, a.y), a23 = pk2(a.z, a.w);
  f32x2 r01, r23;
  if (METRIC == 0) {
    r01 = fma2(qf2, acc01, a01);
    r23 = fma2(qf2, acc23, a23);
  } else {
    const f32x2 one2 = pk2(1.f, 1.f);
    f32x2 m01, m23;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(m01) : "l"(acc01), "l"(a01));
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(m23) : "l"(acc23), "l"(a23));
    r01 = fma2(m01, qf2, one2);
    r23 = fma2(m23, qf2, one2);
  }
  asm("mov.b64 {%0, %1}, %2;" : "=f"(s0), "=f"(s1) : "l"(r01));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(s2), "=f"(s3) : "l"(r23));
}

// min.f32 with three inputs (sm_100: one FMNMX3); a NaN input is dropped unless all are NaN, as with fminf
__device__ __forceinline__ float fmin3(float a, float b, float c) {
  float r;
  asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}

__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}

template <int METRIC, int EW>
__global__ void __launch_bounds__(GemmShape<EW>::THREADS, 1)
exact_gemm_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_x,
                  const __grid_constant__ CUtensorMap tmap_qe, const __grid_constant__ CUtensorMap tmap_xe, GemmParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  using Sh = GemmShape<EW>;
  using Sm = GemmSmem<EW>;
  constexpr int EPI = Sh::EPI, COLS_W = Sh::COLS_W, CHUNKS = Sh::CHUNKS;
  const int S = p.stages;
  const bool a_res = gm_a_resident(p.kblocks, p.narrow);
  const int OB = gm_operand_bytes(p.kblocks, S, p.narrow);
  const uint32_t a_bytes = a_res ? (uint32_t)gm_a_bytes(p.kblocks, p.narrow) : 0u;  // resident query tile, in front of the ring
  const bool nring = a_res && p.narrow;                   // the rows' fold block has its own ring behind the main one
  const int main_kb = nring ? p.kblocks - 1 : p.kblocks;  // K blocks that go through the main ring
  const uint32_t n_base = s_base_of(smem) + a_bytes + (uint32_t)S * (a_res ? GM_B_BYTES : GM_A_BYTES + GM_B_BYTES);
  const uint32_t st_bytes = a_res ? GM_B_BYTES : GM_A_BYTES + GM_B_BYTES;   // one ring slot
  const uint32_t b_off = a_res ? 0u : (uint32_t)GM_A_BYTES;                 // row tile inside a ring slot
  float *slots = reinterpret_cast<float *>(smem + Sm::slots(OB));
  int *cbuf = reinterpret_cast<int *>(smem + Sm::cbuf(OB, p.k));
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + Sm::bars(OB, p.k));
  uint32_t *tmem_ptr_s = reinterpret_cast<uint32_t *>(bars + 2 * GM_STAGES + 5);
  const uint32_t bar_a = smem_u32(bars + 2 * GM_STAGES + 4);
  const uint32_t bar_full = smem_u32(bars), bar_empty = smem_u32(bars + GM_STAGES);
  const uint32_t bar_tfull = smem_u32(bars + 2 * GM_STAGES), bar_tempty = smem_u32(bars + 2 * GM_STAGES + 2);
  const uint32_t bar_nfull = smem_u32(bars + 2 * GM_STAGES + 6), bar_nempty = smem_u32(bars + 2 * GM_STAGES + 8);
  const uint32_t s_base = smem_u32(smem);

  // x = query tile: CTAs that are scheduled together sweep the SAME row tiles in the same order, so a row tile comes
  // from HBM once and from L2 for every other query tile
  const int mt = blockIdx.x, sp = blockIdx.y;
  const int t_begin = sp * p.tiles_per_split;
  const int t_end = min(p.tiles_total, t_begin + p.tiles_per_split);
  const int n_tiles = t_end - t_begin;
  const int warm = min(p.warm_tiles, n_tiles);
  const int seq = n_tiles > 0 ? warm + n_tiles : 0;  // warm-up tiles are visited twice: first only to bound tau

  if (threadIdx.x == 0) {
    for (int i = 0; i < S; i++) {
      mbar_init(bar_full + 8 * i, 1);
      mbar_init(bar_empty + 8 * i, 1);
    }
    for (int i = 0; i < 2; i++) {
      mbar_init(bar_tfull + 8 * i, 1);
      mbar_init(bar_tempty + 8 * i, EW);
    }
    mbar_init(bar_a, 1);
    for (int i = 0; i < GM_NSLOTS; i++) {
      mbar_init(bar_nfull + 8 * i, 1);
      mbar_init(bar_nempty + 8 * i, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // TMEM: 2 accumulator stages x 256 fp32 columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_s)),
                 "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_s;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0, phase = 0;
      if (a_res && seq > 0) {
        mbar_expect_tx(bar_a, a_bytes);
        for (int kb = 0; kb < p.kblocks; kb++) {
          const bool ext = p.narrow && kb == p.kblocks - 1;  // the fold block: its own map, 32 B rows
          tma_load_2d(s_base + kb * GM_A_BYTES, ext ? &tmap_qe : &tmap_q, ext ? 0 : kb * GM_K, mt * GM_M, bar_a);
        }
      }
      for (int idx = 0; idx < seq; idx++) {
        const int tile = t_begin + (idx < warm ? idx : idx - warm);
        if (nring) {  // the rows' fold block of this tile: 8 KB into its own ring
          const int ns = idx % GM_NSLOTS, nph = (idx / GM_NSLOTS) & 1;
          mbar_wait_backoff(bar_nempty + 8 * ns, nph ^ 1);
          mbar_expect_tx(bar_nfull + 8 * ns, GM_BE_BYTES);
          tma_load_2d(n_base + ns * GM_BE_BYTES, &tmap_xe, 0, tile * GM_N, bar_nfull + 8 * ns);
        }
        for (int kb = 0; kb < main_kb; kb++) {
          mbar_wait_backoff(bar_empty + 8 * stage, phase ^ 1);
          const uint32_t sa = s_base + a_bytes + stage * st_bytes, sb = sa + b_off;
          const bool ext = p.narrow && kb == p.kblocks - 1;  // the fold block moves 32 B per row
          mbar_expect_tx(bar_full + 8 * stage, ext ? (a_res ? 0u : (uint32_t)GM_M * 32u) + (uint32_t)GM_N * 32u : st_bytes);
          if (!a_res) tma_load_2d(sa, ext ? &tmap_qe : &tmap_q, ext ? 0 : kb * GM_K, mt * GM_M, bar_full + 8 * stage);
          tma_load_2d(sb, ext ? &tmap_xe : &tmap_x, ext ? 0 : kb * GM_K, tile * GM_N, bar_full + 8 * stage);
          if (++stage == S) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      int stage = 0, phase = 0;
      if (a_res && seq > 0) mbar_wait(bar_a, 0);
      for (int idx = 0; idx < seq; idx++) {
        const int as = idx & 1;
        mbar_wait(bar_tempty + 8 * as, ((idx >> 1) & 1) ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * GM_N;
        for (int kb = 0; kb < main_kb; kb++) {
          mbar_wait(bar_full + 8 * stage, phase);
          tc_fence_after();
          const uint32_t ring = s_base + a_bytes + stage * st_bytes;
          const uint32_t sa = a_res ? s_base + kb * GM_A_BYTES : ring, sb = ring + b_off;
          const bool ext = p.narrow && kb == p.kblocks - 1;
          const uint64_t adesc = ext ? umma_desc_sw32(sa) : umma_desc_sw128(sa);
          const uint64_t bdesc = ext ? umma_desc_sw32(sb) : umma_desc_sw128(sb);
          const int k4n = ext ? 1 : GM_K / 16;  // the fold block holds 16 columns
#pragma unroll
          for (int k4 = 0; k4 < GM_K / 16; k4++)  // UMMA K = 16 fp16 = 32 B: advance the start address field by 2
            if (k4 < k4n) umma_bf16(tmem_d, adesc + 2 * k4, bdesc + 2 * k4, GM_IDESC, (kb | k4) != 0);
          umma_commit(bar_empty + 8 * stage);  // frees the smem stage when these MMAs retire
          if (++stage == S) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (nring) {  // one more k-step: the fold blocks (queries: resident behind the main blocks; rows: own ring)
          const int ns = idx % GM_NSLOTS, nph = (idx / GM_NSLOTS) & 1;
          mbar_wait(bar_nfull + 8 * ns, nph);
          tc_fence_after();
          umma_bf16(tmem_d, umma_desc_sw32(s_base + main_kb * GM_A_BYTES), umma_desc_sw32(n_base + ns * GM_BE_BYTES),
                    GM_IDESC, 1);
          umma_commit(bar_nempty + 8 * ns);
        }
        umma_commit(bar_tfull + 8 * as);  // accumulator ready
      }
    }
  } else {
    // ===== epilogue: one query row per thread, COLS_W columns of every tile per warp =====
    const int ew = warp & 3;                  // TMEM lane quarter this warp may access
    const int r = ew * 32 + lane;             // row of the tile = TMEM lane
    const long long q = (long long)mt * GM_M + r;
    const bool live_q = q < p.B;
    const int e = warp - 2;                   // 0..EW-1 among the epilogue warps
    const int et = e * 32 + lane;             // 0..EPI-1 among the epilogue threads
    const int ch0 = (e >> 2) * CHUNKS;        // first 32-column chunk of this warp's share of the tile
    const int kk = p.k;
    const uint32_t slots_a = smem_u32(slots) + (uint32_t)(et * kk) * 4u;  // this thread's k slot minima
    const uint32_t slots_end = slots_a + (uint32_t)kk * 4u;
    for (int j = 0; j < kk; j++) sts_f1(slots_a + j * 4, __int_as_float(0x7f800000));
    const float margin = live_q ? p.q_margin[q] : 0.f;
    const float qf = live_q ? p.q_inv[q] : 0.f;  // turns the accumulator into -2 q.x (Euclidean) or -q.x/|q| (cosine)
    const float INF = __int_as_float(0x7f800000);
    const f32x2 qf2 = pk2(qf, qf);
    float thr = INF, tau_loc = INF;  // tau_loc: max over this thread's slots when it was last computed
    int sub = 0;
    uint32_t sl_a = slots_a;                   // slot the next block minimum goes to
    // Candidates are staged per thread (cbuf[i][et], conflict-free) and appended to the query's global list a batch at
    // a time: one atomicAdd per batch instead of one per candidate, and the whole warp flushes together so that the
    // round trip to L2 is paid once per flush, not once per lane.
    int *mybuf = cbuf + et;
    int ccnt = 0;
    auto flush = [&]() {
      if (ccnt > 0) {
        const int at = atomicAdd(&p.cand_cnt[q], ccnt);
        for (int i = 0; i < ccnt; i++)
          if (at + i < p.cand_cap) p.cand[q * p.cand_cap + at + i] = mybuf[i * EPI];
        ccnt = 0;
      }
    };
    // Per-row terms (|x|^2 or 1/|x|) of this warp's columns: staged in a buffer only this warp touches (two tiles
    // deep), fetched one tile ahead, so no barrier between the epilogue warps is needed.
    constexpr int AV = COLS_W / 32;           // floats per lane and tile
    const uint32_t aux_w = s_base + (uint32_t)Sm::aux(OB, p.k) + (uint32_t)(e * 2 * COLS_W) * 4u;
    const float *aux_g = p.aux + ch0 * 32 + lane * AV;
    auto aux_fetch = [&](int tile, float (&nx)[AV]) {
      const float *src = aux_g + (long long)tile * GM_N;
#pragma unroll
      for (int i = 0; i < AV; i++) nx[i] = __ldg(src + i);
    };
    auto aux_store = [&](int buf, const float (&nx)[AV]) {
#pragma unroll
      for (int i = 0; i < AV; i++) sts_f1(aux_w + (uint32_t)(buf * COLS_W + lane * AV + i) * 4u, nx[i]);
    };
    // Euclidean: the per-row term is in the accumulator (fold block, see exact_xconv_kernel): score = qf * acc, the
    // smallest score of a block is qf * (its largest accumulator), and no per-row term is staged at all.
    constexpr bool FOLD = METRIC == 0;
    const float inv_qf = live_q && FOLD ? 1.f / qf : 0.f;
    float thr_acc = -INF;  // FOLD: accumulators >= thr_acc may pass (thr in accumulator units, rounded to the safe side)
    if (!FOLD && seq > 0) {
      float n0[AV];
      aux_fetch(t_begin, n0);
      aux_store(0, n0);
    }
    __syncwarp();
    int tile = t_begin;
#pragma unroll 1
    for (int idx = 0; idx < seq; idx++) {
      const int as = idx & 1;
      const int ntile = idx + 1 == warm ? t_begin : tile + 1;  // warm-up tiles are visited twice
      const bool emit = live_q && idx >= warm;
      const uint32_t ax = aux_w + (uint32_t)(as * COLS_W) * 4u;
      bool changed = false;
      float nx[AV];
      if (!FOLD && idx + 1 < seq) aux_fetch(ntile, nx);
      mbar_wait(bar_tfull + 8 * as, (idx >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + as * GM_N + ch0 * 32;
      // EARLY (fold epilogue, two blocks per warp and tile): the 64 accumulators of the tile come into registers at once
      // and the TMEM stage goes back to the MMA issuer BEFORE they are looked at, so the tile is scored under the MMAs
      // of the tile after the next instead of holding them up.
      constexpr bool EARLY = FOLD && CHUNKS <= 2;
      uint32_t va[16], vb[16];  // streaming: the two 16-column halves of a 32-column block, the next half in flight
      uint32_t wa[EARLY ? 2 * CHUNKS : 1][16];
      if constexpr (EARLY) {
#pragma unroll
        for (int h = 0; h < 2 * CHUNKS; h++) tmem_ld16(taddr + h * 16, wa[h]);
#pragma unroll
        for (int h = 0; h < 2 * CHUNKS; h++) tmem_ld_wait16(wa[h]);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_tempty + 8 * as);
      } else {
        tmem_ld16(taddr, va);
      }
#pragma unroll
      for (int cl = 0; cl < CHUNKS; cl++) {
        const uint32_t axc = ax + (uint32_t)(cl * 32) * 4u;
        float g8[8];  // FOLD: the largest accumulator of each group of four columns; else its smallest score
        float bmin;
        unsigned pass = 0;
        bool hot;
        if constexpr (FOLD) {
          if constexpr (!EARLY) {
            tmem_ld_wait16(va);
            tmem_ld16(taddr + cl * 32 + 16, vb);
          }
          uint32_t (&ha)[16] = EARLY ? wa[EARLY ? 2 * cl : 0] : va;
          uint32_t (&hb)[16] = EARLY ? wa[EARLY ? 2 * cl + 1 : 0] : vb;
#pragma unroll
          for (int j4 = 0; j4 < 4; j4++)  // dead rows hold -inf; fmax drops NaN
            g8[j4] = fmaxf(fmax3(__uint_as_float(ha[4 * j4]), __uint_as_float(ha[4 * j4 + 1]), __uint_as_float(ha[4 * j4 + 2])),
                           __uint_as_float(ha[4 * j4 + 3]));
          if constexpr (!EARLY) {
            tmem_ld_wait16(vb);
            if (cl + 1 < CHUNKS) tmem_ld16(taddr + (cl + 1) * 32, va);
          }
#pragma unroll
          for (int j4 = 0; j4 < 4; j4++)
            g8[4 + j4] = fmaxf(fmax3(__uint_as_float(hb[4 * j4]), __uint_as_float(hb[4 * j4 + 1]), __uint_as_float(hb[4 * j4 + 2])),
                               __uint_as_float(hb[4 * j4 + 3]));
          const float bmax = fmax3(fmax3(g8[0], g8[1], g8[2]), fmax3(g8[3], g8[4], g8[5]), fmaxf(g8[6], g8[7]));
          bmin = qf * bmax;  // qf < 0: multiplication by it is monotone, so this IS the smallest score of the block
          hot = __any_sync(FULL, emit && bmax >= thr_acc);
        } else {
          float4 a4[4];  // the per-row terms of the half are fetched under the latency of the accumulator load
#pragma unroll
          for (int j4 = 0; j4 < 4; j4++) a4[j4] = lds_f4(axc + j4 * 16);
          tmem_ld_wait16(va);
          tmem_ld16(taddr + cl * 32 + 16, vb);
#pragma unroll
          for (int j4 = 0; j4 < 4; j4++) {  // 16 independent scores, two per packed FMA; no branch, no dependent chain
            float s0, s1, s2, s3;
            gm_score4<METRIC>(va[4 * j4 + 0], va[4 * j4 + 1], va[4 * j4 + 2], va[4 * j4 + 3], a4[j4], qf2, s0, s1, s2, s3);
            g8[j4] = fminf(fmin3(s0, s1, s2), s3);  // FMNMX3 + FMNMX; like fminf they drop NaN (dead rows)
          }
#pragma unroll
          for (int j4 = 0; j4 < 4; j4++) a4[j4] = lds_f4(axc + (j4 + 4) * 16);
          tmem_ld_wait16(vb);
          if (cl + 1 < CHUNKS) tmem_ld16(taddr + (cl + 1) * 32, va);
#pragma unroll
          for (int j4 = 0; j4 < 4; j4++) {
            float s0, s1, s2, s3;
            gm_score4<METRIC>(vb[4 * j4 + 0], vb[4 * j4 + 1], vb[4 * j4 + 2], vb[4 * j4 + 3], a4[j4], qf2, s0, s1, s2, s3);
            g8[4 + j4] = fminf(fmin3(s0, s1, s2), s3);
          }
          bmin = fmin3(fmin3(g8[0], g8[1], g8[2]), fmin3(g8[3], g8[4], g8[5]), fminf(g8[6], g8[7]));
          hot = __any_sync(FULL, emit && bmin <= thr);
        }
        // Some row of this block passes some lane's filter (warp-uniform test: 32 lanes = 32 queries, so about a third
        // of the blocks get here): only the groups of four columns whose extremum passes are scored again, row by row.
        if (hot) {
#pragma unroll
          for (int j4 = 0; j4 < 8; j4++) {
            if (__any_sync(FULL, emit && (FOLD ? g8[j4] >= thr_acc : g8[j4] <= thr))) {
              // the accumulators are still in TMEM (the stage is released at the end of the tile): four columns again
              uint32_t w[4];
              if constexpr (EARLY) {  // the stage is gone: the accumulators are in registers
#pragma unroll
                for (int i = 0; i < 4; i++) w[i] = wa[EARLY ? 2 * cl + (j4 >> 2) : 0][4 * (j4 & 3) + i];
              } else {
                tmem_ld4_sync(taddr + cl * 32 + j4 * 4, w);
              }
              float s[4];
              if constexpr (FOLD) {
#pragma unroll
                for (int i = 0; i < 4; i++) s[i] = qf * __uint_as_float(w[i]);
              } else {
                const float4 a = lds_f4(axc + j4 * 16);
                gm_score4<METRIC>(w[0], w[1], w[2], w[3], a, qf2, s[0], s[1], s[2], s[3]);
              }
#pragma unroll
              for (int i = 0; i < 4; i++)  // dead rows score +inf / NaN
                pass |= (unsigned)((s[i] <= thr) & (s[i] < INF)) << (4 * j4 + i);
            }
          }
          if (!emit) pass = 0;
          const int np = __popc(pass);
          const int row0 = tile * GM_N + (ch0 + cl) * 32;
          if (ccnt + np > GM_CB) flush();
          if (np > GM_CB) {  // only while the threshold is still loose: straight to the list
            const int at = atomicAdd(&p.cand_cnt[q], np);
            int i = 0;
            while (pass) {
              const int j = __ffs(pass) - 1;
              pass &= pass - 1;
              if (at + i < p.cand_cap) p.cand[q * p.cand_cap + at + i] = row0 + j;
              i++;
            }
          } else {
            while (pass) {
              const int j = __ffs(pass) - 1;
              pass &= pass - 1;
              mybuf[ccnt * EPI] = row0 + j;
              ccnt++;
            }
          }
        }
        // slot `sub mod k` holds the minimum over a set of rows disjoint from every other slot's set
        {
          const float old = lds_f1(sl_a);
          if (bmin < old) {
            sts_f1(sl_a, bmin);
            changed |= old >= tau_loc;  // tau = max(slots) can only move when the slot that held the maximum is lowered
          }
        }
        sl_a += 4;
        if (sl_a == slots_end) sl_a = slots_a;
        sub++;
      }
      if constexpr (!EARLY) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_tempty + 8 * as);
      }
      if (__any_sync(FULL, ccnt >= GM_CB / 2)) flush();  // once per tile: a block that would overflow the stage flushes first
      // next tile's terms into the other buffer (this warp's reads of it ended one tile ago)
      if (!FOLD && idx + 1 < seq) aux_store(as ^ 1, nx);
      __syncwarp();
      if (sub >= kk && live_q) {
        // Now and then (often at first, when bounds improve fast) merge the slots with those of the other threads
        // and CTAs that work on this query: slot j of each covers rows disjoint from slot j' of any other, so the
        // element-wise minimum is again k minima over k disjoint row sets.
        const int done = idx + 1;
        if (sub - CHUNKS < kk || (done & (done - 1)) == 0 || (done & 63) == 0 || idx == seq - 1) {
          uint32_t *sg = p.slots_g + q * kk;
          // reductions without a return value (RED), then the merged values read back eight at a time: two round trips
          // to L2 for k = 10 instead of one per four slots.  A value read here may miss a concurrent update: every value
          // ever stored in sg[j] is the score of a row of class j, so whatever is read is a valid bound.
          for (int j = 0; j < kk; j++) atomicMin(&sg[j], ord_bits(lds_f1(slots_a + j * 4)));
          for (int j0 = 0; j0 < kk; j0 += 8) {
            uint32_t got[8];
#pragma unroll
            for (int i = 0; i < 8; i++) got[i] = j0 + i < kk ? __ldcg(&sg[j0 + i]) : 0xffffffffu;
#pragma unroll
            for (int i = 0; i < 8; i++)
              if (j0 + i < kk && got[i] < ord_bits(lds_f1(slots_a + (j0 + i) * 4)))
                sts_f1(slots_a + (j0 + i) * 4, ord_float(got[i]));
          }
          changed = true;
        }
        if (changed) {
          // k distinct rows score <= max(slots): an upper bound on the k-th smallest score of the whole index
          float tau = lds_f1(slots_a);
          for (int j = 1; j < kk; j++) tau = fmaxf(tau, lds_f1(slots_a + j * 4));
          tau_loc = tau;
          thr = fminf(thr, tau + margin);
          if constexpr (FOLD) {
            // s <= thr  <=>  acc >= thr / qf (qf < 0); two roundings here and one in s = qf * acc: four ulps of slack on
            // the permissive side, the rows themselves are tested with s <= thr
            const float t = thr * inv_qf;
            thr_acc = thr < INF ? t - fabsf(t) * 4.76837158e-7f : -INF;
          }
        }
      }
      tile = ntile;
    }
    flush();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}


// ---- preparation kernels ---------------------------------------------------------------------------------------------
// stats[0] = bits of max |x|^2, stats[1] = bits of min |x|^2 over rows with |x| > 0, stats[2] = bits of max |x_i|
// rows [first, first+count): |x|^2 -> aux (Euclidean) or 1/|x| (cosine), running statistics
__global__ void exact_xstats_kernel(const float *vec, int row_floats, int dim, long long first, long long count,
                                    int metric, float *aux, uint32_t *stats) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= count) return;
  const long long row = first + w;
  const float *src = vec + row * row_floats;
  float acc = 0.f, mx = 0.f;
  for (int i = lane; i < dim; i += 32) {
    const float v = src[i];
    acc = fmaf(v, v, acc);
    mx = fmaxf(mx, fabsf(v));
  }
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    acc += __shfl_xor_sync(FULL, acc, off);
    mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, off));
  }
  if (lane == 0) {
    aux[row] = metric == 0 ? acc : 1.0f / sqrtf(acc);  // a zero vector gives inf: its score is NaN, never a candidate
    atomicMax(&stats[0], __float_as_uint(acc));         // non-negative floats order like their bit patterns
    if (acc > 0.f) atomicMin(&stats[1], __float_as_uint(acc));
    if (mx < __int_as_float(0x7f800000)) atomicMax(&stats[2], __float_as_uint(mx));
  }
}
// rows [first, first+count): fp16 copy of x * 2^sx_exp, zero padded to kpad.
// Euclidean (vext != nullptr): also the FOLD BLOCK of the row, 16 extra K columns of which three are used:
//   t = |x|^2 2^(2 sx - 11) (|x|^2 = aux[row], the fp32 value the scores are defined with) = h1 + 2^-11 h2 + 2^-22 h3
// in fp16 pieces (each remainder is exact in fp32; what is left after three pieces is below 2^-33 t), stored negated.
// With the query's fold columns 2^(sq - sx + 10) (1, 2^-11, 2^-22) (exact_qprep_kernel) the GEMM itself yields
//   acc = 2^(sx+sq) (x.q - |x|^2 / 2),   score = |x|^2 - 2 x.q = -2 2^-(sx+sq) acc,
// so the epilogue takes maxima of raw accumulators instead of loading a per-row term and forming every score.
// A dead row (aux = +inf) gets -inf: its accumulator is -inf, its score +inf.  The scale keeps t < 2^14 (dim <= 2048,
// |x_i| 2^sx < 2^7: exact_after_add).
__global__ void exact_xconv_kernel(const float *vec, int row_floats, int dim, int kpad, long long first,
                                   long long count, int sx_exp, __half *vech, const float *aux, __half *vext) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= count) return;
  const long long row = first + w;
  const float *src = vec + row * row_floats;
  __half *dst = vech + row * kpad;
  for (int i = lane; i < kpad; i += 32) dst[i] = __float2half_rn(i < dim ? ldexpf(src[i], sx_exp) : 0.f);
  if (vext != nullptr && lane < 16) {
    const float a = aux[row];
    float h[3] = {0.f, 0.f, 0.f};
    if (a < __int_as_float(0x7f800000)) {
      float t = ldexpf(a, 2 * sx_exp - 11);
#pragma unroll
      for (int i = 0; i < 3; i++) {
        h[i] = __half2float(__float2half_rn(t));
        t = ldexpf(t - h[i], 11);
      }
    } else {
      h[0] = __int_as_float(0x7f800000);  // +inf or NaN norm: never a candidate
    }
    vext[row * 16 + lane] = __float2half_rn(lane < 3 ? -h[lane] : 0.f);
  }
}
// rows [first, first+n) of the fold block that hold no vector yet: dead
__global__ void exact_fill_vext_kernel(__half *vext, long long first, long long n) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n * 16) vext[first * 16 + i] = __float2half_rn((i & 15) == 0 ? -__int_as_float(0x7f800000) : 0.f);
}
__global__ void exact_fill_kernel(float *p, long long n, float v) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
// queries: fp16 copy of q * 2^sq (sq per query: the largest component lands in [2^13, 2^14)) into Qh[Bpad][kpad],
// the factor that turns an accumulator into the score's dot term, the filter margin 2*eps_q, tau = +inf, count = 0.
//
// Error of a score.  With X = x 2^sx, Q = q 2^sq, u = 2^-11 and nu = 2^-25 (half the fp16 subnormal spacing) every
// converted operand satisfies |fp16(X_i) - X_i| <= u |X_i| + nu, products of fp16 numbers are exact in fp32, and the
// fp32 accumulation of kpad of them loses at most (kpad + 64) 2^-22 of sum |X_i Q_i|.  Hence, in the units of x.q:
//   |D' 2^-(sx+sq) - x.q| <= g |x||q| + nu (1+u) sqrt(kpad) (|x| 2^-sq + |q| 2^-sx) + kpad nu^2 2^-(sx+sq),
//   g = (2u + u^2) + (kpad + 64) 2^-22
// (Cauchy-Schwarz for sum |x_i||q_i| and for the 1-norms).  The Euclidean score |x|^2 - 2 x.q inherits twice that with
// |x| <= xmax; the cosine score 1 - x.q/(|x||q|) inherits it divided by |x||q| with |x| >= xmin.  `fp32` covers the
// difference between the canonical fp32 distance and exact arithmetic, and the epilogue's own roundings.
// Euclidean: the query's three fold columns (see exact_xconv_kernel) are 2^(sq - sx + 10) (1, 2^-11, 2^-22); they have
// to be normal fp16 numbers, which holds for -2 <= sq - sx <= 5.  The query's own scale (largest component in
// [2^7, 2^8)) is moved into that window where it can be (a smaller scale always; a larger one by up to seven binades);
// a query it cannot be moved for (components 2^9 and more above the rows') keeps its own scale, gets zero fold columns
// and an infinite margin: every row passes, its candidate list overflows and the scan path answers it.
__global__ void exact_qprep_kernel(const float *q, long long B, long long Bpad, int dim, int kpad, int metric,
                                   const uint32_t *stats, int sx_exp, __half *qh, float *q_fac, float *q_margin,
                                   uint32_t *slots_g, int k, int32_t *cand_cnt, __half *qext) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= Bpad) return;
  const bool live = w < B;
  float acc = 0.f, mx = 0.f;
  for (int i = lane; i < dim; i += 32) {
    const float v = live ? q[w * dim + i] : 0.f;
    acc = fmaf(v, v, acc);
    mx = fmaxf(mx, fabsf(v));
  }
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    acc += __shfl_xor_sync(FULL, acc, off);
    mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, off));
  }
  int sq_exp = 0;
  if (mx > 0.f && mx < __int_as_float(0x7f800000)) {
    int e;
    frexpf(mx, &e);  // mx = m 2^e, m in [0.5, 1)
    sq_exp = (metric == 0 ? 8 : 14) - e;
  }
  bool fold_ok = true;
  if (metric == 0) {
    const int ideal = sq_exp;
    sq_exp = max(sx_exp - 2, min(sx_exp + 5, sq_exp));
    fold_ok = sq_exp <= ideal + 7;  // beyond that the query's largest component would overflow fp16
    if (!fold_ok) sq_exp = ideal;   // finite operands, no fold columns, infinite margin: every row passes -> scan path
  }
  sq_exp = max(-100 - sx_exp, min(100 - sx_exp, sq_exp));  // keep 2^-(sx+sq) a normal float
  for (int i = lane; i < kpad; i += 32)
    qh[w * kpad + i] = __float2half_rn((live && i < dim) ? ldexpf(q[w * dim + i], sq_exp) : 0.f);
  if (qext != nullptr && lane < 16)
    qext[w * 16 + lane] = __float2half_rn(live && fold_ok && lane < 3 ? ldexpf(1.f, sq_exp - sx_exp + 10 - 11 * lane) : 0.f);
  if (lane == 0) {
    const float qn = sqrtf(acc), xmax = sqrtf(__uint_as_float(stats[0])), xmin = sqrtf(__uint_as_float(stats[1]));
    const float u = 4.8828125e-4f, nu = 2.9802322e-8f;
    const float g = (2.f * u + u * u) * 1.01f + (float)(kpad + 64) * 2.3841858e-7f;
    const float isx = ldexpf(1.f, -sx_exp), isq = ldexpf(1.f, -sq_exp);
    const float sub = nu * (1.f + u) * sqrtf((float)kpad);  // subnormal part, per unit of |.| 2^-scale
    const float fp32 = (float)(dim / 8 + 64) * 5.9604645e-8f;
    const float dsc = ldexpf(1.f, -(sx_exp + sq_exp));
    float eps;
    if (metric == 0) {
      const float e_dot = g * xmax * qn + sub * (xmax * isq + qn * isx) + (float)kpad * nu * nu * dsc;
      // + the fold term: it is summed by the same fp32 accumulation (its magnitude |x|^2 / 2 in the units of x.q joins
      // sum |X_i Q_i|), and three fp16 pieces leave less than 2^-30 of it unrepresented
      const float e_fold = ((float)(kpad + 80) * 2.3841858e-7f + 9.3132257e-10f) * 0.5f * xmax * xmax;
      eps = 2.f * (e_dot + e_fold) + fp32 * (xmax + qn) * (xmax + qn);
      if (!fold_ok) eps = __int_as_float(0x7f800000);
      q_fac[w] = -2.f * dsc;
    } else {
      // divided by |x||q| with |x| >= xmin; a query of norm 0 scores NaN everywhere and selects nothing
      eps = g + sub * (isq / qn + isx / xmin) + (float)kpad * nu * nu * dsc / (xmin * qn) + fp32;
      q_fac[w] = -dsc / qn;
    }
    q_margin[w] = 2.f * eps * 1.01f;
    if (live) cand_cnt[w] = 0;
  }
  for (int j = lane; j < k; j += 32) slots_g[w * k + j] = 0xff800000u;  // ord_bits(+inf)
}

// ---- host side -------------------------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static encode_tiled_fn get_encoder() {
  static encode_tiled_fn fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (encode_tiled_fn)p;
  }
  return fn;
}
// fp16 matrix [rows][kpad], box = 64 elements (128 B) x box_rows, 128 B swizzle; the 16-column fold block (kpad = 16):
// box = 16 elements (32 B) x box_rows, 32 B swizzle.  (A 64-wide box over the 16-wide tensor also works -- the columns
// outside arrive as zeros -- but the counters show the full 128 B per row crossing the crossbar: 30.9 instead of
// 20.6 GB per 10 k x 1 M search, and the kernel became bound by it.)
static int make_tmap(CUtensorMap *m, void *base, uint64_t rows, uint64_t kpad, uint32_t box_rows) {
  encode_tiled_fn enc = get_encoder();
  if (!enc) {
    set_error("exact: cuTensorMapEncodeTiled is not available");
    return B200HNSW_ERR_UNSUPPORTED;
  }
  cuuint64_t gdim[2] = {kpad, rows};
  cuuint64_t gstride[1] = {kpad * 2};
  const bool narrow = kpad == 16;
  cuuint32_t box[2] = {narrow ? 16u : (cuuint32_t)GM_K, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   narrow ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("exact: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return B200HNSW_ERR_CUDA;
  }
  return 0;
}

int exact_after_add(b200exact *x, int64_t first, int64_t count, cudaStream_t s) {
  if (!x->d_maxn2) {
    B200_CUDA(cudaMalloc(&x->d_maxn2, 64));
    const uint32_t init[3] = {0u, 0x7f800000u, 0u};  // max |x|^2, min |x|^2 > 0, max |x_i|
    B200_CUDA(cudaMemsetAsync(x->d_maxn2, 0, 64, s));
    B200_CUDA(cudaMemcpyAsync(x->d_maxn2, init, sizeof init, cudaMemcpyHostToDevice, s));
  }
  if (x->aux_dead_cap != x->cap) {  // storage grew: rows that do not exist yet must never become candidates
    const float dead = x->metric == 0 ? __builtin_inff() : __builtin_nanf("");
    long long nfill = x->cap - first;
    exact_fill_kernel<<<(unsigned)((nfill + 255) / 256), 256, 0, s>>>(x->d_aux + first, nfill, dead);
    if (x->d_vext)
      exact_fill_vext_kernel<<<(unsigned)((nfill * 16 + 255) / 256), 256, 0, s>>>(x->d_vext, first, nfill);
    x->aux_dead_cap = x->cap;
  }
  long long threads = count * 32;
  exact_xstats_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(x->d_vec, x->row_floats, x->dim, first, count,
                                                                         x->metric, x->d_aux, x->d_maxn2);
  // the fp16 copy is scaled by a power of two chosen from the largest component seen so far (three binades of
  // headroom below fp16's 2^15); a later batch that does not fit re-converts everything under a new scale
  uint32_t st[3];
  B200_CUDA(cudaMemcpyAsync(st, x->d_maxn2, sizeof st, cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  float mx;
  memcpy(&mx, &st[2], 4);
  long long conv_first = first, conv_count = count;
  // Euclidean: the fold block carries |x 2^sx|^2 2^-11 in fp16, so the components stay below 2^7 (largest one in
  // [2^5, 2^6) when the scale is chosen, one binade of slack for later batches); fp16's relative precision does not
  // depend on the scale, and the subnormal terms of the error bound (exact_qprep_kernel) take the scale as it is
  const float fit = x->metric == B200HNSW_EUCLIDEAN ? 128.f : 32768.f;
  const int top = x->metric == B200HNSW_EUCLIDEAN ? 6 : 12;
  if (!x->have_scale || ldexpf(mx, x->sx_exp) > fit) {
    int e = 0;
    if (mx > 0.f) frexpf(mx, &e);  // mx = m 2^e, m in [0.5, 1)
    x->sx_exp = mx > 0.f ? top - e : 0;
    if (x->sx_exp > 100) x->sx_exp = 100;
    if (x->sx_exp < -100) x->sx_exp = -100;
    x->have_scale = true;
    conv_first = 0;
    conv_count = first + count;
  }
  threads = conv_count * 32;
  exact_xconv_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(x->d_vec, x->row_floats, x->dim, x->kpad,
                                                                        conv_first, conv_count, x->sx_exp, x->d_vech,
                                                                        x->d_aux, x->d_vext);
  B200_CUDA(cudaGetLastError());
  return 0;
}

template <int EW>
static int launch_gemm(b200exact *x, GemmParams &p, int k, const CUtensorMap &tq, const CUtensorMap &tqe, int m_tiles,
                       int splits, cudaStream_t s) {
  const size_t smem = GemmSmem<EW>::total(gm_operand_bytes(p.kblocks, p.stages, p.narrow), k);
  auto kern = x->metric == 0 ? exact_gemm_kernel<0, EW> : exact_gemm_kernel<1, EW>;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)m_tiles, (unsigned)splits);
  const CUtensorMap *tx = (const CUtensorMap *)x->tmap_x;
  kern<<<grid, GemmShape<EW>::THREADS, smem, s>>>(tq, tx[0], tqe, tx[1], p);
  B200_CUDA(cudaGetLastError());
  return 0;
}

int exact_candidates_tensor(b200exact *x, const float *d_queries, int64_t B, int k, int32_t *d_cand,
                            int32_t *d_cand_cnt, int cand_cap, cudaStream_t s) {
  if (k > GM_MAX_K) return B200HNSW_ERR_UNSUPPORTED;  // the slot table is sized for k <= 64: the scan handles the rest
  // widest epilogue whose slot tables leave room for at least three operand stages
  // (measured: 16 warps win while the epilogue is the bottleneck, dim <= 256; at dim 1024 the tensor pipe is, and the
  // looser per-thread bounds of 16 warps only add candidates: 2.6 vs 2.4 ms on 250 k x 1024)
  const bool fold = x->metric == B200HNSW_EUCLIDEAN;  // one more K block: the 16-column fold block
  const int kblocks = x->kpad / GM_K + (fold ? 1 : 0);
  int ew = x->kpad / GM_K <= 4 ? 16 : 8, stages = GM_STAGES;
  if (const char *v = getenv("B200EXACT_EPI_WARPS")) ew = atoi(v) == 8 ? 8 : 16;  // tuning experiments
  auto total = [&](int w, int st) {
    const int ob = gm_operand_bytes(kblocks, st, fold ? 1 : 0);
    return w == 16 ? GemmSmem<16>::total(ob, k) : GemmSmem<8>::total(ob, k);
  };
  if (ew == 16 && total(16, 3) > x->smem_optin) ew = 8;
  if (total(ew, stages) > x->smem_optin) stages = 3;
  if (total(ew, stages) > x->smem_optin) return B200HNSW_ERR_UNSUPPORTED;
  const int chunks = ew == 16 ? GemmShape<16>::CHUNKS : GemmShape<8>::CHUNKS;
  const int64_t Bpad = (B + GM_M - 1) / GM_M * GM_M;
  auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
  const size_t need = al(2 * (size_t)Bpad * x->kpad) + 2 * al(4 * (size_t)Bpad) + al(4 * (size_t)Bpad * k) +
                      al(2 * (size_t)Bpad * 16);
  if (need > x->gemm_ws_bytes) {
    if (x->d_gemm_ws) cudaFree(x->d_gemm_ws);
    x->d_gemm_ws = nullptr;
    x->gemm_ws_bytes = 0;
    B200_CUDA(cudaMalloc(&x->d_gemm_ws, need));
    x->gemm_ws_bytes = need;
  }
  char *w = (char *)x->d_gemm_ws;
  __half *qh = (__half *)w;
  w += al(2 * (size_t)Bpad * x->kpad);
  float *q_inv = (float *)w;
  w += al(4 * (size_t)Bpad);
  float *q_margin = (float *)w;
  w += al(4 * (size_t)Bpad);
  uint32_t *slots_g = (uint32_t *)w;
  w += al(4 * (size_t)Bpad * k);
  __half *qext = fold ? (__half *)w : nullptr;
  exact_qprep_kernel<<<(unsigned)((Bpad * 32 + 255) / 256), 256, 0, s>>>(d_queries, B, Bpad, x->dim, x->kpad, x->metric,
                                                                          x->d_maxn2, x->sx_exp, qh, q_inv, q_margin, slots_g,
                                                                          k, d_cand_cnt, qext);
  B200_CUDA(cudaGetLastError());
  int rc;
  if (!x->tmap_x) x->tmap_x = calloc(2, sizeof(CUtensorMap));  // [0] rows, [1] their fold block
  if (x->tmap_rows != x->cap) {
    CUtensorMap *tx = (CUtensorMap *)x->tmap_x;
    if ((rc = make_tmap(&tx[0], x->d_vech, (uint64_t)x->cap, (uint64_t)x->kpad, GM_N))) return rc;
    if (fold && (rc = make_tmap(&tx[1], x->d_vext, (uint64_t)x->cap, 16, GM_N))) return rc;
    if (!fold) tx[1] = tx[0];
    x->tmap_rows = x->cap;
  }
  CUtensorMap tq, tqe;
  if ((rc = make_tmap(&tq, qh, (uint64_t)Bpad, (uint64_t)x->kpad, GM_M))) return rc;
  if (fold) {
    if ((rc = make_tmap(&tqe, qext, (uint64_t)Bpad, 16, GM_M))) return rc;
  } else {
    tqe = tq;
  }
  GemmParams p{};
  p.aux = x->d_aux;
  p.metric = x->metric;
  p.kblocks = kblocks;
  p.narrow = fold ? 1 : 0;
  p.tiles_total = (int)((x->n + GM_N - 1) / GM_N);
  p.warm_tiles = (k + chunks - 1) / chunks;  // a thread fills `chunks` slots per tile
  p.stages = stages;
  p.k = k;
  p.B = B;
  p.q_inv = q_inv;
  p.q_margin = q_margin;
  p.slots_g = slots_g;
  p.cand = d_cand;
  p.cand_cnt = d_cand_cnt;
  p.cand_cap = cand_cap;
  const int m_tiles = (int)(Bpad / GM_M);
  // one CTA per SM (shared memory); splits of the row range so that the grid covers the SMs about once, but never so
  // short that the warm-up tiles dominate
  const int min_tiles = 8 * p.warm_tiles + 8;
  const int max_splits = std::max(1, std::min(p.tiles_total / min_tiles, 64));
  int splits = 1;
  double best_eff = 0;
  for (int sp = 1; sp <= max_splits; sp++) {  // fill whole waves of sm_count CTAs
    long long ctas = (long long)m_tiles * sp;
    double eff = (double)ctas / (double)(((ctas + x->sm_count - 1) / x->sm_count) * x->sm_count);
    if (eff > best_eff + 0.02) {
      best_eff = eff;
      splits = sp;
    }
  }
  p.tiles_per_split = (p.tiles_total + splits - 1) / splits;
  splits = (p.tiles_total + p.tiles_per_split - 1) / p.tiles_per_split;
  return ew == 16 ? launch_gemm<16>(x, p, k, tq, tqe, m_tiles, splits, s)
                  : launch_gemm<8>(x, p, k, tq, tqe, m_tiles, splits, s);
}

}  // namespace b200
