// fp32 GEMM family on the 5th-generation tensor cores: error-compensated TF32 ("3xTF32") on
// tcgen05.mma.kind::tf32 with fp32 accumulators in TMEM.
//
// The fp32 path of the reference-parity contract (1e-5 relative) cannot use bf16, and FFMA tops out near
// 20 TFLOP/s on the MNIST layers.  Every fp32 operand x is split as x = hi + lo with hi = trunc_tf32(x) (what
// the tensor core reads from a raw fp32 word) and lo = x - hi (exact in fp32, |lo| < 2^-10 |x|), and the
// product is evaluated as
//     A.B  ~=  A_hi.B_hi  +  (A_lo.B_hi + A_hi.B_lo)            (dropped: A_lo.B_lo ~ 2^-20)
// with the bracketed correction in its OWN TMEM accumulator, added in the epilogue -- the small terms are not
// rounded against the large running sum.
//
// Pipeline per CTA (persistent, one per SM, 512 threads):
//   warp 0      TMA producer: raw fp32 A / B tiles (128-byte swizzle, 32 fp32 per row) into the `hi` slots
//   warps 8-15  converters: read each 16-byte chunk of the stage and store its lo part (A: same offset of a lo
//               slot; B: behind B hi in the stage).  The split is elementwise, so it is done on the raw swizzled
//               bytes: no layout math.
//   warp 1      MMA issuer: 2 x (BLOCK_K / 8) tcgen05.mma per stage (A_hi.[B_hi | B_lo] with N = 2 BN, A_lo.B_hi)
//   warp 2      TMEM allocator; warps 4-7 epilogue (TMEM -> 32x32 transpose -> fused functor -> 16-byte stores)
// Same three operand layouts as the bf16 family (descriptor major-ness instead of transposes in memory).
#include <cuda.h>

#include <cstdlib>
#include <map>
#include <mutex>
#include <tuple>

#include "ddnn_umma.cuh"

namespace ddnn {
namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 32;  // 32 fp32 = 128 bytes = one SWIZZLE_128B row
constexpr int UMMA_K = 8;    // kind::tf32
constexpr int NUM_THREADS = 512;
constexpr int CONV_THREADS = 256;  // warps 8-15

struct Tf32Params {
  int M, N;
  int m_blocks, n_blocks, num_kb, splits, kb_per_split, group_m;
  int64_t zrows;  // row offset per split (split-K partials)
  float* db_part;  // dW split-K with one tile per CTA: column sums of the B operand, [split][db_np] (else null)
  int db_np;
  // Cluster split-K (template flag CL): the `splits` CTAs of a thread-block cluster own ONE output tile, each reduces
  // its share of the batch, and the partial tiles are summed through distributed shared memory -- no partials in
  // HBM, no reducer launch.  The CTAs then apply the epilogue themselves: SGD on W / bias (apply) or fp32 store.
  float* cl_w;      // apply: W_master (updated in place); else: dW_out
  float* cl_b;      // apply: bias; else: db_out
  float cl_lr;
  int cl_apply;
};

constexpr int CL_PITCH_PAD = 4;  // floats of padding per partial-tile row (keeps float4 alignment, spreads the banks)

__device__ __forceinline__ float4 ld_dsmem4(uint32_t local_addr, uint32_t cta) {
  uint32_t ra;
  float4 v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(local_addr), "r"(cta));
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(ra)
               : "memory");
  return v;
}

__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accum) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
      : "memory");
}

// kind::tf32 reads the top 19 bits of each fp32 operand word (the low 13 mantissa bits are ignored: measured -- with
// raw fp32 words as the `hi` operand and lo = x - trunc(x) the results stay at the 1e-6 level, a rounding unit
// would have shown 2^-11).  So the raw tile IS the hi operand and only lo has to be written.
__device__ __forceinline__ float trunc_tf32(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }
__device__ __forceinline__ float rn_tf32(float x) {  // nearest (ties away) on the bit pattern, finite inputs
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}

// K-major tile [rows][32 fp32] (row = 128 B, 8-row swizzle atoms 1024 B apart): step k (8 elements) by 32 bytes.
// MN-major tile: 32-bit MN-major operands exist only in the SWIZZLE_128B_BASE32B layout (32-byte swizzle units,
// 4-row period; TMA: CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B).  32-element MN atoms, each [32 k][32 mn] (k row =
// 128 B, 4 KB per atom = LBO); 4-k groups 512 B apart (SBO); step k (8 rows) by 1024 bytes.
template <bool MN_MAJOR>
__device__ __forceinline__ uint64_t tf32_desc(uint32_t tile_base, int k) {
  if (MN_MAJOR) return make_desc_layout(tile_base + k * (UMMA_K * 128), BLOCK_K * 128, 512, 1u);
  return make_desc(tile_base + k * (UMMA_K * 4), 0, 1024);
}
// Instruction descriptor, kind::tf32: D=f32 [4,6)=1, A=tf32 [7,10)=2, B=tf32 [10,13)=2, majors, N>>3, M>>4.
template <bool A_MN, bool B_MN, int BN>
__device__ __forceinline__ constexpr uint32_t tf32_idesc() {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
         ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BLOCK_M >> 4) << 24);
}

template <int BN>
struct TCfg {
  static constexpr uint32_t A_BYTES = BLOCK_M * BLOCK_K * 4;  // 16 KB (one of hi / lo)
  static constexpr uint32_t B_BYTES = BN * BLOCK_K * 4;
  // A ring of stages [A raw = A hi][B raw = B hi][B lo] and a 2-slot side ring for A lo.  The raw TMA tiles are the
  // hi operands as they are.  B hi and B lo are adjacent so that ONE MMA with N = 2 BN computes A_hi.[B_hi | B_lo]
  // into the adjacent main | correction accumulator columns (8 instead of 12 MMAs per k-block, a third fewer A
  // reads from shared memory).  An A lo slot lives from the conversion of a k-block to the retirement of its MMAs.
  static constexpr uint32_t STAGE_BYTES = A_BYTES + 2 * B_BYTES;
  static constexpr int STAGES = BN == 128 ? 3 : (BN == 64 ? 5 : 7);
  static constexpr int LO_SLOTS = 2;                          // [A lo] each
  // The tensor core adds into its fp32 accumulator with truncation: a BIASED error that grows with the number of
  // accumulation steps and survives in sums whose terms cancel (bias gradients came out 1e-5 off on the 4096-row
  // MNIST step, against 3e-6 with FFMA).  Forward / dX tiles therefore rotate their k-blocks over NACC
  // (main | correction) accumulator sets, summed in fp32 by the epilogue: chains 1 / NACC as long.  512 columns
  // = 2 tile stages x NACC sets x 2 BN for every BN.  dW tiles are split-K (short chains) and use one set; their
  // second stage holds the bias-gradient accumulators.
  static constexpr int NACC_FWD = BN == 32 ? 4 : (BN == 64 ? 2 : 1);
  static constexpr uint32_t TMEM_COLS = 512;
  static constexpr uint32_t STG_PITCH = 36;
  static constexpr uint32_t STG_BYTES = 4 * 32 * STG_PITCH * 4;
  static constexpr uint32_t BAR_BYTES = ((3 * STAGES + LO_SLOTS + 4) * 8 + 16 + 15) / 16 * 16;  // keeps stg 16-B aligned
  static constexpr uint32_t ONES_BYTES = 1024;  // all-ones fp32 A operand of the fused bias-gradient MMA (one k-step)
  static constexpr size_t SMEM_BYTES =
      1024 + (size_t)STAGES * STAGE_BYTES + (size_t)LO_SLOTS * A_BYTES + BAR_BYTES + STG_BYTES + 128 + ONES_BYTES;
};

// Epilogue of one 128 x BN tile for the warp owning TMEM lanes [32q, 32q+32): main + correction accumulators,
// 32x32 transpose through shared memory, then lane = (row r8 = lane/4, 8 consecutive columns) feeds the functor
// with two 4-column groups per row (N % 4 == 0: a group is all in or all out).
template <int BN, typename Epi>
__device__ __forceinline__ void tf32_epilogue(const Epi& epi, int M, int N, float* stg, uint32_t tmem_main,
                                              uint32_t tmem_corr, int nsets, int row0, int n0, int64_t zoff, int q,
                                              int lane) {
  const int nrows = min(32, M - row0);  // warp-uniform
  const int cg = lane & 3, r8 = lane >> 2;
  const uint32_t lane_off = (uint32_t)(q * 32) << 16;
#pragma unroll 1
  for (int c = 0; c < BN / 32; ++c) {
    float v[32], w[32];
    tmem_ld32(tmem_main + c * 32 + lane_off, v);
    tmem_ld32(tmem_corr + c * 32 + lane_off, w);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] += w[j];
    for (int sset = 1; sset < nsets; ++sset) {  // the other accumulator sets of this tile (2 BN columns apart)
      float u[32];
      tmem_ld32(tmem_main + sset * 2 * BN + c * 32 + lane_off, u);
      tmem_ld32(tmem_corr + sset * 2 * BN + c * 32 + lane_off, w);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] += u[j] + w[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<float4*>(&stg[lane * TCfg<BN>::STG_PITCH + 4 * j]) =
          make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    __syncwarp();
    const int col = n0 + c * 32 + cg * 8;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int r = it * 8 + r8;
      if (r >= nrows) continue;
      const float4 x0 = *reinterpret_cast<const float4*>(&stg[r * TCfg<BN>::STG_PITCH + cg * 8]);
      const float4 x1 = *reinterpret_cast<const float4*>(&stg[r * TCfg<BN>::STG_PITCH + cg * 8 + 4]);
      if (col < N) epi.apply4((int64_t)row0 + r + zoff, col, x0, 4, true);
      if (col + 4 < N) epi.apply4((int64_t)row0 + r + zoff, col + 4, x1, 4, true);
    }
    __syncwarp();
  }
}

template <bool A_MN, bool B_MN, int BN, typename Epi, bool CL = false>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tf32_gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                 const Tf32Params p, const Epi epi) {
  using C = TCfg<BN>;
  constexpr bool DW = A_MN && B_MN;
  // accumulator sets per tile (see TCfg); split-K tiles (dW, cluster split-K) have short chains and use one
  constexpr int NACC = (DW || CL) ? 1 : C::NACC_FWD;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t tiles = (raw + 1023u) & ~1023u;
  const uint32_t lo_ring = tiles + C::STAGES * C::STAGE_BYTES;
  const uint32_t bars = lo_ring + C::LO_SLOTS * C::A_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };                      // TMA bytes landed
  auto conv_bar = [&](int s) { return bars + 8u * (C::STAGES + s); };        // hi / lo written (all converter warps)
  auto empty_bar = [&](int s) { return bars + 8u * (2 * C::STAGES + s); };   // MMAs of the stage retired
  auto lofree_bar = [&](int u) { return bars + 8u * (3 * C::STAGES + u); };  // MMAs that read lo slot u retired
  auto tfull_bar = [&](int a) { return bars + 8u * (3 * C::STAGES + C::LO_SLOTS + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (3 * C::STAGES + C::LO_SLOTS + 2 + a); };
  const uint32_t tmem_slot = bars + 8u * (3 * C::STAGES + C::LO_SLOTS + 4);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = p.m_blocks * p.n_blocks * p.splits;
  // CL: one (tile, split) per CTA; the CTAs of a cluster (consecutive blockIdx.x) share the tile and differ in the split
  const uint32_t cl_rank = CL ? cluster_ctarank() : 0u;
  const int tile_first = CL ? (int)cl_rank * (p.m_blocks * p.n_blocks) + (int)(blockIdx.x / p.splits) : (int)blockIdx.x;
  const int tile_step = CL ? total_tiles : (int)gridDim.x;
  pdl_launch_dependents();  // the next kernel of the stream may set up while this one runs

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(conv_bar(s), CONV_THREADS / 32);
      mbar_init(empty_bar(s), 1);
    }
    for (int u = 0; u < C::LO_SLOTS; ++u) mbar_init(lofree_bar(u), 1);
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, C::TMEM_COLS);
    tmem_relinquish();
  }
  // Bias gradient of the split-K dW GEMM (A_MN && B_MN, one tile per CTA): db[n] = sum_r dZ[r][n] = (ones . dZ)[0][n]
  // from the dZ tiles already in shared memory.  All-ones is layout-invariant, so 1 KB (one k-step of one 32-wide
  // MN atom) stands for the whole 128 x 32 operand: LBO = 0 aliases the MN atoms, the descriptor never advances.
  const uint32_t ones_base = (bars + C::BAR_BYTES + C::STG_BYTES + 127u) & ~127u;
  const bool fuse_db = A_MN && B_MN && (CL || p.db_part != nullptr);
  if (fuse_db && warp == 3) {
    float* o = reinterpret_cast<float*>(smem_raw + (ones_base - raw));
    for (int i = lane; i < (int)(C::ONES_BYTES / 4); i += 32) o[i] = 1.0f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  pdl_wait();  // barriers, TMEM and descriptors are ready: now wait for the producer of our operands

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int tile = tile_first; tile < total_tiles; tile += tile_step) {
        const TileCoord t = decode_tile(tile, p);
        for (int kb = t.kb0; kb < t.kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          mbar_expect_tx(full_bar(stage), C::A_BYTES + C::B_BYTES);
          const uint32_t a_dst = tiles + stage * C::STAGE_BYTES;
          const uint32_t b_dst = a_dst + C::A_BYTES;
          if (!A_MN) {
            tma_load_2d(a_dst, &tma_a, full_bar(stage), kb * BLOCK_K, t.m_blk * BLOCK_M);
          } else {
#pragma unroll
            for (int i = 0; i < BLOCK_M / 32; ++i)
              tma_load_2d(a_dst + i * (BLOCK_K * 128), &tma_a, full_bar(stage), t.m_blk * BLOCK_M + i * 32,
                          kb * BLOCK_K);
          }
          if (!B_MN) {
            tma_load_2d(b_dst, &tma_b, full_bar(stage), kb * BLOCK_K, t.n_blk * BN);
          } else {
#pragma unroll
            for (int i = 0; i < BN / 32; ++i)
              tma_load_2d(b_dst + i * (BLOCK_K * 128), &tma_b, full_bar(stage), t.n_blk * BN + i * 32, kb * BLOCK_K);
          }
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = tf32_idesc<A_MN, B_MN, BN>(), idesc2 = tf32_idesc<A_MN, B_MN, 2 * BN>();
      constexpr uint32_t A_STEP = A_MN ? UMMA_K * 128 : UMMA_K * 4, B_STEP = B_MN ? UMMA_K * 128 : UMMA_K * 4;
      const uint64_t desc_a = tf32_desc<A_MN>(tiles, 0);                    // A hi slot of stage 0, k-step 0
      const uint64_t desc_b = tf32_desc<B_MN>(tiles + C::A_BYTES, 0);       // B hi slot of stage 0
      const uint64_t desc_al = tf32_desc<A_MN>(lo_ring, 0);                 // A lo of slot 0
      const uint64_t ones_desc = make_desc_layout(ones_base, 0, 512, 1u);
      uint32_t stage = 0, phase = 0, as = 0, aphase = 0, u = 0;
      for (int tile = tile_first; tile < total_tiles; tile += tile_step) {
        const TileCoord t = decode_tile(tile, p);
        mbar_wait(tempty_bar(as), aphase ^ 1);
        tc_fence_after();
        const uint32_t d_stage = tmem_base + as * (NACC * 2 * BN);
        for (int kb = t.kb0; kb < t.kb1; ++kb) {
          mbar_wait(conv_bar(stage), phase);  // hi / lo of this stage are in place
          tc_fence_after();
          // all descriptors of the ring differ only in the 16-byte-granular start-address field
          const uint64_t so = (uint64_t)((stage * C::STAGE_BYTES) >> 4), lo = (uint64_t)((u * C::A_BYTES) >> 4);
          const uint64_t ah = desc_a + so, al = desc_al + lo;
          const uint64_t bh = desc_b + so;  // B hi; B lo follows it: N = 2 BN covers both
          const int kj = kb - t.kb0;
          const uint32_t d_main = d_stage + (uint32_t)(kj % NACC) * (2 * BN), d_corr = d_main + BN;
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            const uint32_t acc = (kj >= NACC || k > 0) ? 1u : 0u;  // the first k-block of each set overwrites
            const uint64_t ka = (uint64_t)k * (A_STEP >> 4), kbo = (uint64_t)k * (B_STEP >> 4);
            umma_tf32(d_main, ah + ka, bh + kbo, idesc2, acc);  // main | corr (+)= A_hi . [B_hi | B_lo]
            umma_tf32(d_corr, al + ka, bh + kbo, idesc, 1u);    // corr += A_lo . B_hi
            if (fuse_db && t.m_blk == 0)  // bias gradient: ones . [dZ_hi | dZ_lo] into the idle second accumulator stage
              umma_tf32(tmem_base + 2 * BN, ones_desc, bh + kbo, idesc2, (kj > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(empty_bar(stage));
          umma_commit(lofree_bar(u));
          u ^= 1;
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(tfull_bar(as));
        as ^= 1;
        if (as == 0) aphase ^= 1;
      }
    }
    __syncwarp();
  } else if (warp >= 8) {
    // ===================== converters: lo = x - trunc_tf32(x) from the raw stage bytes =====================
    const int ct = threadIdx.x - 256;  // 0..CONV_THREADS-1
    constexpr int A_CH = C::A_BYTES / 16, B_CH = C::B_BYTES / 16;
    constexpr int PER = (A_CH + B_CH + CONV_THREADS - 1) / CONV_THREADS;
    static_assert(A_CH % CONV_THREADS == 0, "a thread's j-th chunk is in A or in B for all threads alike");
    uint32_t stage = 0, phase = 0, u = 0, uphase = 0;
    for (int tile = tile_first; tile < total_tiles; tile += tile_step) {
      const TileCoord t = decode_tile(tile, p);
      for (int kb = t.kb0; kb < t.kb1; ++kb) {
        mbar_wait(full_bar(stage), phase);
        mbar_wait(lofree_bar(u), uphase ^ 1);  // the MMAs of the k-block that last used this lo slot have retired
        uint8_t* base = smem_raw + (tiles + stage * C::STAGE_BYTES - raw);
        uint8_t* alo = smem_raw + (lo_ring + u * C::A_BYTES - raw);  // A lo slot; B lo sits behind B raw in the stage
        float4 x[PER];
        // all loads first (PER independent 16-byte loads in flight), then the split, then the stores
#pragma unroll
        for (int j = 0; j < PER; ++j) {
          const int i = ct + j * CONV_THREADS;
          const bool is_a = j * CONV_THREADS < A_CH;
          if (is_a || i < A_CH + B_CH)
            x[j] = *reinterpret_cast<const float4*>(base + i * 16);  // [A][B] are contiguous in a stage
        }
#pragma unroll
        for (int j = 0; j < PER; ++j) {
          const int i = ct + j * CONV_THREADS;
          const bool is_a = j * CONV_THREADS < A_CH;
          if (is_a || i < A_CH + B_CH) {
            float4 h, l;
            h.x = trunc_tf32(x[j].x); h.y = trunc_tf32(x[j].y); h.z = trunc_tf32(x[j].z); h.w = trunc_tf32(x[j].w);
            // lo has up to 13 significant bits and the tensor core will truncate it to 10: round it here (to
            // nearest), otherwise every lo is cut towards zero -- a bias that survives in sums whose terms
            // cancel (bias gradients: 1e-5 on the 4096-row MNIST step)
            l.x = rn_tf32(x[j].x - h.x); l.y = rn_tf32(x[j].y - h.y); l.z = rn_tf32(x[j].z - h.z); l.w = rn_tf32(x[j].w - h.w);
            // the raw tile stays in place as the hi operand
            *reinterpret_cast<float4*>(is_a ? alo + i * 16 : base + C::B_BYTES + i * 16) = l;
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to tcgen05.mma
        __syncwarp();
        if (lane == 0) mbar_arrive(conv_bar(stage));
        u ^= 1;
        if (u == 0) uphase ^= 1;
        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int q = warp & 3;
    float* stg = reinterpret_cast<float*>(smem_raw + (bars + C::BAR_BYTES - raw)) + q * 32 * C::STG_PITCH;
    uint32_t as = 0, aphase = 0;
    for (int tile = tile_first; tile < total_tiles; tile += tile_step) {
      const TileCoord t = decode_tile(tile, p);
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const uint32_t d_main = tmem_base + as * (NACC * 2 * BN);
      if (CL) {
        // the partial tile (main + correction) goes to shared memory, over the operand ring: every MMA that read the
        // ring has retired (tfull), and this CTA loads nothing more
        float* part = reinterpret_cast<float*>(smem_raw + (tiles - raw));
        constexpr int PITCH = BN + CL_PITCH_PAD;
        const int row = q * 32 + lane;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          float v[32], w[32];
          tmem_ld32(d_main + c * 32 + ((uint32_t)(q * 32) << 16), v);
          tmem_ld32(d_main + BN + c * 32 + ((uint32_t)(q * 32) << 16), w);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(&part[row * PITCH + c * 32 + 4 * j]) =
                make_float4(v[4 * j] + w[4 * j], v[4 * j + 1] + w[4 * j + 1], v[4 * j + 2] + w[4 * j + 2],
                            v[4 * j + 3] + w[4 * j + 3]);
        }
        if (DW && t.m_blk == 0 && q == 0) {  // bias-gradient partial: TMEM lane 0 of the ones-accumulators
          float* dbp = part + BLOCK_M * PITCH;
#pragma unroll 1
          for (int c = 0; c < BN / 32; ++c) {
            float v[32], w[32];
            tmem_ld32(tmem_base + 2 * BN + c * 32, v);
            tmem_ld32(tmem_base + 3 * BN + c * 32, w);
            tmem_ld_wait();
            if (lane == 0) {
#pragma unroll
              for (int j = 0; j < 32; ++j) dbp[c * 32 + j] = v[j] + w[j];
            }
          }
        }
      } else {
      tf32_epilogue<BN, Epi>(epi, p.M, p.N, stg, d_main, d_main + BN, min(NACC, t.kb1 - t.kb0),
                             t.m_blk * BLOCK_M + q * 32, t.n_blk * BN,
                             (int64_t)t.ks * p.zrows, q, lane);
      }
      if (!CL && fuse_db && t.m_blk == 0 && q == 0) {
        // every row of the ones-accumulators holds the column sums; TMEM lane 0 belongs to lane 0 of this warp
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          float v[32], w[32];
          tmem_ld32(tmem_base + 2 * BN + c * 32, v);
          tmem_ld32(tmem_base + 3 * BN + c * 32, w);
          tmem_ld_wait();
          const int col = t.n_blk * BN + c * 32;
          if (lane == 0) {
            float* o = p.db_part + (int64_t)t.ks * p.db_np + col;
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (col + 4 * j < p.N)  // N % 4 == 0
                *reinterpret_cast<float4*>(o + 4 * j) = make_float4(v[4 * j] + w[4 * j], v[4 * j + 1] + w[4 * j + 1],
                                                                    v[4 * j + 2] + w[4 * j + 2], v[4 * j + 3] + w[4 * j + 3]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(as));
      as ^= 1;
      if (as == 0) aphase ^= 1;
    }
  }

  tc_fence_before();
  if (CL) {
    // ---- split-K reduction across the cluster through distributed shared memory, then the epilogue ----
    cluster_sync_all();  // every CTA's partial tile is in its shared memory
    constexpr int PITCH = BN + CL_PITCH_PAD;
    const int S = p.splits;
    const TileCoord t = decode_tile(tile_first, p);
    const uint32_t part_addr = tiles;  // same offset in every CTA of the cluster
    const int rows_per_cta = BLOCK_M / S;  // S divides 128
    const int q4 = BN / 4;
    for (int idx = threadIdx.x; idx < rows_per_cta * q4; idx += NUM_THREADS) {
      const int r = (int)cl_rank * rows_per_cta + idx / q4, c4 = idx % q4;
      const int row = t.m_blk * BLOCK_M + r, col = t.n_blk * BN + c4 * 4;
      if (row >= p.M || col >= p.N) continue;  // N % 4 == 0: a group of 4 columns is all in or all out
      // the weight this thread will update: its global load goes out first, under the remote loads
      float* dst = DW ? p.cl_w + (int64_t)row * p.N + col : nullptr;
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      if (DW && p.cl_apply) w = *reinterpret_cast<const float4*>(dst);
      // all S <= 16 remote loads in flight at once, then summed in CTA order (deterministic)
      float4 v[16];
#pragma unroll
      for (int u = 0; u < 16; ++u)
        v[u] = (u < S) ? ld_dsmem4(part_addr + (uint32_t)(r * PITCH + c4 * 4) * 4u, (uint32_t)u)
                       : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 s = v[0];
#pragma unroll
      for (int u = 1; u < 16; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
      if (!DW) {  // forward / dX: the layout's own epilogue (bias + activation, activation derivative) on the sum
        epi.apply4((int64_t)row, col, s, 4, true);
        continue;
      }
      if (p.cl_apply) {
        w.x -= p.cl_lr * s.x; w.y -= p.cl_lr * s.y; w.z -= p.cl_lr * s.z; w.w -= p.cl_lr * s.w;
        s = w;
      }
      *reinterpret_cast<float4*>(dst) = s;
    }
    if (DW && t.m_blk == 0 && cl_rank == 0) {  // bias gradient of this column block
      for (int c4 = threadIdx.x; c4 < q4; c4 += NUM_THREADS) {
        const int col = t.n_blk * BN + c4 * 4;
        if (col >= p.N) continue;
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int src = 0; src < S; ++src) {
          const float4 v = ld_dsmem4(part_addr + (uint32_t)(BLOCK_M * PITCH + c4 * 4) * 4u, (uint32_t)src);
          s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        float* dst = p.cl_b + col;
        if (p.cl_apply) {
          float4 b = *reinterpret_cast<const float4*>(dst);
          b.x -= p.cl_lr * s.x; b.y -= p.cl_lr * s.y; b.z -= p.cl_lr * s.z; b.w -= p.cl_lr * s.w;
          s = b;
        }
        *reinterpret_cast<float4*>(dst) = s;
      }
    }
    cluster_sync_all();  // nobody leaves while a neighbour still reads its shared memory
  } else {
    __syncthreads();
  }
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// ---- host side ------------------------------------------------------------------------
int ctas_available() {
  int n = umma_num_sms() - umma_sm_reserve();
  return n < 1 ? 1 : n;
}

int operand_maps(bool a_mn, const float* A, bool b_mn, const float* Bm, int64_t M, int64_t N, int64_t Kred, int bn,
                 CUtensorMap* ta, CUtensorMap* tb) {
  int rc = a_mn ? umma_get_tmap(A, 4, M, Kred, 32, 32, 1, ta) : umma_get_tmap(A, 4, Kred, M, 32, BLOCK_M, 0, ta);
  if (rc != DDNN_OK) return rc;
  return b_mn ? umma_get_tmap(Bm, 4, N, Kred, 32, 32, 1, tb) : umma_get_tmap(Bm, 4, Kred, N, 32, bn, 0, tb);
}

void fill_tiles(Tf32Params& p, int64_t M, int64_t N, int64_t Kred, int bn, int splits) {
  p.M = (int)M;
  p.N = (int)N;
  p.m_blocks = (int)cdiv(M, BLOCK_M);
  p.n_blocks = (int)cdiv(N, bn);
  p.num_kb = (int)cdiv(Kred, BLOCK_K);
  if (splits < 1) splits = 1;
  p.kb_per_split = (int)cdiv(p.num_kb, splits);
  p.splits = (int)cdiv(p.num_kb, p.kb_per_split);
  p.group_m = p.m_blocks;
  p.zrows = 0;
  p.db_part = nullptr;
  p.db_np = 0;
}

// output-tile width: the widest BN that still gives (nearly) every SM a tile; these GEMMs are small, so
// parallelism across CTAs matters more than operand reuse inside one
int pick_bn(int64_t M, int64_t N) {
  const long mb = cdiv(M, BLOCK_M);
  const long want = (long)ctas_available() * 4 / 5;
  if (N >= 128 && mb * cdiv(N, 128) >= want) return 128;
  if (N >= 64 && mb * cdiv(N, 64) >= want) return 64;
  return 32;
}

template <bool A_MN, bool B_MN, int BN, typename Epi>
int launch_one(const CUtensorMap& ta, const CUtensorMap& tb, const Tf32Params& p, const Epi& epi, cudaStream_t st) {
  auto kern = tf32_gemm_kernel<A_MN, B_MN, BN, Epi>;
  if (int rc = ensure_dynamic_smem((const void*)kern, TCfg<BN>::SMEM_BYTES)) return rc;
  const int total = p.m_blocks * p.n_blocks * p.splits;
  const int ctas = ctas_available();
  DDNN_CUDA_CHECK(launch_pdl(kern, dim3(total < ctas ? total : ctas), dim3(NUM_THREADS), TCfg<BN>::SMEM_BYTES, st, ta,
                             tb, p, epi));
  count_launch();
  return DDNN_OK;
}
template <bool A_MN, bool B_MN, typename Epi>
int launch_bn(int bn, const CUtensorMap& ta, const CUtensorMap& tb, const Tf32Params& p, const Epi& epi,
              cudaStream_t st) {
  if (bn == 128) return launch_one<A_MN, B_MN, 128, Epi>(ta, tb, p, epi, st);
  if (bn == 64) return launch_one<A_MN, B_MN, 64, Epi>(ta, tb, p, epi, st);
  return launch_one<A_MN, B_MN, 32, Epi>(ta, tb, p, epi, st);
}

// Cluster split-K plan: S CTAs per output tile, S in {16, 8, 4, 2} (S divides the 128 tile rows).  Wanted: one wave
// (tiles * S <= SMs) that uses at least `min_ctas` SMs, >= 2 k-blocks per split, S >= `min_s` (accumulation chains
// within DDNN_TF32_MAX_CHAIN), all clusters co-resident (a B200 holds 7 clusters of more than 8 CTAs).  Returns 0
// when no S fits (the caller keeps its other plan).
template <bool A_MN, bool B_MN, int BN, typename Epi>
int cluster_plan(long tiles, long num_kb, long min_s, long min_ctas) {
  auto kern = tf32_gemm_kernel<A_MN, B_MN, BN, Epi, true>;
  if (ensure_dynamic_smem((const void*)kern, TCfg<BN>::SMEM_BYTES) != DDNN_OK) return 0;
  static const int dbg_max_s = getenv("DDNN_DBG_CLUSTER_S") ? atoi(getenv("DDNN_DBG_CLUSTER_S")) : 16;  // measurement knob
  for (int S : {16, 8, 4, 2}) {
    if (S > dbg_max_s) continue;
    if (tiles * S > ctas_available() || tiles * S < min_ctas || num_kb / S < 2 || S < min_s) continue;
    if (cdiv(num_kb, cdiv(num_kb, S)) != S) continue;  // the k-blocks do not split S ways
    if (S > 8 && cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
      cudaGetLastError();
      continue;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(tiles * S));
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = TCfg<BN>::SMEM_BYTES;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = S;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int max_clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&max_clusters, (const void*)kern, &cfg) != cudaSuccess) {
      cudaGetLastError();
      continue;
    }
    if (max_clusters >= tiles) return S;  // all clusters co-resident: one wave
  }
  return 0;
}

template <bool A_MN, bool B_MN, int BN, typename Epi>
int launch_cluster(const CUtensorMap& ta, const CUtensorMap& tb, const Tf32Params& p, const Epi& epi, cudaStream_t st) {
  auto kern = tf32_gemm_kernel<A_MN, B_MN, BN, Epi, true>;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)(p.m_blocks * p.n_blocks * p.splits));
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = TCfg<BN>::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = p.splits;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_pdl.load(std::memory_order_relaxed) ? 2 : 1;
  DDNN_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, p, epi));
  count_launch();
  return DDNN_OK;
}

// plan look-up once per (device, shape): the occupancy query is a driver call
struct PlanCache {
  std::mutex mu;
  std::map<std::tuple<int, int64_t, int64_t, int64_t>, int> plans;
  template <typename F>
  int get(int64_t B, int64_t K, int64_t N, F&& compute) {
    int dev = 0;
    cudaGetDevice(&dev);
    const auto key = std::make_tuple(dev, B, K, N);
    {
      std::lock_guard<std::mutex> g(mu);
      auto it = plans.find(key);
      if (it != plans.end()) return it->second;
    }
    const int S = compute();
    std::lock_guard<std::mutex> g(mu);
    if (plans.size() > 4096) plans.clear();
    plans[key] = S;
    return S;
  }
};

// Forward / dX with few output tiles and a long reduction (MNIST layer 1: 32 tiles of 128 x 128, 25 k-blocks): the
// plain plan shrinks the tile width until ~every SM has a tile and walks all k-blocks serially in each CTA, which is
// what the launch is bound by (25 x ~0.5 us).  With the widest tile and the reduction split over a cluster, each
// CTA walks 1/S of the k-blocks; the partial tiles are summed through distributed shared memory and the layout's
// epilogue (bias + activation / activation derivative) runs on the sum.  *taken = 0: no plan, run the plain kernel.
template <bool A_MN, bool B_MN, typename Epi>
int forward_like_cluster(PlanCache& cache, const float* A, const float* Bm, int64_t M, int64_t N, int64_t Kred,
                         const Epi& epi, int* taken, cudaStream_t st) {
  *taken = 0;
  const int bn = N >= 128 ? 128 : (N >= 64 ? 64 : 32);
  const long tiles = (long)cdiv(M, BLOCK_M) * cdiv(N, bn), num_kb = cdiv(Kred, BLOCK_K);
  if (num_kb < 8 || tiles * 2 > ctas_available()) return DDNN_OK;
  const long min_ctas = (long)ctas_available() * 3 / 5;
  const int S = cache.get(M, Kred, N, [&] {
    return bn == 128 ? cluster_plan<A_MN, B_MN, 128, Epi>(tiles, num_kb, 1, min_ctas)
                     : (bn == 64 ? cluster_plan<A_MN, B_MN, 64, Epi>(tiles, num_kb, 1, min_ctas)
                                 : cluster_plan<A_MN, B_MN, 32, Epi>(tiles, num_kb, 1, min_ctas));
  });
  if (S == 0) return DDNN_OK;
  CUtensorMap ta, tb;
  int rc = operand_maps(A_MN, A, B_MN, Bm, M, N, Kred, bn, &ta, &tb);
  if (rc != DDNN_OK) return rc;
  Tf32Params p{};
  fill_tiles(p, M, N, Kred, bn, S);
  if (p.splits != S) return DDNN_OK;
  *taken = 1;
  if (bn == 128) return launch_cluster<A_MN, B_MN, 128, Epi>(ta, tb, p, epi, st);
  if (bn == 64) return launch_cluster<A_MN, B_MN, 64, Epi>(ta, tb, p, epi, st);
  return launch_cluster<A_MN, B_MN, 32, Epi>(ta, tb, p, epi, st);
}

std::atomic<int> g_cluster_dw{1};

}  // namespace

void tf32_set_cluster_dw(int on) { g_cluster_dw.store(on ? 1 : 0); }

static int dw_bn(int64_t N);

// dW + db with the split-K reduction inside the kernel (thread-block cluster, distributed shared memory): writes the
// final result -- SGD on W_master / bias (apply) or the fp32 gradients -- in ONE launch, no workspace traffic.
// *taken = 0 when the shape has no cluster plan (the caller then runs the workspace path).
int tf32_dw_cluster(const float* X, const float* dZ, float* W_master, float* bias, float* dW_out, float* db_out, float lr,
                    int apply, int64_t B, int64_t K, int64_t N, int* taken, cudaStream_t st) {
  *taken = 0;
  if (!g_cluster_dw.load()) return DDNN_OK;
  float* w = apply ? W_master : dW_out;
  float* b = apply ? bias : db_out;
  if ((reinterpret_cast<uintptr_t>(w) & 15u) || (reinterpret_cast<uintptr_t>(b) & 15u)) return DDNN_OK;
  const int bn = dw_bn(N);
  static_assert(BLOCK_M * (128 + CL_PITCH_PAD) * 4 + 128 * 4 <= TCfg<128>::STAGES * (int)TCfg<128>::STAGE_BYTES &&
                BLOCK_M * (64 + CL_PITCH_PAD) * 4 + 64 * 4 <= TCfg<64>::STAGES * (int)TCfg<64>::STAGE_BYTES &&
                BLOCK_M * (32 + CL_PITCH_PAD) * 4 + 32 * 4 <= TCfg<32>::STAGES * (int)TCfg<32>::STAGE_BYTES,
                "the partial tile must fit in the operand ring it overlays");
  static PlanCache cache;
  const long tiles = (long)cdiv(K, BLOCK_M) * cdiv(N, bn), num_kb = cdiv(B, BLOCK_K);
  const long min_s = cdiv(B, DDNN_TF32_MAX_CHAIN);
  const int S = cache.get(B, K, N, [&] {
    return bn == 128 ? cluster_plan<true, true, 128, EpiStoreF32>(tiles, num_kb, min_s, 0)
                     : (bn == 64 ? cluster_plan<true, true, 64, EpiStoreF32>(tiles, num_kb, min_s, 0)
                                 : cluster_plan<true, true, 32, EpiStoreF32>(tiles, num_kb, min_s, 0));
  });
  if (S == 0) return DDNN_OK;
  CUtensorMap ta, tb;
  int rc = operand_maps(true, X, true, dZ, K, N, B, bn, &ta, &tb);
  if (rc != DDNN_OK) return rc;
  Tf32Params p{};
  fill_tiles(p, K, N, B, bn, S);
  if (p.splits != S) return DDNN_OK;  // the k-blocks do not split S ways: leave it to the workspace path
  p.cl_w = w;
  p.cl_b = b;
  p.cl_lr = lr;
  p.cl_apply = apply ? 1 : 0;
  *taken = 1;
  const EpiStoreF32 unused{nullptr, 0};
  if (bn == 128) return launch_cluster<true, true, 128, EpiStoreF32>(ta, tb, p, unused, st);
  if (bn == 64) return launch_cluster<true, true, 64, EpiStoreF32>(ta, tb, p, unused, st);
  return launch_cluster<true, true, 32, EpiStoreF32>(ta, tb, p, unused, st);
}

// TMA needs 16-byte row pitches (widths % 4) and the boxes are 32 elements wide
bool tf32_shape_ok(int64_t B, int64_t K, int64_t N) {
  return B > 0 && K >= 32 && N >= 32 && K % 4 == 0 && N % 4 == 0;
}

// longest un-split reduction the forward / dX GEMM with an (M x N) output accepts: DDNN_TF32_MAX_CHAIN per
// accumulator set of the tile width pick_bn chooses
int64_t tf32_max_reduction(int64_t M, int64_t N) {
  const int bn = pick_bn(M, N);
  return DDNN_TF32_MAX_CHAIN * (bn == 32 ? TCfg<32>::NACC_FWD : (bn == 64 ? TCfg<64>::NACC_FWD : TCfg<128>::NACC_FWD));
}

int tf32_forward(const float* X, const float* W, const float* bias, float* A, int64_t B, int64_t K, int64_t N, int act,
                 cudaStream_t st) {
  // D[B, N] = X[B, K] . W[K, N]: A operand K-major, B operand MN-major (W is [K][N] like grpc_node.py:51)
  if (g_cluster_dw.load()) {
    static PlanCache cache;
    int taken = 0;
    EpiFwd<float> epi{bias, A, N, act};
    int rc = forward_like_cluster<false, true, EpiFwd<float>>(cache, X, W, B, N, K, epi, &taken, st);
    if (rc || taken) return rc;
  }
  const int bn = pick_bn(B, N);
  CUtensorMap ta, tb;
  int rc = operand_maps(false, X, true, W, B, N, K, bn, &ta, &tb);
  if (rc != DDNN_OK) return rc;
  Tf32Params p{};
  fill_tiles(p, B, N, K, bn, 1);
  EpiFwd<float> epi{bias, A, N, act};
  return launch_bn<false, true, EpiFwd<float>>(bn, ta, tb, p, epi, st);
}

int tf32_dx(const float* dZ, const float* W, const float* Aprev, float* dZprev, int64_t B, int64_t K, int64_t N,
            int act_prev, cudaStream_t st) {
  // D[B, K] = dZ[B, N] . W[K, N]^T: both operands K-major (the reduction runs along W's rows)
  if (g_cluster_dw.load()) {
    static PlanCache cache;
    int taken = 0;
    EpiDx<float> epi{Aprev, dZprev, K, act_prev};
    int rc = forward_like_cluster<false, false, EpiDx<float>>(cache, dZ, W, B, K, N, epi, &taken, st);
    if (rc || taken) return rc;
  }
  const int bn = pick_bn(B, K);
  CUtensorMap ta, tb;
  int rc = operand_maps(false, dZ, false, W, B, K, N, bn, &ta, &tb);
  if (rc != DDNN_OK) return rc;
  Tf32Params p{};
  fill_tiles(p, B, K, N, bn, 1);
  EpiDx<float> epi{Aprev, dZprev, K, act_prev};
  return launch_bn<false, false, EpiDx<float>>(bn, ta, tb, p, epi, st);
}

static int dw_bn(int64_t N) {
  static const int dbg = getenv("DDNN_DBG_DW_BN") ? atoi(getenv("DDNN_DBG_DW_BN")) : 0;  // measurement knob
  if (dbg == 128 || dbg == 64 || dbg == 32) return N >= dbg ? dbg : 32;
  return N >= 128 ? 128 : (N >= 64 ? 64 : 32);
}

int tf32_dw_splits(int64_t B, int64_t K, int64_t N) {
  const int bn = dw_bn(N);
  const long tiles = (long)cdiv(K, BLOCK_M) * cdiv(N, bn);
  const long num_kb = cdiv(B, BLOCK_K);
  long want = umma_num_sms() / tiles;  // one wave
  long max_by_k = num_kb / 2;          // >= 2 k-blocks per split
  if (max_by_k < 1) max_by_k = 1;
  long s = want < max_by_k ? want : max_by_k;
  const long min_by_chain = cdiv(B, DDNN_TF32_MAX_CHAIN);  // accuracy first: bounded accumulation chains
  if (s < min_by_chain) s = min_by_chain;
  if (s < 1) s = 1;
  const long per = cdiv(num_kb, s);
  return (int)cdiv(num_kb, per);
}

int tf32_dw_partials(const float* X, const float* dZ, float* dw_part, int splits, int64_t B, int64_t K, int64_t N,
                     float* db_part, int* db_fused, cudaStream_t st) {
  if (db_fused) *db_fused = 0;
  // D[K, N] = X[B, K]^T . dZ[B, N]: both operands MN-major, reduction over the batch; partials [split][K][Np]
  const int bn = dw_bn(N);
  CUtensorMap ta, tb;
  int rc = operand_maps(true, X, true, dZ, K, N, B, bn, &ta, &tb);
  if (rc != DDNN_OK) return rc;
  Tf32Params p{};
  fill_tiles(p, K, N, B, bn, splits);
  if (p.splits != splits) {
    set_error("tf32_dw_partials: split plan mismatch (%d vs %d)", p.splits, splits);
    return DDNN_ERR_ARG;
  }
  p.zrows = K;
  const int64_t np = (N + 3) & ~(int64_t)3;
  // one tile per CTA: the bias gradient rides along on the idle accumulator stage
  if (db_part && db_fused && (long)p.m_blocks * p.n_blocks * p.splits <= (long)ctas_available()) {
    p.db_part = db_part;
    p.db_np = (int)np;
    *db_fused = 1;
  }
  EpiStoreF32 epi{dw_part, np};
  return launch_bn<true, true, EpiStoreF32>(bn, ta, tb, p, epi, st);
}

}  // namespace ddnn
