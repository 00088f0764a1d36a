// Mahalanobis pass 1 on the 5th-generation tensor cores: the Gram statistics as a tcgen05.mma contraction.
//
// Replaces CoreMahalanobis.py:98-100 of the reference (the (nb,p,p) outer-product tensor and its sum) for
// d = 64, 128 and 256: per sub-segment  s = sum(xc - a),  G = sum (xc - a)(xc - a)^T  (see maha_stats.cu).
//
// Data flow per CTA (one per SM, persistent over sub-segments): 16 producer warps + 1 MMA warp.
//   producers: global --LDG.128 (a warp instruction reads 256 contiguous bytes of each of two rows; one or two 32 KB
//   chunks per SM stay in flight ahead of the chunk being converted)--> registers: clip, subtract the start-of-call mean,
//   split every value into a TF32 "hi" part and the TF32-rounded remainder "lo" (integer round-to-nearest, 2 instructions)
//   --STS.128--> shared-memory tiles in the canonical MN-major tf32 layout (SWIZZLE_128B_BASE32B: 128-byte feature slabs,
//   4-row atoms, 32-byte chunks XOR-ed with row & 3; verified on hardware with tools/umma_probe.cu); fence.proxy.async,
//   one mbarrier arrive per warp on the slot's "full" barrier.
//   MMA warp (one lane): waits "full", issues tcgen05.mma.kind::tf32 with A = B = descriptors of the same tile (the Gram
//   is T^T T), fp32 accumulator in TMEM; tcgen05.commit releases the slot ("empty") and, after the last chunk of a
//   sub-segment, signals the producers to drain the accumulator with tcgen05.ld and store the sub-segment's partial.
//   3xTF32 accuracy: hi*hi + hi*lo + lo*hi.
//     d = 64:  lo is stored as two more feature slabs of the hi tile, ONE M = 128, N = 64 MMA per 8 rows gives
//              [hi^T hi; lo^T hi] and the epilogue forms G = HH + LH + LH^T (a third of the tensor work and half of the
//              shared-memory reads of three M = 64 MMAs, which also run the pipe at half rate);
//     d = 128: three M = N = 128 MMAs per 8 rows;
//     d = 256: two accumulators (rows 0-127 x 256 columns, rows 128-255 x columns 128-255: the block below the
//              diagonal is mirrored on store), 384 of the 512 TMEM columns, six MMAs per 8 rows.
//
// d = 64 and d = 256 additionally stage the raw rows in shared memory: a loader warp issues one cp.async.bulk per
// 32 KB chunk into a ring (3 slots at d = 64, 2 at d = 256; mbarrier complete_tx), the producers read their float4 from
// there; the tile ring then has two slots.  (d = 128 was measured faster with register prefetch.)
//
// Bounds (B200, measured, 16Mi rows): d = 64: 1.01 ms = 4.2 TB/s, 65 % of the measured HBM copy rate.  Shared-memory
// bandwidth is the limiter: per 32 KB chunk 32 KB bulk-copy writes + 32 KB raw reads + 64 KB tile writes + 96 KB
// operand reads by the MMAs = 224 KB at 128 B/clk = 1750 clk against 1424 clk of HBM time (without the raw staging the
// kernel moved 160 KB per chunk but stalled on HBM latency: 1.07 ms).  d = 128: 3.4 TB/s.  d = 256: 2.7 TB/s; the operand
// reads of the six MMAs per 8 rows (240 KB per 32 KB chunk) make shared-memory bandwidth the limiter there as well.
#include "maha.cuh"

namespace sb {
namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}

// round to nearest (ties away from zero) to the 10-bit TF32 mantissa with integer arithmetic: what cvt.rna.tf32.f32
// computes for finite inputs, in two instructions (ptxas expands the cvt into about six on sm_100a)
__device__ __forceinline__ uint32_t round_tf32(float x) { return (__float_as_uint(x) + 0x1000u) & 0xffffe000u; }

__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  }
}

// shared-memory matrix descriptor: MN-major tf32 tile, layout SWIZZLE_128B_BASE32B (= 1), descriptor version 1
__device__ __forceinline__ uint64_t umma_desc(uint32_t start, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((start >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3fff) << 32) | (1ull << 46) | (1ull << 61);
}

// instruction descriptor: D = fp32, A = B = tf32, both MN-major, shape M x N (K = 8)
__host__ __device__ constexpr uint32_t umma_idesc(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc),
      "r"(accum) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,"
      "%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}

// d = 64: rows per chunk and raw-ring depth.  Measured (16 Mi rows, round 2): 128 rows / 3 slots 0.94 ms; 64 rows with
// 6 or 8 slots (more, smaller copies in flight) 1.12 / 1.15 ms - the per-chunk barrier traffic costs more than the
// deeper ring gives.
#ifndef SB_P1_ROWS64
#define SB_P1_ROWS64 128
#endif
#ifndef SB_P1_RAW64
#define SB_P1_RAW64 3
#endif
// d = 64 is HBM-latency bound: its rows are staged raw in shared memory by bulk copies (cp.async.bulk, a loader warp;
// three 32 KB slots in flight per SM cost no registers) and the tile ring gives up one slot for that.
__host__ __device__ constexpr bool umma_raw_staged(int D) { return D == 64 || D == 256; }  // measured: d = 128 is faster without
__host__ __device__ constexpr int umma_tile_stages(int D) { return umma_raw_staged(D) ? 2 : 3; }
constexpr int kRawStagesMax = 8;
__host__ __device__ constexpr int umma_raw_stages(int D) { return D == 64 ? SB_P1_RAW64 : 2; }  // d = 256 has 5 KB of constants more
// rows per chunk: 8192 / D, so every tile is 32 KB and every thread converts four float4 per chunk
__host__ __device__ constexpr int umma_chunk_rows(int D) { return D == 64 ? SB_P1_ROWS64 : 8192 / D; }

__host__ __device__ constexpr size_t umma_smem_bytes(int D, int terms) {
  return (size_t)umma_tile_stages(D) * (terms == 3 ? 2 : 1) * umma_chunk_rows(D) * D * 4 +
         (umma_raw_staged(D) ? (size_t)umma_raw_stages(D) * umma_chunk_rows(D) * D * 4 : 0) +
         (D > 128 ? (size_t)5 * D * 4 : 0) /* mean, clip bounds, input transform when they do not fit in registers */ +
         1024 /* alignment slack */;
}

constexpr int kUmmaThreads = 512;            // producer threads (16 warps); one more warp issues the MMAs
constexpr int kUmmaBlock = kUmmaThreads + 32;
__host__ __device__ constexpr int umma_block(int D) { return kUmmaBlock + (umma_raw_staged(D) ? 32 : 0); }  // + loader warp

// STD: the rows are standardised on load, x' = (x - in_mu) * in_is per feature with the two roundings of the stand-alone
// standardise kernel (combiner.cu), so that fusing the transformer into the detector feeds it the same values.
template <int D, bool CLIP, int TERMS, bool STD>
__global__ void __launch_bounds__(umma_block(D), 1)
pass1_gram_umma(const float* __restrict__ X, int64_t ld, SubLayout lay, const float* __restrict__ mean_f,
                const float* __restrict__ lo_f, const float* __restrict__ hi_f, const float* __restrict__ in_mu,
                const float* __restrict__ in_is, float* __restrict__ partials) {
  constexpr int PS = D + D * D;
  constexpr int NSLAB = D / 32;                 // 128-byte feature slabs per row
  constexpr uint32_t LBO = 512;                 // between feature slabs of one 4-row atom
  // Symmetry: hi*lo and lo*hi are transposes of each other, so only  HH = sum hi^T hi  and  LH = sum lo^T hi  are
  // contracted and the epilogue forms  G = HH + LH + LH^T.
  //   d = 64  (STACK): lo is stored as two more feature slabs of the hi tile; ONE M = 128, N = 64 MMA per 8 rows gives
  //                    [HH; LH] in TMEM lanes 0-63 / 64-127 (M = 128 also runs the tensor pipe at full rate, M = 64 at half)
  // (d = 128 with two accumulators HH, LH was measured slower than the plain three products: the longer epilogue is not
  // hidden behind the MMAs of a 900-row sub-segment.)
  constexpr bool STACK = D == 64 && TERMS == 3;
  constexpr bool RAW = umma_raw_staged(D);
  constexpr int kRawStages = umma_raw_stages(D);
  constexpr int kUmmaStages = umma_tile_stages(D);
  constexpr int RAWBYTES = umma_chunk_rows(D) * D * 4;
  constexpr uint32_t SBO = (STACK ? 2 : 1) * NSLAB * 512;  // between 4-row groups
  constexpr int kChunkRows = umma_chunk_rows(D);
  constexpr int NROW = kChunkRows / 32;         // rows per thread and chunk: (tid >> 4) + 32*j
  constexpr int TILE = kChunkRows * D * 4;      // bytes of one tile (32 KB)
  constexpr int NT = TERMS == 3 ? 2 : 1;        // tiles per stage (hi, lo)
  constexpr int NPIECE = D / 64;                // float4 pieces per thread: (tid & 15) + 16*i of row tid >> 4
  constexpr uint32_t LO_OFF = STACK ? NSLAB * LBO : TILE;  // from a hi element to its lo element
  constexpr int TMEM_COLS = D == 256 ? 512 : D;  // power of two >= columns used (D = 256 uses 384)
  extern __shared__ uint8_t umma_smem_raw[];
  uint8_t* tiles = (uint8_t*)(((uintptr_t)umma_smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* raw = tiles + kUmmaStages * NT * TILE;  // [kRawStages][chunk rows][D] fp32, row major (RAW only)
  float* ms = reinterpret_cast<float*>(raw + (RAW ? kRawStages * RAWBYTES : 0));
  float* los = ms + D;
  float* his = los + D;
  __shared__ uint64_t bar_full[kUmmaStages];   // 16 producer warps have filled the slot
  __shared__ uint64_t bar_empty[kUmmaStages];  // the MMAs reading the slot have completed (tcgen05.commit)
  __shared__ uint64_t bar_rfull[kRawStagesMax];   // the bulk copy of a raw chunk has landed
  __shared__ uint64_t bar_rempty[kRawStagesMax];  // 16 producer warps have converted the raw chunk
  __shared__ uint64_t bar_acc;                 // all MMAs of the sub-segment have completed
  __shared__ uint64_t bar_tfree;               // 16 producer warps have drained the accumulator
  __shared__ uint32_t tmem_base_s;
  __shared__ float ssum_s[D];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int krow = tid >> 4, piece0 = tid & 15;  // a warp instruction reads 256 contiguous bytes of each of 2 rows
  constexpr bool HOIST = NPIECE <= 2;  // this thread's mean / clip bounds stay in registers (no shared copy)
  if (!HOIST) {
    for (int f = tid; f < D; f += umma_block(D)) {
      ms[f] = mean_f[f];
      los[f] = CLIP ? lo_f[f] : 0.f;
      his[f] = CLIP ? hi_f[f] : 0.f;
      his[D + f] = STD ? in_mu[f] : 0.f;
      his[2 * D + f] = STD ? in_is[f] : 0.f;
    }
  }
  if (tid == 0) {
    for (int s = 0; s < kUmmaStages; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 16;" ::"r"(smem_u32(&bar_full[s])));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar_empty[s])));
    }
    for (int s = 0; s < kRawStages; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar_rfull[s])));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 16;" ::"r"(smem_u32(&bar_rempty[s])));
    }
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar_acc)));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 16;" ::"r"(smem_u32(&bar_tfree)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = tmem_base_s;
  const uint32_t tiles_u32 = smem_u32(tiles);

  if (RAW && warp == kUmmaThreads / 32 + 1) {
    // ===== loader warp: one lane walks the (sub-segment, chunk) sequence and keeps the raw ring full
    if (lane == 0) {
      uint32_t itl = 0;
      for (int sub = blockIdx.x; sub < lay.nsub; sub += gridDim.x) {
        const int64_t r0 = sub_begin(lay, sub), r1 = sub_begin(lay, sub + 1);
        const int nchunks = (int)((r1 - r0 + kChunkRows - 1) / kChunkRows);
        for (int c = 0; c < nchunks; ++c, ++itl) {
          const uint32_t slot = itl % kRawStages;
          if (itl >= (uint32_t)kRawStages) mbar_wait(smem_u32(&bar_rempty[slot]), ((itl / kRawStages) - 1) & 1);
          const int64_t row = r0 + (int64_t)c * kChunkRows;
          const int nrows = (int)(r1 - row < kChunkRows ? r1 - row : kChunkRows);
          const uint32_t bar = smem_u32(&bar_rfull[slot]);
          const uint32_t dst = smem_u32(raw + slot * RAWBYTES);
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)(nrows * D * 4)) : "memory");
          if (ld == D) {
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst), "l"(X + row * ld), "r"((uint32_t)(nrows * D * 4)), "r"(bar) : "memory");
          } else {
            for (int k = 0; k < nrows; ++k)
              asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                           ::"r"(dst + (uint32_t)(k * D * 4)), "l"(X + (row + k) * ld), "r"((uint32_t)(D * 4)), "r"(bar) : "memory");
          }
        }
      }
    }
  } else if (warp == kUmmaThreads / 32) {
    // ===== MMA warp: one lane issues; walks the same (sub-segment, chunk) sequence as the producers
    if (lane == 0) {
      uint32_t itm = 0, tfree_parity = 0;
      bool first_sub = true;
      for (int sub = blockIdx.x; sub < lay.nsub; sub += gridDim.x) {
        const int64_t r0 = sub_begin(lay, sub), r1 = sub_begin(lay, sub + 1);
        const int nchunks = (int)((r1 - r0 + kChunkRows - 1) / kChunkRows);
        if (nchunks == 0) continue;
        if (!first_sub) {  // the previous accumulator must have been read out
          mbar_wait(smem_u32(&bar_tfree), tfree_parity);
          tfree_parity ^= 1u;
        }
        first_sub = false;
        for (int c = 0; c < nchunks; ++c, ++itm) {
          const uint32_t stage = itm % kUmmaStages;
          mbar_wait(smem_u32(&bar_full[stage]), (itm / kUmmaStages) & 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t hbase = tiles_u32 + stage * NT * TILE, lbase = hbase + TILE;
          constexpr uint64_t kstep = (uint64_t)((2u * SBO) >> 4);  // 8 rows = two 4-row groups, in descriptor units
          if (STACK) {
            constexpr uint32_t idesc = umma_idesc(128, 64);
            uint64_t dh = umma_desc(hbase, LBO, SBO);
#pragma unroll 4
            for (int ks = 0; ks < kChunkRows / 8; ++ks, dh += kstep) umma_tf32(tmem, dh, dh, idesc, (c > 0 || ks > 0) ? 1u : 0u);
          } else if (D <= 128) {
            constexpr uint32_t idesc = umma_idesc(D, D);
            uint64_t dh = umma_desc(hbase, LBO, SBO), dl = umma_desc(lbase, LBO, SBO);
#pragma unroll 4
            for (int ks = 0; ks < kChunkRows / 8; ++ks, dh += kstep, dl += kstep) {
              umma_tf32(tmem, dh, dh, idesc, (c > 0 || ks > 0) ? 1u : 0u);
              if (TERMS == 3) {
                umma_tf32(tmem, dh, dl, idesc, 1u);
                umma_tf32(tmem, dl, dh, idesc, 1u);
              }
            }
          } else {
            // rows 0-127 against all 256 columns -> TMEM columns [0,256); rows 128-255 against columns 128-255 -> [256,384)
            constexpr uint32_t id0 = umma_idesc(128, 256), id1 = umma_idesc(128, 128);
            uint64_t dh0 = umma_desc(hbase, LBO, SBO), dl0 = umma_desc(lbase, LBO, SBO);
            uint64_t dh1 = umma_desc(hbase + 4 * LBO, LBO, SBO), dl1 = umma_desc(lbase + 4 * LBO, LBO, SBO);
#pragma unroll
            for (int ks = 0; ks < kChunkRows / 8; ++ks, dh0 += kstep, dl0 += kstep, dh1 += kstep, dl1 += kstep) {
              const uint32_t acc = (c > 0 || ks > 0) ? 1u : 0u;
              umma_tf32(tmem, dh0, dh0, id0, acc);
              umma_tf32(tmem + 256, dh1, dh1, id1, acc);
              if (TERMS == 3) {
                umma_tf32(tmem, dh0, dl0, id0, 1u);
                umma_tf32(tmem, dl0, dh0, id0, 1u);
                umma_tf32(tmem + 256, dh1, dl1, id1, 1u);
                umma_tf32(tmem + 256, dl1, dh1, id1, 1u);
              }
            }
          }
          // the ring slot is free again once these MMAs have read it
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar_empty[stage])) : "memory");
          if (c == nchunks - 1)
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar_acc)) : "memory");
        }
      }
    }
  } else {
  // ===== producer warps (16): load, clip, shift, split, store tiles; epilogue at the end of each sub-segment
  uint32_t it = 0;        // chunks issued so far by this CTA (ring position)
  // byte offset of (row krow, piece piece0) in a tile: slab = piece >> 3, 16-byte column (piece & 7) of the 128-byte row,
  // 32-byte chunks XOR-ed with row & 3
  const uint32_t b0 = 16u * (piece0 & 7);
  const uint32_t off0 = (uint32_t)(krow >> 2) * SBO + (uint32_t)(piece0 >> 3) * LBO + (uint32_t)(krow & 3) * 128u +
                        ((((b0 >> 5) ^ (uint32_t)(krow & 3)) << 5) | (b0 & 31u));
  float4 m_r[HOIST ? NPIECE : 1], l_r[HOIST ? NPIECE : 1], h_r[HOIST ? NPIECE : 1];
  float4 su_r[HOIST && STD ? NPIECE : 1], si_r[HOIST && STD ? NPIECE : 1];
  if (HOIST) {
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (STD) {
#pragma unroll
      for (int i = 0; i < NPIECE; ++i) {
        su_r[i] = *reinterpret_cast<const float4*>(in_mu + 4 * (piece0 + 16 * i));
        si_r[i] = *reinterpret_cast<const float4*>(in_is + 4 * (piece0 + 16 * i));
      }
    }
#pragma unroll
    for (int i = 0; i < NPIECE; ++i) {
      m_r[i] = *reinterpret_cast<const float4*>(mean_f + 4 * (piece0 + 16 * i));
      l_r[i] = CLIP ? *reinterpret_cast<const float4*>(lo_f + 4 * (piece0 + 16 * i)) : zero4;
      h_r[i] = CLIP ? *reinterpret_cast<const float4*>(hi_f + 4 * (piece0 + 16 * i)) : zero4;
    }
  }
  uint32_t acc_parity = 0;
  for (int sub = blockIdx.x; sub < lay.nsub; sub += gridDim.x) {
    const int64_t r0 = sub_begin(lay, sub), r1 = sub_begin(lay, sub + 1);
    const int nchunks = (int)((r1 - r0 + kChunkRows - 1) / kChunkRows);
    float4 sums[NPIECE];
#pragma unroll
    for (int i = 0; i < NPIECE; ++i) sums[i] = make_float4(0.f, 0.f, 0.f, 0.f);

    // this thread's rows of a chunk: krow + 32*j; its float4 pieces: piece0 + 16*i
    auto load_chunk = [&](int c, float4 (&v)[NROW][NPIECE]) {
#pragma unroll
      for (int j = 0; j < NROW; ++j) {
        const int64_t row = r0 + (int64_t)c * kChunkRows + krow + 32 * j;
        const bool valid = row < r1;
        const float4* p = reinterpret_cast<const float4*>(X + row * ld);
#pragma unroll
        for (int i = 0; i < NPIECE; ++i) v[j][i] = valid ? ldg_stream_f4(p + piece0 + 16 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    };
    // clip, shift, split into hi/lo and store chunk c (held in v4) into its ring slot; then signal the MMA warp
    auto convert_chunk = [&](int c, const float4 (&v4)[NROW][NPIECE]) {
      const uint32_t stage = it % kUmmaStages;
      const uint32_t th = tiles_u32 + stage * NT * TILE + off0;  // hi tile (+ this thread's offset); lo at + LO_OFF
      if (it >= (uint32_t)kUmmaStages) mbar_wait(smem_u32(&bar_empty[stage]), ((it / kUmmaStages) - 1) & 1);
#pragma unroll
      for (int j = 0; j < NROW; ++j) {
        const bool valid = r0 + (int64_t)c * kChunkRows + krow + 32 * j < r1;
#pragma unroll
        for (int i = 0; i < NPIECE; ++i) {
          float v[4] = {v4[j][i].x, v4[j][i].y, v4[j][i].z, v4[j][i].w};
          float mm[4], ll[4], hh[4];
          if (HOIST) {
            mm[0] = m_r[i].x; mm[1] = m_r[i].y; mm[2] = m_r[i].z; mm[3] = m_r[i].w;
            if (CLIP) {
              ll[0] = l_r[i].x; ll[1] = l_r[i].y; ll[2] = l_r[i].z; ll[3] = l_r[i].w;
              hh[0] = h_r[i].x; hh[1] = h_r[i].y; hh[2] = h_r[i].z; hh[3] = h_r[i].w;
            }
          } else {
            const int piece = piece0 + 16 * i;
            const float4 m4 = *reinterpret_cast<const float4*>(ms + 4 * piece);
            mm[0] = m4.x; mm[1] = m4.y; mm[2] = m4.z; mm[3] = m4.w;
            if (CLIP) {
              const float4 l4 = *reinterpret_cast<const float4*>(los + 4 * piece);
              const float4 h4 = *reinterpret_cast<const float4*>(his + 4 * piece);
              ll[0] = l4.x; ll[1] = l4.y; ll[2] = l4.z; ll[3] = l4.w;
              hh[0] = h4.x; hh[1] = h4.y; hh[2] = h4.z; hh[3] = h4.w;
            }
          }
          if (STD) {
            float4 u4, s4;
            if (HOIST) { u4 = su_r[i]; s4 = si_r[i]; }
            else {
              u4 = *reinterpret_cast<const float4*>(his + D + 4 * (piece0 + 16 * i));
              s4 = *reinterpret_cast<const float4*>(his + 2 * D + 4 * (piece0 + 16 * i));
            }
            v[0] = __fmul_rn(__fsub_rn(v[0], u4.x), s4.x);
            v[1] = __fmul_rn(__fsub_rn(v[1], u4.y), s4.y);
            v[2] = __fmul_rn(__fsub_rn(v[2], u4.z), s4.z);
            v[3] = __fmul_rn(__fsub_rn(v[3], u4.w), s4.w);
          }
          uint32_t hi4[4], lo4[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            if (CLIP) v[e] = fminf(fmaxf(v[e], ll[e]), hh[e]);
            const float dv = valid ? v[e] - mm[e] : 0.f;
            if (TERMS == 3) {
              // split by truncation: the tensor core reads the top 19 bits of a tf32 operand and ignores the rest, so
              // "hi" is the value as it is and "lo" what the truncation drops (exact; itself truncated by the hardware)
              hi4[e] = __float_as_uint(dv);
              lo4[e] = __float_as_uint(dv - __uint_as_float(hi4[e] & 0xffffe000u));
            } else {
              hi4[e] = round_tf32(dv);  // a single term keeps round-to-nearest: its error is the whole error
            }
            v[e] = dv;
          }
          sums[i].x += v[0]; sums[i].y += v[1]; sums[i].z += v[2]; sums[i].w += v[3];
          // row k = krow + 32 j sits 8 j four-row groups after row krow; piece + 16 i sits 2 i slabs after piece
          const uint32_t a = th + (uint32_t)(8 * j) * SBO + (uint32_t)(2 * i) * LBO;
          sts128(a, hi4[0], hi4[1], hi4[2], hi4[3]);
          if (TERMS == 3) sts128(a + LO_OFF, lo4[0], lo4[1], lo4[2], lo4[3]);
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy stores -> visible to the tensor core
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bar_full[stage])) : "memory");
      ++it;
    };

    if (RAW) {
      // rows arrive in shared memory by bulk copy; this thread's float4 of row k sits at raw[k][4 piece0 ..]: a warp
      // reads 2 x 256 contiguous bytes, conflict free
      for (int c = 0; c < nchunks; ++c) {
        const uint32_t slot = it % kRawStages;
        mbar_wait(smem_u32(&bar_rfull[slot]), (it / kRawStages) & 1);
        float4 buf[NROW][NPIECE];
        const float4* rp = reinterpret_cast<const float4*>(raw + slot * RAWBYTES);
#pragma unroll
        for (int j = 0; j < NROW; ++j)
#pragma unroll
          for (int i = 0; i < NPIECE; ++i) buf[j][i] = rp[(krow + 32 * j) * (D / 4) + piece0 + 16 * i];
        convert_chunk(c, buf);
        __syncwarp();
        if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bar_rempty[slot])) : "memory");
      }
    } else {
      float4 bufA[NROW][NPIECE], bufB[NROW][NPIECE];
      if (nchunks > 0) load_chunk(0, bufA);
      for (int c = 0; c < nchunks; c += 2) {
        if (c + 1 < nchunks) load_chunk(c + 1, bufB);
        convert_chunk(c, bufA);
        if (c + 1 < nchunks) {
          if (c + 2 < nchunks) load_chunk(c + 2, bufA);
          convert_chunk(c + 1, bufB);
        }
      }
    }

    // ---- end of the sub-segment: row sums, then the accumulator
    float* part = partials + (size_t)sub * PS;
    for (int f = tid; f < D; f += kUmmaThreads) ssum_s[f] = 0.f;
    asm volatile("bar.sync 1, %0;" ::"n"(kUmmaThreads) : "memory");
#pragma unroll
    for (int i = 0; i < NPIECE; ++i) {
      float4 a = sums[i];
      // the two half-warps hold two different rows of the same pieces
      a.x += __shfl_xor_sync(0xffffffffu, a.x, 16);
      a.y += __shfl_xor_sync(0xffffffffu, a.y, 16);
      a.z += __shfl_xor_sync(0xffffffffu, a.z, 16);
      a.w += __shfl_xor_sync(0xffffffffu, a.w, 16);
      if (lane < 16) {
        const int f = 4 * (piece0 + 16 * i);
        atomicAdd(&ssum_s[f], a.x); atomicAdd(&ssum_s[f + 1], a.y); atomicAdd(&ssum_s[f + 2], a.z); atomicAdd(&ssum_s[f + 3], a.w);
      }
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kUmmaThreads) : "memory");
    for (int f = tid; f < D; f += kUmmaThreads) part[f] = ssum_s[f];

    float* G = part + D;
    if (nchunks > 0) {
      mbar_wait(smem_u32(&bar_acc), acc_parity);
      acc_parity ^= 1u;
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      uint32_t r[32];
      const int quarter = warp & 3, cblk = warp >> 2;  // TMEM lanes 32*quarter.., and which columns this warp drains
      if (STACK) {
        // G = HH + LH + LH^T: LH goes through shared memory (the ring is idle: every MMA of the sub-segment has completed
        // and no producer passes the closing barrier before all have finished reading)
        constexpr int LDS_ = D + 1;
        float* lh = reinterpret_cast<float*>(tiles);
        const int c0 = 32 * cblk;
        {
          // lanes 0-63: HH rows, lanes 64-127: LH rows; two column blocks of 32
          if (cblk < 2 && quarter >= 2) {
            tmem_ld32(tmem + ((uint32_t)(32 * quarter) << 16) + c0, r);
            const int m = 32 * (quarter - 2) + lane;
#pragma unroll
            for (int e = 0; e < 32; ++e) lh[m * LDS_ + c0 + e] = __uint_as_float(r[e]);
          }
          asm volatile("bar.sync 1, %0;" ::"n"(kUmmaThreads) : "memory");
          if (cblk < 2 && quarter < 2) {
            tmem_ld32(tmem + ((uint32_t)(32 * quarter) << 16) + c0, r);
            const int m = 32 * quarter + lane;
            float4* dst = reinterpret_cast<float4*>(G + (size_t)m * D + c0);
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              float o[4];
#pragma unroll
              for (int e = 0; e < 4; ++e)
                o[e] = __uint_as_float(r[4 * q + e]) + (lh[m * LDS_ + c0 + 4 * q + e] + lh[(c0 + 4 * q + e) * LDS_ + m]);
              dst[q] = make_float4(o[0], o[1], o[2], o[3]);
            }
          }
        }
      } else if (D == 64) {
        // M = 64: accumulator row m sits in TMEM lane (m & 15) + 32 * (m >> 4): lanes 0-15 of every quarter
        if (cblk < 2) {
          const int c0 = 32 * cblk;
          tmem_ld32(tmem + ((uint32_t)(32 * quarter) << 16) + c0, r);
          if (lane < 16) {
            float4* dst = reinterpret_cast<float4*>(G + (size_t)(16 * quarter + lane) * D + c0);
#pragma unroll
            for (int q = 0; q < 8; ++q)
              dst[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]), __uint_as_float(r[4 * q + 2]),
                                   __uint_as_float(r[4 * q + 3]));
          }
        }
      } else {
        const int m = 32 * quarter + lane;  // M = 128: row m in lane m
        constexpr int NC0 = D == 256 ? 256 : D;
#pragma unroll 1
        for (int c0 = 32 * cblk; c0 < NC0; c0 += 128) {
          tmem_ld32(tmem + ((uint32_t)(32 * quarter) << 16) + c0, r);
          float4* dst = reinterpret_cast<float4*>(G + (size_t)m * D + c0);
#pragma unroll
          for (int q = 0; q < 8; ++q)
            dst[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]), __uint_as_float(r[4 * q + 2]),
                                 __uint_as_float(r[4 * q + 3]));
          if (D == 256 && c0 >= 128) {  // mirror the block above the diagonal into rows 128-255, columns 0-127
#pragma unroll
            for (int e = 0; e < 32; ++e) G[(size_t)(c0 + e) * D + m] = __uint_as_float(r[e]);
          }
        }
        if (D == 256) {
          const int c0 = 32 * cblk;  // 128 columns of the second accumulator: one block per warp
          tmem_ld32(tmem + ((uint32_t)(32 * quarter) << 16) + 256 + c0, r);
          float4* dst = reinterpret_cast<float4*>(G + (size_t)(128 + m) * D + 128 + c0);
#pragma unroll
          for (int q = 0; q < 8; ++q)
            dst[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]), __uint_as_float(r[4 * q + 2]),
                                 __uint_as_float(r[4 * q + 3]));
        }
      }
    } else {
      for (int e = tid; e < D * D; e += kUmmaThreads) G[e] = 0.f;
    }
    // tell the MMA warp that this warp has drained its part of the accumulator; keep the producers together so that
    // ssum_s is not reused early
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncwarp();
    if (nchunks > 0 && lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bar_tfree)) : "memory");
    asm volatile("bar.sync 1, %0;" ::"n"(kUmmaThreads) : "memory");
  }
  }  // producer role
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS));
  }
}

template <int D>
int launch_umma_d(const sb_maha* h, const float* X, int64_t ld, const MahaWs& ws, cudaStream_t st) {
  const SubLayout& L = h->lay;
  int grid = L.nsub < kNumSMs ? L.nsub : kNumSMs;
#define SB_UMMA_LAUNCH2(CLIPV, TERMSV, STDV)                                                                            \
  do {                                                                                                                  \
    const size_t smem = umma_smem_bytes(D, TERMSV);                                                                     \
    SB_CUDA(sb::func_attr_cached(pass1_gram_umma<D, CLIPV, TERMSV, STDV>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                 (int)smem));                                                                           \
    pass1_gram_umma<D, CLIPV, TERMSV, STDV><<<grid, umma_block(D), smem, st>>>(                                         \
        X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, h->d_in_mu, h->d_in_is, ws.partials);                              \
  } while (0)
#define SB_UMMA_LAUNCH(CLIPV, TERMSV)                                                                                   \
  do {                                                                                                                  \
    if (h->d_in_mu) SB_UMMA_LAUNCH2(CLIPV, TERMSV, true); else SB_UMMA_LAUNCH2(CLIPV, TERMSV, false);                   \
  } while (0)
  if (h->use_clip) {
    if (h->gram_terms == 3) SB_UMMA_LAUNCH(true, 3); else SB_UMMA_LAUNCH(true, 1);
  } else {
    if (h->gram_terms == 3) SB_UMMA_LAUNCH(false, 3); else SB_UMMA_LAUNCH(false, 1);
  }
#undef SB_UMMA_LAUNCH
#undef SB_UMMA_LAUNCH2
  return SB_OK;
}

}  // namespace

// returns SB_OK after launching, or 1 if this shape is not handled by the tcgen05 kernel
int launch_pass1_umma(const sb_maha* h, const float* X, int64_t ld, const MahaWs& ws, cudaStream_t st) {
  if (ld % 4 != 0 || ((uintptr_t)X & 15) != 0) return 1;
  switch (h->d) {
    case 64: return launch_umma_d<64>(h, X, ld, ws, st);
    case 128: return launch_umma_d<128>(h, X, ld, ws, st);
    case 256: return launch_umma_d<256>(h, X, ld, ws, st);
  }
  return 1;
}

}  // namespace sb
