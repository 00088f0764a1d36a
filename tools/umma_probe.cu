// Probe for tcgen05.mma kind::tf32 with both operands read from ONE shared-memory tile (Gram: D = T^T T).
// Not part of the library: it pins down, on real hardware, the descriptor fields and the TMEM accumulator layout
// that csrc/maha_gram_umma.cu relies on.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -o umma_probe umma_probe.cu
//
// Tile: X[k][f], k = 0..31 rows, f = 0..D-1 features, stored "MN-major" (feature contiguous) in the canonical
// SWIZZLE_128B layout: 32-feature slabs of 128 B per row, 8-row groups of 1024 B, 16-byte chunks XOR-ed with (k & 7):
//   addr(k, f) = (k >> 3) * SBO + (f >> 5) * LBO + (k & 7) * 128 + ((((f & 31) >> 2) ^ (k & 7)) << 4) + (f & 3) * 4
// Usage: umma_probe D M lbo sbo amajor bmajor [layout_type] [swap_lbo_sbo]
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1)
probe(const float* __restrict__ X, int D, int M, int N, uint32_t lbo, uint32_t sbo, int amajor, int bmajor, int ltype,
      uint32_t desc_lbo, uint32_t desc_sbo, float* __restrict__ out /* [128][512] */, int ncols, int init_pattern, int do_mma) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t mbar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // tile at the 1024-aligned start of dynamic smem
  uint8_t* tile = (uint8_t*)(((uintptr_t)smem + 1023) & ~(uintptr_t)1023);
  const int tile_bytes = 4 * sbo > (uint32_t)(D / 32) * lbo ? 4 * sbo : (D / 32) * lbo;
  for (int i = tid; i < 140 * 256; i += blockDim.x) ((uint32_t*)tile)[i] = 0;
  (void)tile_bytes;
  __syncthreads();
  for (int idx = tid; idx < 32 * D; idx += blockDim.x) {
    const int k = idx / D, f = idx - k * D;
    uint32_t a;
    if (ltype == 1) {  // SWIZZLE_128B_BASE32B (the only MN-major layout for tf32): 4-row atoms, 32-byte chunks ^ (k & 3)
      const uint32_t b = (f & 31) * 4;
      a = (k >> 2) * sbo + (f >> 5) * lbo + (k & 3) * 128 + ((((b >> 5) ^ (k & 3)) << 5) | (b & 31));
    } else {
      a = (k >> 3) * sbo + (f >> 5) * lbo + (k & 7) * 128 + ((((f & 31) >> 2) ^ (k & 7)) << 4) + (f & 3) * 4;
    }
    *(float*)(tile + a) = X[idx];
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(ncols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;");  // generic-proxy smem writes -> visible to the tensor core (async proxy)
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = tmem_base_s;
  if (init_pattern) {
    for (int c0 = 0; c0 < ncols; c0 += 32) {
      uint32_t r[32];
      for (int i = 0; i < 32; ++i) r[i] = __float_as_uint((float)(1000 * (32 * warp + lane) + c0 + i));
      const uint32_t taddr = tmem + ((uint32_t)(32 * warp) << 16) + c0;
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,"
          "%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};\n\ttcgen05.wait::st.sync.aligned;"
          ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
          "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
          "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
          "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
  }
  if (tid == 0 && do_mma) {
    // instruction descriptor: c=F32 (1<<4), a=TF32 (2<<7), b=TF32 (2<<10), majors, N>>3 at 17, M>>4 at 24
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)amajor << 15) | ((uint32_t)bmajor << 16) |
                           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    const uint32_t base = smem_u32(tile);
    for (int kg = 0; kg < 4; ++kg) {
      const uint32_t start = base + kg * (ltype == 1 ? 2 * sbo : sbo);
      uint64_t desc = (uint64_t)((start >> 4) & 0x3fff) | ((uint64_t)((desc_lbo >> 4) & 0x3fff) << 16) |
                      ((uint64_t)((desc_sbo >> 4) & 0x3fff) << 32) | (1ull << 46) | ((uint64_t)ltype << 61);
      const uint32_t accum = (kg > 0) || init_pattern;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(desc), "l"(desc), "r"(idesc),
          "r"(accum));
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar)));
  }
  if (tid == 0 && !do_mma) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&mbar)));
  // wait for the MMAs
  {
    uint32_t done = 0;
    while (!done) {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(done) : "r"(smem_u32(&mbar)), "r"(0));
    }
  }
  asm volatile("tcgen05.fence::after_thread_sync;");
  // dump: warp w reads its 32 lanes, 32 columns at a time
  for (int c0 = 0; c0 < ncols; c0 += 32) {
    uint32_t r[32];
    const uint32_t taddr = tmem + ((uint32_t)(32 * warp) << 16) + c0;
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
    for (int i = 0; i < 32; ++i) out[(32 * warp + lane) * 512 + c0 + i] = __uint_as_float(r[i]);
  }
  if (tid == 0) { out[127 * 512 + 511] = 12345.f; out[127 * 512 + 510] = __uint_as_float(tmem); }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(ncols));
}

int main(int argc, char** argv) {
  const int D = argc > 1 ? atoi(argv[1]) : 64;
  const int M = argc > 2 ? atoi(argv[2]) : D;
  const uint32_t lbo = argc > 3 ? atoi(argv[3]) : 1024;
  const uint32_t sbo = argc > 4 ? atoi(argv[4]) : 1024 * (D / 32);
  const int amajor = argc > 5 ? atoi(argv[5]) : 1, bmajor = argc > 6 ? atoi(argv[6]) : 1;
  const int ltype = argc > 7 ? atoi(argv[7]) : 2;
  const int swap = argc > 8 ? atoi(argv[8]) : 0;
  const int init_pattern = argc > 9 ? atoi(argv[9]) : 0;
  const int do_mma = argc > 10 ? atoi(argv[10]) : 1;
  const int N = D;
  int ncols = 32;
  while (ncols < N) ncols <<= 1;
  std::vector<float> X(32 * D);
  srand(1);
  for (auto& v : X) v = (float)((rand() % 7) - 3);  // small integers: exact in tf32, exact sums
  float *dX, *dOut;
  CK(cudaMalloc(&dX, X.size() * 4));
  CK(cudaMalloc(&dOut, 128 * 512 * 4));
  CK(cudaMemcpy(dX, X.data(), X.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(dOut, 0, 128 * 512 * 4));
  const size_t smem = 150 * 1024;
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  probe<<<1, 128, smem>>>(dX, D, M, N, lbo, sbo, amajor, bmajor, ltype, swap ? sbo : lbo, swap ? lbo : sbo, dOut, ncols, init_pattern, do_mma);
  CK(cudaDeviceSynchronize());
  std::vector<float> out(128 * 512);
  CK(cudaMemcpy(out.data(), dOut, out.size() * 4, cudaMemcpyDeviceToHost));
  // expected Gram
  std::vector<double> G((size_t)D * D, 0.0);
  for (int k = 0; k < 32; ++k)
    for (int a = 0; a < D; ++a)
      for (int b = 0; b < D; ++b) G[(size_t)a * D + b] += (double)X[k * D + a] * X[k * D + b];
  // hypotheses for the lane of accumulator row m
  auto lane128 = [](int m) { return m; };
  auto lane64 = [](int m) { return (m & 15) + 32 * (m >> 4); };
  for (int hyp = 0; hyp < 2; ++hyp) {
    int bad = 0, badT = 0;
    for (int m = 0; m < M && m < D; ++m)
      for (int n = 0; n < N; ++n) {
        const int ln = hyp == 0 ? lane128(m) : lane64(m);
        if (ln >= 128) { ++bad; ++badT; continue; }
        if (out[ln * 512 + n] != (float)G[(size_t)m * D + n]) ++bad;
      }
    printf("D=%d M=%d N=%d lbo=%u sbo=%u amaj=%d bmaj=%d ltype=%d swap=%d | hypothesis %s: %d mismatches of %d\n", D, M, N, lbo,
           sbo, amajor, bmajor, ltype, swap, hyp == 0 ? "lane=m" : "lane=(m&15)+32*(m>>4)", bad, (M < D ? M : D) * N);
  }
  int nz = 0;
  for (int l = 0; l < 128; ++l)
    for (int c = 0; c < ncols; ++c) nz += out[l * 512 + c] != 0.f;
  { uint32_t tb; memcpy(&tb, &out[127 * 512 + 510], 4); printf("  marker %g tmem_base 0x%08x\n", out[127 * 512 + 511], tb); }
  printf("  out[1][0..3] = %g %g %g %g out[33][2] = %g\n", out[512], out[513], out[514], out[515], out[33 * 512 + 2]);
  printf("  nonzero TMEM entries: %d; sample out[0][0..3] = %g %g %g %g expected G[0][0..3] = %g %g %g %g\n", nz, out[0], out[1],
         out[2], out[3], G[0], G[1], G[2], G[3]);
  return 0;
}
