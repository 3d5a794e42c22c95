// Development probe for the int8 tcgen05 path (not part of the library): one CTA pair (cta_group::2) or one CTA
// (cta_group::1) multiplies MN-major int8 planes loaded by TMA tensor copies with the 128-byte swizzle and checks the
// int32 accumulators read back from tensor memory against the host.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/umma_probe tools/umma_probe.cu && /tmp/umma_probe [rows] [pair]
// Every mbarrier wait is bounded, so a wrong descriptor ends in an error code, not a hang.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

constexpr int KC = 64;            // rows of a stage
constexpr int COLS = 256;         // plane width in bytes (= columns of X)
constexpr int HALF = 128;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_wait_bounded(uint64_t* bar, unsigned parity) {
  const uint32_t addr = smem_u32(bar);
  for (int it = 0; it < 2000000; ++it) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    if (ok) return true;
  }
  return false;
}
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// MN-major, 128-byte swizzle: atoms of 8 rows x 128 bytes; K groups of 8 rows are SBO apart, 128-byte column groups LBO apart
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
  d |= (uint64_t)1 << 46;          // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;          // SWIZZLE_128B
  return d;
}
__host__ __device__ constexpr uint32_t idesc_i8(int M, int N) {
  return (2u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

template <int PAIR>
__global__ void __launch_bounds__(192) k_probe(const __grid_constant__ CUtensorMap tmap, int rows, int32_t* out, int* err) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* tiles = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);     // [2 planes][PAIR ? 1 : 2 halves][KC][128]
  __shared__ __align__(8) uint64_t full_bar, done_bar;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0;
  constexpr int NHALF = PAIR ? 1 : 2;                 // column halves this CTA stages per plane
  constexpr uint32_t TILE = KC * HALF;                // 8 KB
  if (threadIdx.x == 0) {
    mbar_init(&full_bar, PAIR ? 2 : 1);
    mbar_init(&done_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    if (PAIR) asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 256;" :: "r"(smem_u32(&tmem_base_s)));
    else asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" :: "r"(smem_u32(&tmem_base_s)));
    if (PAIR) asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    else asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if (PAIR) cluster_sync(); else __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_base_s;
  const int nchunks = (rows + KC - 1) / KC;
  bool ok = true;

  for (int ch = 0; ch < nchunks && ok; ++ch) {
    const unsigned parity = ch & 1;
    if (warp == 0 && lane == 0) {
      // producer: plane 0 = A (row offset 0), plane 1 = B (row offset `rows` in the same tensor)
      const uint32_t bytes = 2 * NHALF * TILE;
      if (PAIR) {
        // both CTAs signal the leader's barrier
        const uint32_t bar = smem_u32(&full_bar) & 0xFEFFFFFFu;
        if (rank == 0) mbar_expect_tx(&full_bar, 2 * bytes);
        else {
          // arrive (without tx) on the leader's barrier
          asm volatile("{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, 0;\n\tmbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}"
                       :: "r"(smem_u32(&full_bar)) : "memory");
        }
        for (int pl = 0; pl < 2; ++pl) {
          const uint32_t dst = smem_u32(tiles + pl * TILE);
          const int c0 = (int)rank * HALF, c1 = pl * rows + ch * KC;
          asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                       :: "r"(dst), "l"(&tmap), "r"(bar), "r"(c0), "r"(c1) : "memory");
        }
      } else {
        mbar_expect_tx(&full_bar, bytes);
        for (int pl = 0; pl < 2; ++pl)
          for (int h = 0; h < 2; ++h) {
            const uint32_t dst = smem_u32(tiles + (pl * 2 + h) * TILE);
            const int c0 = h * HALF, c1 = pl * rows + ch * KC;
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                         :: "r"(dst), "l"(&tmap), "r"(smem_u32(&full_bar)), "r"(c0), "r"(c1) : "memory");
          }
      }
    }
    if (warp == 1 && rank == 0) {
      // MMA issuer (leader CTA)
      ok = mbar_wait_bounded(&full_bar, parity);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (ok && lane == 0) {
        const uint32_t a_base = smem_u32(tiles), b_base = smem_u32(tiles + NHALF * TILE);
        for (int ks = 0; ks < KC / 32; ++ks) {
          const uint64_t ad = umma_desc(a_base + ks * 32 * HALF, TILE, 8 * HALF);
          const uint64_t bd = umma_desc(b_base + ks * 32 * HALF, TILE, 8 * HALF);
          const uint32_t acc = (ch > 0 || ks > 0) ? 1u : 0u;
          if (PAIR) {
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n\t}"
                         :: "r"(tmem_base), "l"(ad), "l"(bd), "r"(idesc_i8(256, 256)), "r"(acc), "r"(0u) : "memory");
          } else {
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                         :: "r"(tmem_base), "l"(ad), "l"(bd), "r"(idesc_i8(128, 256)), "r"(acc), "r"(0u) : "memory");
          }
        }
        // completion of everything issued so far -> done_bar (in both CTAs of a pair)
        if (PAIR) asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                               :: "r"(smem_u32(&done_bar)), "h"((uint16_t)3) : "memory");
        else asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&done_bar)) : "memory");
      }
      __syncwarp();
    }
    // everyone waits for the MMAs of this chunk before the stage is overwritten (single stage: probe only)
    ok = ok && mbar_wait_bounded(&done_bar, parity);
    ok = __syncthreads_and(ok);
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (!ok) { if (threadIdx.x == 0) atomicExch(err, 1 + (int)rank); }
  if (ok && warp >= 2) {
    // epilogue: warps 2..5 -> lane quarters (warp % 4)
    const int q = warp & 3;
    for (int c0 = 0; c0 < 256; c0 += 32) {
      uint32_t v[32];
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                   "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                   "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                   : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                     "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                     "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                     "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                   : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      const int row = (int)rank * 128 + q * 32 + lane;
      for (int j = 0; j < 32; ++j) out[(size_t)row * 256 + c0 + j] = (int32_t)v[j];
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if (PAIR) cluster_sync(); else __syncthreads();
  if (warp == 1) {
    if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 256;" :: "r"(tmem_base));
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" :: "r"(tmem_base));
  }
}

// Issue rate of tcgen05.mma kind::i8 (M = 256 pair, N = 256, K = 32) from resident shared-memory tiles, for MN-major and
// K-major operand descriptors (the data is whatever the tiles hold: timing only).
__global__ void __launch_bounds__(64) k_rate(int iters, int kmajor, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* tiles = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t done_bar;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  for (int i = threadIdx.x; i < 8 * 8192 / 4; i += 64) ((uint32_t*)tiles)[i] = 0x01010101u * (uint32_t)(i & 3);
  if (threadIdx.x == 0) { mbar_init(&done_bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base_s)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  cluster_sync();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_base_s;
  long long t0 = 0, t1 = 0;
  if (warp == 0 && rank == 0 && lane == 0) {
    const uint32_t idesc = kmajor ? ((2u << 4) | (1u << 7) | (1u << 10) | ((256u >> 3) << 17) | ((256u >> 4) << 24)) : idesc_i8(256, 256);
    t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint32_t a_base = smem_u32(tiles) + (uint32_t)(it & 3) * 8192u, b_base = smem_u32(tiles) + (uint32_t)(4 + (it & 3)) * 8192u;
      for (int ks = 0; ks < 2; ++ks) {
        // K-major SW128: rows of 128 bytes (128 k), 8-row groups 1024 bytes apart; K advances 32 bytes inside the row
        const uint64_t ad = kmajor ? umma_desc(a_base + ks * 32, 16, 1024) : umma_desc(a_base + ks * 32 * HALF, 8192, 8 * HALF);
        const uint64_t bd = kmajor ? umma_desc(b_base + ks * 32, 16, 1024) : umma_desc(b_base + ks * 32 * HALF, 8192, 8 * HALF);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n\t}"
                     :: "r"(tmem_base + (uint32_t)(it & 1) * 256u), "l"(ad), "l"(bd), "r"(idesc), "r"(1u), "r"(0u) : "memory");
      }
    }
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 :: "r"(smem_u32(&done_bar)), "h"((uint16_t)3) : "memory");
  }
  mbar_wait_bounded(&done_bar, 0);
  if (warp == 0 && rank == 0 && lane == 0) { t1 = clock64(); cycles[0] = t1 - t0; }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  cluster_sync();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" :: "r"(tmem_base));
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int rate_test() {
  long long* dc; CK(cudaMalloc(&dc, 8));
  const size_t smem = 8 * 8192 + 1024;
  CK(cudaFuncSetAttribute(k_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  for (int kmajor = 0; kmajor < 2; ++kmajor)
    for (int pairs = 1; pairs <= 72; pairs *= 72) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(2 * pairs); cfg.blockDim = dim3(64); cfg.dynamicSmemBytes = smem;
      cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      const int iters = 20000;
      CK(cudaLaunchKernelEx(&cfg, k_rate, iters, kmajor, dc));
      CK(cudaDeviceSynchronize());
      long long c = 0; CK(cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost));
      printf("rate %s pairs=%d: %.1f clk per MMA (M256 N256 K32)\n", kmajor ? "K-major" : "MN-major", pairs, (double)c / (2.0 * iters));
    }
  return 0;
}

int main(int argc, char** argv) {
  if (argc > 1 && atoi(argv[1]) == 0) return rate_test();
  const int rows = argc > 1 ? atoi(argv[1]) : 256;
  const int pair = argc > 2 ? atoi(argv[2]) : 0;
  std::vector<int8_t> h((size_t)2 * rows * COLS);
  uint32_t s = 12345;
  for (auto& v : h) { s = s * 1664525u + 1013904223u; v = (int8_t)(s >> 24); }
  int8_t* d; int32_t* dout; int* derr;
  CK(cudaMalloc(&d, h.size() + 64 * COLS));
  CK(cudaMemset(d, 0, h.size() + 64 * COLS));
  CK(cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice));
  CK(cudaMalloc(&dout, 256 * 256 * 4)); CK(cudaMemset(dout, 0xff, 256 * 256 * 4));
  CK(cudaMalloc(&derr, 4)); CK(cudaMemset(derr, 0, 4));
  void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
  if (!fn) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
  CUtensorMap tm;
  // NOTE: plane B starts at row `rows`; rows of the last chunk of plane A beyond `rows` would read plane B's rows, so the
  // probe needs rows % KC == 0 unless the planes are padded -- they are separated by using the tensor extent 2*rows and
  // masking is NOT done here: keep rows a multiple of KC.
  cuuint64_t dims[2] = {(cuuint64_t)COLS, (cuuint64_t)(2 * rows)};
  cuuint64_t strides[1] = {(cuuint64_t)COLS};
  cuuint32_t box[2] = {HALF, KC};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = ((EncodeFn)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
  const size_t smem = 4 * KC * HALF + 1024;
  if (pair) {
    CK(cudaFuncSetAttribute(k_probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2); cfg.blockDim = dim3(192); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    CK(cudaLaunchKernelEx(&cfg, k_probe<1>, tm, rows, dout, derr));
  } else {
    CK(cudaFuncSetAttribute(k_probe<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_probe<0><<<1, 192, smem>>>(tm, rows, dout, derr);
  }
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  int herr = 0; CK(cudaMemcpy(&herr, derr, 4, cudaMemcpyDeviceToHost));
  std::vector<int32_t> got(256 * 256);
  CK(cudaMemcpy(got.data(), dout, got.size() * 4, cudaMemcpyDeviceToHost));
  const int M = pair ? 256 : 128;
  long bad = 0;
  for (int i = 0; i < M; ++i)
    for (int j = 0; j < 256; ++j) {
      long acc = 0;
      for (int k = 0; k < rows; ++k) acc += (long)h[(size_t)k * COLS + i] * (long)h[(size_t)(rows + k) * COLS + j];
      if ((int32_t)acc != got[(size_t)i * 256 + j]) { if (bad < 5) printf("mismatch (%d,%d): want %ld got %d\n", i, j, acc, got[(size_t)i * 256 + j]); ++bad; }
    }
  printf("probe rows=%d pair=%d err=%d mismatches=%ld of %d\n", rows, pair, herr, bad, M * 256);
  return (herr || bad) ? 2 : 0;
}
