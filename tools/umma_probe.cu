// Probe of the tcgen05 (UMMA) conventions the banded plane-sweep kernel relies on; run on a B200:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/umma_probe tools/umma_probe.cu && /tmp/umma_probe
// Checks D[M x N] (fp32, TMEM) = A[M x K] . B[N x K]^T for bf16 operands staged by hand (st.shared, no TMA) in the
// K-major SWIZZLE_NONE canonical layout ("core matrix" = 8 rows x 16 bytes contiguous; SBO = byte distance between
// 8-row groups, LBO = byte distance between consecutive 16-byte K chunks), for M = 128 and M = 64, and reports the
// accuracy of the 3-term bf16 split (a_hi b_hi + a_lo b_hi + a_hi b_lo) against fp64.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3fff);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version 1 (Blackwell)
  return d;                // base_offset 0, lbo_mode 0, layout_type 0 = SWIZZLE_NONE
}

__device__ __forceinline__ uint32_t make_idesc_bf16(int M, int N) {
  uint32_t d = 0;
  d |= 1u << 4;                 // D format f32
  d |= 1u << 7;                 // A format bf16
  d |= 1u << 10;                // B format bf16
  d |= (uint32_t)(N >> 3) << 17;
  d |= (uint32_t)(M >> 4) << 24;
  return d;                     // a_major = b_major = 0 (K-major), no negate, dense
}

// A: [M][K] bf16 row-major in global, B: [N][K]; D out: [128 lanes][N] raw TMEM dump (lane-major)
__global__ void __launch_bounds__(128) probe_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, int M, int N,
                                                    int K, int pad16) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint64_t s_bar;
  const int t = threadIdx.x, warp = t >> 5;
  const int kchunks = K / 8;
  const uint32_t a_plane = (uint32_t)M * 16 + pad16 * 16, b_plane = (uint32_t)N * 16 + pad16 * 16;  // LBO
  uint8_t* sA = smem;
  uint8_t* sB = smem + kchunks * a_plane;
  // stage: chunk kc of row r at (kc * plane + r * 16)
  for (int i = t; i < M * kchunks; i += 128) {
    const int r = i / kchunks, kc = i % kchunks;
    *reinterpret_cast<uint4*>(sA + kc * a_plane + r * 16) = *reinterpret_cast<const uint4*>(A + (size_t)r * K + kc * 8);
  }
  for (int i = t; i < N * kchunks; i += 128) {
    const int r = i / kchunks, kc = i % kchunks;
    *reinterpret_cast<uint4*>(sB + kc * b_plane + r * 16) = *reinterpret_cast<const uint4*>(B + (size_t)r * K + kc * 8);
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(256));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  asm volatile("fence.proxy.async.shared::cta;");  // generic-proxy smem writes -> visible to the tensor core (async proxy)
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = s_tmem;
  if (t == 0) {
    const uint32_t idesc = make_idesc_bf16(M, N);
    for (int ks = 0; ks < K / 16; ++ks) {
      const uint64_t da = make_desc(smem_u32(sA) + 2 * ks * a_plane, a_plane, 128);
      const uint64_t db = make_desc(smem_u32(sB) + 2 * ks * b_plane, b_plane, 128);
      const uint32_t acc = ks > 0;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc));
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_bar)) : "memory");
  }
  // everyone waits for the MMAs
  {
    uint32_t done = 0;
    while (!done) {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}\n"
                   : "=r"(done) : "r"(smem_u32(&s_bar)), "r"(0));
    }
  }
  asm volatile("tcgen05.fence::after_thread_sync;");
  // dump: warp w owns lanes 32w..32w+31; 32 columns per load
  for (int c0 = 0; c0 < N; c0 += 32) {
    uint32_t v[32];
    const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,"
        "%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(addr));
    asm volatile("tcgen05.wait::ld.sync.aligned;");
    for (int j = 0; j < 32 && c0 + j < N; ++j) D[(size_t)t * N + c0 + j] = __uint_as_float(v[j]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256));
}


// Variant with the A operand in TMEM ("TS" form): thread (lane = row m) stores its row's K elements packed two bf16 per
// 32-bit column, 8 columns per K = 16 step, at columns a_col0 + 8 * ks; D at column 0.
__global__ void __launch_bounds__(128) probe_ts_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, int N, int K) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint64_t s_bar;
  const int t = threadIdx.x, warp = t >> 5;
  const int kchunks = K / 8;
  const uint32_t b_plane = (uint32_t)N * 16 + 16;
  uint8_t* sB = smem;
  for (int i = t; i < N * kchunks; i += 128) {
    const int r = i / kchunks, kc = i % kchunks;
    *reinterpret_cast<uint4*>(sB + kc * b_plane + r * 16) = *reinterpret_cast<const uint4*>(B + (size_t)r * K + kc * 8);
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(256));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  asm volatile("fence.proxy.async.shared::cta;");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = s_tmem;
  const uint32_t a_col0 = 128;
  for (int ks = 0; ks < K / 16; ++ks) {
    uint32_t r[8];
    const uint32_t* src = reinterpret_cast<const uint32_t*>(A + (size_t)t * K + ks * 16);
    for (int c = 0; c < 8; ++c) r[c] = src[c];
    const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16) + a_col0 + ks * 8;
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(addr), "r"(r[0]), "r"(r[1]),
                 "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]));
  }
  asm volatile("tcgen05.wait::st.sync.aligned;");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  if (t == 0) {
    const uint32_t idesc = make_idesc_bf16(128, N);
    for (int ks = 0; ks < K / 16; ++ks) {
      const uint64_t db = make_desc(smem_u32(sB) + 2 * ks * b_plane, b_plane, 128);
      const uint32_t acc = ks > 0;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(tmem), "r"(tmem + a_col0 + ks * 8), "l"(db), "r"(idesc), "r"(acc));
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_bar)) : "memory");
  }
  {
    uint32_t done = 0;
    while (!done) {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}\n"
                   : "=r"(done) : "r"(smem_u32(&s_bar)), "r"(0));
    }
  }
  asm volatile("tcgen05.fence::after_thread_sync;");
  for (int c0 = 0; c0 < N; c0 += 32) {
    uint32_t v[32];
    const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,"
        "%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(addr));
    asm volatile("tcgen05.wait::ld.sync.aligned;");
    for (int j = 0; j < 32 && c0 + j < N; ++j) D[(size_t)t * N + c0 + j] = __uint_as_float(v[j]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256));
}

static float bf16_round(float x) { return __bfloat162float(__float2bfloat16(x)); }

int main() {
  const int K = 64;
  for (int M : {128, 64}) {
    for (int N : {64, 128, 256}) {
      for (int pad16 : {0, 1}) {
        std::vector<__nv_bfloat16> hA((size_t)M * K), hB((size_t)N * K);
        std::vector<float> fA(hA.size()), fB(hB.size());
        srand(7);
        for (size_t i = 0; i < hA.size(); ++i) { fA[i] = bf16_round((rand() % 2001 - 1000) / 500.0f); hA[i] = __float2bfloat16(fA[i]); }
        for (size_t i = 0; i < hB.size(); ++i) { fB[i] = bf16_round((rand() % 2001 - 1000) / 500.0f); hB[i] = __float2bfloat16(fB[i]); }
        __nv_bfloat16 *dA, *dB;
        float* dD;
        CK(cudaMalloc(&dA, hA.size() * 2)); CK(cudaMalloc(&dB, hB.size() * 2)); CK(cudaMalloc(&dD, 128 * N * 4));
        CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemset(dD, 0xff, 128 * N * 4));
        const size_t smem = (size_t)(K / 8) * ((M + N) * 16 + 2 * pad16 * 16) + 1024;
        CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        probe_kernel<<<1, 128, smem>>>(dA, dB, dD, M, N, K, pad16);
        CK(cudaDeviceSynchronize());
        std::vector<float> hD((size_t)128 * N);
        CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
        // reference
        std::vector<double> ref((size_t)M * N);
        for (int m = 0; m < M; ++m)
          for (int n = 0; n < N; ++n) {
            double s = 0;
            for (int k = 0; k < K; ++k) s += (double)fA[(size_t)m * K + k] * fB[(size_t)n * K + k];
            ref[(size_t)m * N + n] = s;
          }
        // find the lane of each row: try identity and the M=64 candidates
        double err_id = 0;
        for (int m = 0; m < M; ++m)
          for (int n = 0; n < N; ++n) err_id = fmax(err_id, fabs(hD[(size_t)m * N + n] - ref[(size_t)m * N + n]));
        printf("M=%d N=%d pad=%d: max|D[lane=row] - ref| = %.3e", M, N, pad16, err_id);
        if (M == 64 && err_id > 1e-2) {
          // search which lane holds row m (column 0..N-1 all match)
          printf("  lane map:");
          for (int m = 0; m < 64; m += 9) {
            int found = -1;
            for (int l = 0; l < 128; ++l) {
              bool ok = true;
              for (int n = 0; n < N && ok; ++n) ok = fabs(hD[(size_t)l * N + n] - ref[(size_t)m * N + n]) < 1e-3;
              if (ok) { found = l; break; }
            }
            printf(" row%d->lane%d", m, found);
          }
        }
        printf("\n");
        cudaFree(dA); cudaFree(dB); cudaFree(dD);
      }
    }
  }
  // A operand from TMEM
  for (int N : {64, 128}) {
    const int M = 128;
    std::vector<__nv_bfloat16> hA((size_t)M * K), hB((size_t)N * K);
    std::vector<float> fA(hA.size()), fB(hB.size());
    srand(9);
    for (size_t i = 0; i < hA.size(); ++i) { fA[i] = bf16_round((rand() % 2001 - 1000) / 500.0f); hA[i] = __float2bfloat16(fA[i]); }
    for (size_t i = 0; i < hB.size(); ++i) { fB[i] = bf16_round((rand() % 2001 - 1000) / 500.0f); hB[i] = __float2bfloat16(fB[i]); }
    __nv_bfloat16 *dA, *dB;
    float* dD;
    CK(cudaMalloc(&dA, hA.size() * 2)); CK(cudaMalloc(&dB, hB.size() * 2)); CK(cudaMalloc(&dD, 128 * N * 4));
    CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
    const size_t smem = (size_t)(K / 8) * (N * 16 + 16) + 1024;
    CK(cudaFuncSetAttribute(probe_ts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    probe_ts_kernel<<<1, 128, smem>>>(dA, dB, dD, N, K);
    CK(cudaDeviceSynchronize());
    std::vector<float> hD((size_t)128 * N);
    CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
    double err = 0;
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < N; ++n) {
        double sref = 0;
        for (int k = 0; k < K; ++k) sref += (double)fA[(size_t)m * K + k] * fB[(size_t)n * K + k];
        err = fmax(err, fabs(hD[(size_t)m * N + n] - sref));
      }
    printf("A in TMEM: M=128 N=%d: max|D - ref| = %.3e\n", N, err);
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
  }
  // accuracy of the 3-term bf16 split for a C = 256 dot product of unit-variance fp32 values
  {
    const int C = 256, T = 20000;
    double worst = 0, worst2 = 0, scale = 0;
    srand(11);
    auto rnd = []() { double u = 0; for (int i = 0; i < 12; ++i) u += rand() / (double)RAND_MAX; return (float)(u - 6.0); };
    for (int it = 0; it < T; ++it) {
      double exact = 0, split3 = 0, split1 = 0;
      for (int c = 0; c < C; ++c) {
        const float a = rnd(), b = rnd();
        const float ah = bf16_round(a), al = bf16_round(a - ah), bh = bf16_round(b), bl = bf16_round(b - bh);
        exact += (double)a * b;
        split3 += (double)ah * bh + (double)al * bh + (double)ah * bl;
        split1 += (double)ah * bh;
      }
      worst = fmax(worst, fabs(split3 - exact));
      worst2 = fmax(worst2, fabs(split1 - exact));
      scale = fmax(scale, fabs(exact));
    }
    printf("bf16 split over C=256: max|3-term - exact| = %.3e, max|1-term - exact| = %.3e, max|exact| = %.3e -> rel %.2e / %.2e\n",
           worst, worst2, scale, worst / scale, worst2 / scale);
  }
  return 0;
}
