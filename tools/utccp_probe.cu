// Probe of tcgen05.cp (shared memory -> TMEM, SASS UTCCP) for the key tile of psweep_ws.cu; run on a B200:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/utccp_probe tools/utccp_probe.cu && /tmp/utccp_probe
// Source: 128 rows x 32 bytes in UMMA's K-major SWIZZLE_NONE canonical layout (core matrix = 8 rows x 16 bytes contiguous;
// LBO = bytes between the two 16-byte K chunks, SBO = bytes between 8-row groups).  Expected: TMEM lane r, columns c..c+7 =
// the eight 32-bit words of row r (chunk 0 then chunk 1), i.e. what tcgen05.st.32x32b.x8 of the same words produces.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((addr >> 4) & 0x3fff) | ((uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32) |
         ((uint64_t)1 << 46);
}

// variant 0: groups 256 B apart (each group = [chunk 2][row 8][16 B]), chunks 128 B apart
// variant 1: all chunk-0 core matrices first (groups 128 B apart), chunk 1 at +2048
__global__ void __launch_bounds__(128) probe(uint32_t* out, int variant, int col0) {
  __shared__ __align__(1024) uint8_t sm[4096];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint64_t s_bar;
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const uint32_t lbo = variant == 0 ? 128 : 2048, sbo = variant == 0 ? 256 : 128;
  for (int i = t; i < 128 * 8; i += 128) {
    const int r = i >> 3, wd = i & 7;               // row, 32-bit word of its 32 bytes
    const int g = r >> 3, rr = r & 7, ch = wd >> 2, w4 = wd & 3;
    *reinterpret_cast<uint32_t*>(sm + g * sbo + ch * lbo + rr * 16 + w4 * 4) = ((uint32_t)r << 8) | (uint32_t)wd;
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(64));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic st.shared -> async-proxy read by the copy
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s_tmem;
  if (t == 0) {
    const uint64_t d = make_desc(smem_u32(sm), lbo, sbo);
    asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(tmem + col0), "l"(d) : "memory");
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_bar)) : "memory");
  }
  {
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}\n" : "=r"(done) : "r"(smem_u32(&s_bar)), "r"(0) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(tmem + col0 + ((uint32_t)(warp * 32) << 16)));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int i = 0; i < 8; ++i) out[(warp * 32 + lane) * 8 + i] = r[i];
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64));
}

int main() {
  uint32_t* d;
  CK(cudaMalloc(&d, 128 * 8 * 4));
  uint32_t h[128 * 8];
  for (int variant = 0; variant < 2; ++variant) {
    for (int col0 = 0; col0 <= 8; col0 += 8) {
      CK(cudaMemset(d, 0xff, sizeof(h)));
      probe<<<1, 128>>>(d, variant, col0);
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost));
      int ok = 0;
      for (int r = 0; r < 128; ++r)
        for (int w = 0; w < 8; ++w) ok += h[r * 8 + w] == (((uint32_t)r << 8) | (uint32_t)w);
      printf("variant %d col0 %d: %d / 1024 words where expected\n", variant, col0, ok);
      if (ok != 1024) {
        for (int r = 0; r < 128; r += 9) {
          printf("  lane %3d:", r);
          for (int w = 0; w < 8; ++w) printf(" (%u,%u)", h[r * 8 + w] >> 8, h[r * 8 + w] & 255);
          printf("\n");
        }
      }
    }
  }
  return 0;
}
