// Development probe (not part of the library): how fast can k SMs write to a peer GPU's memory over NVLink,
//   mode 0: 16-byte st.global from 1024 threads per CTA (loads from local memory with ld.global.cg),
//   mode 1: cp.async.bulk local global -> shared, then cp.async.bulk shared -> peer global, `chunk` bytes per copy, 16 warps x 2 buffers.
// Built and driven by tools/peer_rate_probe.py.
#include <cuda_runtime.h>
#include <stdint.h>

namespace {
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(1024, 1) st_kernel(float4* dst, const float4* src, size_t n16) {
  for (size_t i = (size_t)blockIdx.x * 1024 + threadIdx.x; i < n16; i += (size_t)gridDim.x * 1024) dst[i] = __ldcg(src + i);
}

constexpr int kWarps = 16, kBufs = 2, kMaxChunk = 4096;

__global__ void __launch_bounds__(kWarps * 32, 1) bulk_kernel(char* dst, const char* src, size_t bytes, int chunk) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ unsigned long long bars[kWarps * kBufs];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane != 0) return;
  unsigned char* mybuf = smem + (size_t)warp * kBufs * kMaxChunk;
  for (int b = 0; b < kBufs; ++b) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(smem_u32(&bars[warp * kBufs + b])));
  asm volatile("fence.mbarrier_init.release.cluster;\n");
  const size_t nchunks = bytes / chunk;
  const size_t stride = (size_t)gridDim.x * kWarps;
  size_t c = (size_t)blockIdx.x * kWarps + warp;
  unsigned it = 0;
  // prologue: load for buffer 0
  if (c < nchunks) {
    const unsigned bar = smem_u32(&bars[warp * kBufs]);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(chunk) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(smem_u32(mybuf)),
                 "l"(src + c * chunk), "r"(chunk), "r"(bar) : "memory");
  }
  for (; c < nchunks; c += stride, ++it) {
    const unsigned b = it % kBufs, nb = (it + 1) % kBufs;
    const size_t cn = c + stride;
    if (cn < nchunks) {
      // buffer nb was last read by the store issued one iteration ago: wait until at most 0 older stores still read smem
      asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");
      const unsigned bar = smem_u32(&bars[warp * kBufs + nb]);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(chunk) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(smem_u32(mybuf + nb * kMaxChunk)),
                   "l"(src + cn * chunk), "r"(chunk), "r"(bar) : "memory");
    }
    const unsigned bar = smem_u32(&bars[warp * kBufs + b]), parity = (it / kBufs) & 1;
    unsigned done = 0;
    while (!done)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(dst + c * chunk), "r"(smem_u32(mybuf + b * kMaxChunk)), "r"(chunk) : "memory");
    asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
  }
  asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory");
}
}  // namespace

extern "C" int probe_launch(void* dst, const void* src, size_t bytes, int ctas, int mode, int chunk, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 0) {
    st_kernel<<<ctas, 1024, 0, st>>>((float4*)dst, (const float4*)src, bytes / 16);
  } else {
    if (chunk < 16 || chunk > kMaxChunk || chunk % 16) return -1;
    const int smem = kWarps * kBufs * kMaxChunk;
    static bool set = false;
    if (!set) { cudaFuncSetAttribute(bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); set = true; }
    bulk_kernel<<<ctas, kWarps * 32, smem, st>>>((char*)dst, (const char*)src, bytes, chunk);
  }
  return (int)cudaGetLastError();
}
