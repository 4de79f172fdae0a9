// Compact transport of the sampling mask for callers whose outputs go back to HOST memory (pipeline.HostPlanesweepPipeline, the
// end-to-end leg of bench.py).  The mask of PlanesweepCorrelation.forward (planesweep_corr.py:99-102, :189-191) is a float32 tensor
// holding only 0.0 and 1.0: half of the bytes the host path sends over PCIe.  Here the device packs it to one byte per element
// (mask_pack_kernel), the byte tensor crosses PCIe, and a host function queued on a CUDA stream (cudaLaunchHostFunc) expands it
// into the caller's float32 buffer with streaming stores on a small worker pool -- the host buffer ends up bit-identical to a
// float32 device-to-host copy of the mask.  No counterpart in the reference (it keeps the tensors on one device).
#include <cuda_runtime.h>
#include <emmintrin.h>
#include <unistd.h>

#include <condition_variable>
#include <cstdint>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "../../include/rmvd_b200.h"

namespace rmvd {
namespace {

// four mask values per thread: one 16-byte load (512 B per warp), one 4-byte store (128 B per warp)
__global__ void __launch_bounds__(256) mask_pack_kernel(const float* __restrict__ mask, uint8_t* __restrict__ out, size_t n4, size_t n) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 m = __ldcs(reinterpret_cast<const float4*>(mask) + i);
    const unsigned v = (m.x != 0.f ? 1u : 0u) | (m.y != 0.f ? 0x100u : 0u) | (m.z != 0.f ? 0x10000u : 0u) | (m.w != 0.f ? 0x1000000u : 0u);
    reinterpret_cast<unsigned*>(out)[i] = v;
  }
  if (blockIdx.x == 0 && threadIdx.x < (unsigned)(n - 4 * n4)) {
    const size_t i = 4 * n4 + threadIdx.x;
    out[i] = mask[i] != 0.f ? 1 : 0;
  }
}

// one bit per element: word j, bit i = mask[32 j + i] != 0 (a warp reads 128 contiguous bytes and writes one word per trip)
__global__ void __launch_bounds__(256) mask_pack_bits_kernel(const float* __restrict__ mask, unsigned* __restrict__ out, size_t words, size_t n) {
  const int lane = threadIdx.x & 31;
  const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
  for (size_t j = warp; j < words; j += nwarps) {
    const size_t i = j * 32 + lane;
    const float m = i < n ? __ldcs(mask + i) : 0.f;
    const unsigned bits = __ballot_sync(0xffffffffu, m != 0.f);
    if (lane == 0) out[j] = bits;
  }
}

// dst[i] = (float)src[i] for src in {0, 1} (any byte value converts exactly); streaming stores where dst is 16-byte aligned
void expand_range(const uint8_t* src, float* dst, size_t n) {
  size_t i = 0;
  while (i < n && (reinterpret_cast<uintptr_t>(dst + i) & 15u)) { dst[i] = (float)src[i]; ++i; }
  const __m128i zero = _mm_setzero_si128();
  for (; i + 16 <= n; i += 16) {
    const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i));
    const __m128i lo = _mm_unpacklo_epi8(b, zero), hi = _mm_unpackhi_epi8(b, zero);
    _mm_stream_ps(dst + i, _mm_cvtepi32_ps(_mm_unpacklo_epi16(lo, zero)));
    _mm_stream_ps(dst + i + 4, _mm_cvtepi32_ps(_mm_unpackhi_epi16(lo, zero)));
    _mm_stream_ps(dst + i + 8, _mm_cvtepi32_ps(_mm_unpacklo_epi16(hi, zero)));
    _mm_stream_ps(dst + i + 12, _mm_cvtepi32_ps(_mm_unpackhi_epi16(hi, zero)));
  }
  for (; i < n; ++i) dst[i] = (float)src[i];
  _mm_sfence();
}

// dst[i] = bit i of the packed words (i in [0, n), n counted from a byte boundary of `src`): one table entry of eight floats per byte
struct BitLut {
  alignas(16) float f[256][8];
  BitLut() {
    for (int b = 0; b < 256; ++b)
      for (int k = 0; k < 8; ++k) f[b][k] = (float)((b >> k) & 1);
  }
};

void expand_bits_range(const uint8_t* src, float* dst, size_t n) {
  static const BitLut lut;
  const bool aligned = (reinterpret_cast<uintptr_t>(dst) & 15u) == 0;
  size_t i = 0;
  if (aligned) {
    for (; i + 8 <= n; i += 8) {
      const float* e = lut.f[src[i >> 3]];
      _mm_stream_ps(dst + i, _mm_load_ps(e));
      _mm_stream_ps(dst + i + 4, _mm_load_ps(e + 4));
    }
  } else {
    for (; i + 8 <= n; i += 8) {
      const float* e = lut.f[src[i >> 3]];
      _mm_storeu_ps(dst + i, _mm_load_ps(e));
      _mm_storeu_ps(dst + i + 4, _mm_load_ps(e + 4));
    }
  }
  for (; i < n; ++i) dst[i] = (float)((src[i >> 3] >> (i & 7)) & 1);
  _mm_sfence();
}

// Worker pool of the expansion: created on first use, at most kMaxWorkers threads, one job at a time (jobs arrive from one CUDA
// callback thread or from synchronous callers; a mutex serialises them).
constexpr int kMaxWorkers = 32;

class ExpandPool {
 public:
  static ExpandPool& get() {
    static ExpandPool* p = new ExpandPool();   // never destroyed: CUDA callbacks may still run at process exit
    return *p;
  }

  void run(const uint8_t* src, float* dst, size_t n, int threads, bool bits = false) {
    if (threads < 1) threads = 1;
    if (threads > kMaxWorkers) threads = kMaxWorkers;
    const size_t min_part = (size_t)1 << 16;   // below 64 K elements per part a thread hand-off costs more than it saves
    size_t parts = (n + min_part - 1) / min_part;
    if (parts < 1) parts = 1;
    if (parts > (size_t)threads) parts = (size_t)threads;
    if (parts == 1) { bits ? expand_bits_range(src, dst, n) : expand_range(src, dst, n); return; }
    std::lock_guard<std::mutex> job(job_mutex_);
    {
      std::unique_lock<std::mutex> lk(m_);
      if (pid_ != getpid()) {                          // forked child: the parent's worker threads do not exist here
        for (auto& t : workers_) t.detach();
        workers_.clear();
        pid_ = getpid();
      }
      try {
        while ((int)workers_.size() < (int)parts - 1) {
          const int id = (int)workers_.size();
          workers_.emplace_back([this, id] { worker(id); });
        }
      } catch (...) {                                  // no more threads to be had: nothing throws across the C ABI, fewer parts do
        parts = workers_.size() + 1;
      }
      if (parts == 1) {
        lk.unlock();
        bits ? expand_bits_range(src, dst, n) : expand_range(src, dst, n);
        return;
      }
      src_ = src; dst_ = dst; n_ = n; bits_ = bits; parts_ = (int)parts; pending_ = (int)parts - 1;
      ++generation_;
    }
    cv_.notify_all();
    part(0);
    std::unique_lock<std::mutex> lk(m_);
    done_.wait(lk, [this] { return pending_ == 0; });
  }

 private:
  void part(int k) {
    // parts are multiples of 16 elements so that every part but the first keeps the destination's alignment phase
    const size_t per = ((n_ + parts_ - 1) / parts_ + 15) & ~(size_t)15;       // multiple of 16: whole bytes of a bit-packed source too
    const size_t b = per * k, e = b + per < n_ ? b + per : n_;
    if (b >= e) return;
    if (bits_) expand_bits_range(src_ + (b >> 3), dst_ + b, e - b);
    else expand_range(src_ + b, dst_ + b, e - b);
  }

  void worker(int id) {
    unsigned long long seen = 0;
    for (;;) {
      int parts;
      {
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [&] { return generation_ != seen; });
        seen = generation_;
        parts = parts_;
      }
      if (id + 1 < parts) {
        part(id + 1);
        std::lock_guard<std::mutex> lk(m_);
        if (--pending_ == 0) done_.notify_one();
      }
    }
  }

  std::mutex job_mutex_, m_;
  std::condition_variable cv_, done_;
  std::vector<std::thread> workers_;
  const uint8_t* src_ = nullptr;
  float* dst_ = nullptr;
  size_t n_ = 0;
  bool bits_ = false;
  int parts_ = 0, pending_ = 0;
  unsigned long long generation_ = 0;
  pid_t pid_ = getpid();
};

struct ExpandJob {
  const uint8_t* src;
  float* dst;
  size_t n;
  int threads;
  bool bits;
};

void CUDART_CB expand_callback(void* p) {
  ExpandJob* j = static_cast<ExpandJob*>(p);
  ExpandPool::get().run(j->src, j->dst, j->n, j->threads, j->bits);
  delete j;
}

}  // namespace
}  // namespace rmvd

extern "C" int rmvd_mask_pack_u8(const float* mask, uint8_t* packed, size_t n, void* stream) {
  if (!mask || !packed) return RMVD_E_NULL;
  if (n == 0) return RMVD_OK;
  if (reinterpret_cast<uintptr_t>(mask) & 15u || reinterpret_cast<uintptr_t>(packed) & 3u) return RMVD_E_SHAPE;
  const size_t n4 = n / 4;
  size_t blocks = (n4 + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  rmvd::mask_pack_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(mask, packed, n4, n);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_host_expand_mask(const uint8_t* packed_host, float* mask_host, size_t n, int threads) {
  if (!packed_host || !mask_host) return RMVD_E_NULL;
  if (n == 0) return RMVD_OK;
  rmvd::ExpandPool::get().run(packed_host, mask_host, n, threads);
  return RMVD_OK;
}

extern "C" int rmvd_host_expand_mask_async(const uint8_t* packed_host, float* mask_host, size_t n, int threads, void* stream) {
  if (!packed_host || !mask_host) return RMVD_E_NULL;
  if (n == 0) return RMVD_OK;
  rmvd::ExpandJob* j = new (std::nothrow) rmvd::ExpandJob{packed_host, mask_host, n, threads, false};
  if (!j) return (int)cudaErrorMemoryAllocation;
  const cudaError_t e = cudaLaunchHostFunc((cudaStream_t)stream, rmvd::expand_callback, j);
  if (e != cudaSuccess) delete j;
  return (int)e;
}

extern "C" int rmvd_mask_pack_bits(const float* mask, uint32_t* packed, size_t n, void* stream) {
  if (!mask || !packed) return RMVD_E_NULL;
  if (n == 0) return RMVD_OK;
  const size_t words = (n + 31) / 32;
  size_t blocks = (words + 7) / 8;                     // 8 warps per CTA, one word per warp and trip
  if (blocks > 148 * 16) blocks = 148 * 16;
  rmvd::mask_pack_bits_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(mask, packed, words, n);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_host_expand_mask_bits(const uint32_t* packed_host, float* mask_host, size_t n, int threads) {
  if (!packed_host || !mask_host) return RMVD_E_NULL;
  if (n == 0) return RMVD_OK;
  rmvd::ExpandPool::get().run(reinterpret_cast<const uint8_t*>(packed_host), mask_host, n, threads, true);
  return RMVD_OK;
}

extern "C" int rmvd_host_expand_mask_bits_async(const uint32_t* packed_host, float* mask_host, size_t n, int threads, void* stream) {
  if (!packed_host || !mask_host) return RMVD_E_NULL;
  if (n == 0) return RMVD_OK;
  rmvd::ExpandJob* j = new (std::nothrow) rmvd::ExpandJob{reinterpret_cast<const uint8_t*>(packed_host), mask_host, n, threads, true};
  if (!j) return (int)cudaErrorMemoryAllocation;
  const cudaError_t e = cudaLaunchHostFunc((cudaStream_t)stream, rmvd::expand_callback, j);
  if (e != cudaSuccess) delete j;
  return (int)e;
}
