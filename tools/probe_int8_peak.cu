// tools/probe_int8_peak.cu -- measured int8 tensor-core peak of this GPU, with our own tcgen05.mma.kind::i8 issue loop.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/probe_int8_peak tools/probe_int8_peak.cu && /tmp/probe_int8_peak
//
// bench.py's roofline for tc_sign_gemm_kernel needs an int8 denominator; MEASURED_PEAKS.json only carries a cuBLAS bf16
// figure ("int8 = 2 x bf16" is nominal, not measured).  This program keeps the operands RESIDENT: one CTA per SM, A
// (128 x 32 B) and B (256 x 32 B) tiles written to shared memory once, and one thread issuing M=128, N=256, K=32
// instructions back to back into two 256-column accumulators -- no loads, no epilogue, nothing but the tensor pipe.
// It prints the rate for a short burst and for a run of a few seconds (power cap reached), dense, one JSON line each.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                                              \
  do {                                                                                                     \
    cudaError_t e_ = (x);                                                                                  \
    if (e_ != cudaSuccess) {                                                                               \
      printf("CUDA error %s at line %d: %s\n", cudaGetErrorName(e_), __LINE__, cudaGetErrorString(e_));    \
      exit(2);                                                                                             \
    }                                                                                                      \
  } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t idesc_i8(int m, int n) {
  return (2u << 4) | (1u << 7) | (0u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// iters x (4 k-steps x 2 accumulators) instructions of 128 x 256 x 32 per CTA
__global__ void __launch_bounds__(128) peak_kernel(int iters, int n_cols) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sa = smem;                // 128 rows x 128 B
  uint8_t* sb = smem + 128 * 128;    // 256 rows x 128 B
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < (128 + 256) * 128 / 4; i += 128) ((uint32_t*)smem)[i] = 0x01ff0201u * (uint32_t)(i + 1);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  if (tid == 0) {
    const uint32_t id = idesc_i8(128, n_cols);
    uint32_t parity = 0;
    for (int it = 0; it < iters; it++) {
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const uint64_t ad = desc_sw128(smem_u32(sa) + 32 * k), bd = desc_sw128(smem_u32(sb) + 32 * k);
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(ad), "l"(bd), "r"(id), "r"(1)
            : "memory");
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem + 256), "l"(ad), "l"(bd), "r"(id), "r"(1)
            : "memory");
      }
      if ((it & 15) == 15 || it == iters - 1) {  // bound the queue: wait for everything issued so far
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        uint32_t ok = 0;
        const long long t0 = clock64();
        while (!ok) {
          asm volatile(
              "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
              : "=r"(ok)
              : "r"(smem_u32(&bar)), "r"(parity)
              : "memory");
          if (clock64() - t0 > 4000000000LL) __trap();
        }
        parity ^= 1;
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

int main() {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    printf("no CUDA device\n");
    return 2;
  }
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  const size_t smem = (128 + 256) * 128 + 1024;
  CK(cudaFuncSetAttribute(peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int n_cols : {256, 128}) {
    for (int iters : {2000, 20000, 4000000}) {
      peak_kernel<<<sms, 128, smem>>>(200, n_cols);  // warm-up
      CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(e0));
      peak_kernel<<<sms, 128, smem>>>(iters, n_cols);
      CK(cudaEventRecord(e1));
      CK(cudaDeviceSynchronize());
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      const double macs = (double)sms * iters * 8.0 * 128.0 * n_cols * 32.0;
      printf("{\"probe\": \"int8_peak\", \"kind\": \"tcgen05.mma.cta_group::1.kind::i8 dense, operands resident in shared memory\", "
             "\"M\": 128, \"N\": %d, \"K\": 32, \"ctas\": %d, \"instructions_per_cta\": %d, \"ms\": %.3f, \"int8_tops\": %.1f}\n",
             n_cols, sms, iters * 8, ms, 2.0 * macs / (ms * 1e-3) / 1e12);
    }
  }
  return 0;
}
