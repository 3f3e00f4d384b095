// Issue rate of tcgen05.mma (cta_group::1, kind::f16, bf16 in / fp32 out) with both operands in shared memory:
// clocks per M128 x N x K16 instruction for N = 32 .. 256, K-major SW128 operands that stay resident (no TMA traffic).
// What it answers: are the small-N layers (Cout = 32 / 64) bound by the tensor array (128*N*16 MACs / 4096 per clk) or by
// the operand fetch from shared memory (A: 4 KB per instruction whatever N is)?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/build/umma_bench scripts/umma_bench.cu && scripts/build/umma_bench
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t kmajor_sw128(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ void umma(uint32_t d_tmem, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n"
               ::"r"(d_tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}

template <int N, int AMODE>      // AMODE 0: the same A tile every time, 1: four A tiles in rotation (taps of a conv)
__global__ void __launch_bounds__(128, 1) umma_kernel(int iters, long long* clocks) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  for (int i = threadIdx.x; i < (4 * 16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  const uint32_t bar_a = smem_u32(&bar);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  if (threadIdx.x == 0) {
    const uint64_t da0 = kmajor_sw128(smem_u32(smem)), db0 = kmajor_sw128(smem_u32(smem) + 4 * 16384);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint64_t da = da0 + (uint64_t)((AMODE ? (it & 3) * 16384 : 0) >> 4);
#pragma unroll
      for (int k = 0; k < 4; ++k) umma(tmem, da + (uint64_t)(k * 2), db0 + (uint64_t)(k * 2), idesc, 1u);
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_a) : "memory");
    uint32_t ok = 0;
    while (!ok) {
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(bar_a), "r"(0) : "memory");
    }
    const long long t1 = clock64();
    if (blockIdx.x == 0) clocks[0] = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
  }
}

template <int N, int AMODE>
int run(long long* dclk) {
  const int iters = 4096, smem = 4 * 16384 + 32768 + 1024;
  CK(cudaFuncSetAttribute(umma_kernel<N, AMODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  for (int rep = 0; rep < 2; ++rep) umma_kernel<N, AMODE><<<148, 128, smem>>>(iters, dclk);
  CK(cudaDeviceSynchronize());
  long long c = 0;
  CK(cudaMemcpy(&c, dclk, sizeof(c), cudaMemcpyDeviceToHost));
  const double per = (double)c / (iters * 4.0);
  printf("M128 x N%-3d x K16, A %-12s %7.1f clk per MMA   ideal (4096 MAC/clk) %5.1f   operand bytes per clk %6.1f\n", N,
         AMODE ? "4 tiles" : "1 tile", per, 128.0 * N * 16 / 4096.0, (128 * 32 + N * 32) / per);
  return 0;
}

int main() {
  long long* dclk;
  CK(cudaMalloc(&dclk, sizeof(long long)));
  if (run<32, 0>(dclk) || run<64, 0>(dclk) || run<128, 0>(dclk) || run<256, 0>(dclk)) return 1;
  if (run<32, 1>(dclk) || run<64, 1>(dclk) || run<128, 1>(dclk) || run<256, 1>(dclk)) return 1;
  return 0;
}
