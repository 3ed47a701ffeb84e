// Probe (not part of the library): does a SWIZZLE_128B K-major tcgen05 operand still read correctly when its start
// address is shifted by whole 128-byte rows inside the 1024-byte swizzle atom (a filter-tap shift over an NHWC patch)?
// And what does one M128 N64 K16 MMA cost for aligned / row-shifted SW128 rows and for no-swizzle rows?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I perseus_b200/csrc tools/probe_umma_shift.cu -o gpurun_out/probe
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "common.cuh"

using namespace pb200;

constexpr int kRowsA = 160;

__device__ __forceinline__ float fa(int i, int k) { return static_cast<float>((i * 3 + k) % 7 - 3); }
__device__ __forceinline__ float fb(int n, int k) { return static_cast<float>((n + 2 * k) % 5 - 2); }

// mode: 0 = base_offset 0, 1 = base_offset = (start >> 7) & 7
__global__ void __launch_bounds__(128) probe_kernel(int shift, int mode, float* out) {
  __shared__ __align__(1024) uint8_t sA[kRowsA * 128];
  __shared__ __align__(1024) uint8_t sB[64 * 128];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int e = threadIdx.x; e < kRowsA * 64; e += 128) {
    const int i = e / 64, k = e % 64;
    const int off = i * 128 + (((k / 8) ^ (i % 8)) * 16) + (k % 8) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sA + off) = __float2bfloat16(fa(i, k));
  }
  for (int e = threadIdx.x; e < 64 * 64; e += 128) {
    const int n = e / 64, k = e % 64;
    const int off = n * 128 + (((k / 8) ^ (n % 8)) * 16) + (k % 8) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sB + off) = __float2bfloat16(fb(n, k));
  }
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_slot, 64);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_slot;
  if (warp == 0) {
    if (elect_one()) {
      const uint32_t a_addr = smem_u32(sA) + shift * 128;
      const uint32_t bo = mode ? ((a_addr >> 7) & 7u) : 0u;
      const uint32_t a_hi = desc_hi_swizzled(128) | (bo << 17);
      const uint32_t b_hi = desc_hi_swizzled(128);
      const uint32_t a_lo = desc_lo(a_addr, 16), b_lo = desc_lo(smem_u32(sB), 16);
      for (int k = 0; k < 4; ++k)
        umma_bf16_split(tm, a_lo + 2 * k, a_hi, b_lo + 2 * k, b_hi, umma_idesc_bf16(128, 64), k ? 1u : 0u);
      umma_commit(&bar);
    }
    __syncwarp();
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int c = 0; c < 2; ++c) {
    uint32_t v[32];
    tmem_ld_32x32(tm + (static_cast<uint32_t>(warp * 32) << 16) + c * 32, v);
    tmem_ld_wait_regs(v);
    for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * 64 + c * 32 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tm, 64);
  }
}

// variant: 0 SW128 aligned, 1 SW128 shifted one row, 2 no-swizzle (16-byte row pitch, SBO 128, LBO 16) aligned,
//          3 no-swizzle shifted 16 B, 4 no-swizzle shifted 48 B; N = 64 or 128
__global__ void __launch_bounds__(128) time_kernel(int variant, int N, int reps, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  for (int e = threadIdx.x; e < 48 * 1024 / 4; e += 128) reinterpret_cast<uint32_t*>(smem)[e] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_slot, 128);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_slot;
  if (warp == 0) {
    if (elect_one()) {
      const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
      uint32_t a_lo, a_hi;
      if (variant < 2) {
        const uint32_t a_addr = base + (variant == 1 ? 128 : 0);
        a_lo = desc_lo(a_addr, 16);
        a_hi = desc_hi_swizzled(128) | (((a_addr >> 7) & 7u) << 17);
      } else {
        const uint32_t a_addr = base + (variant == 3 ? 16 : (variant == 4 ? 48 : 0));
        a_lo = desc_lo(a_addr, 16);
        a_hi = desc_hi_interleaved(128);
      }
      const uint32_t b_lo = desc_lo(base + 24 * 1024, 16), b_hi = desc_hi_swizzled(128);
      const uint32_t idesc = umma_idesc_bf16(128, N);
      const long long t0 = clock64();
      for (int r = 0; r < reps; ++r) umma_bf16_split(tm, a_lo, a_hi, b_lo, b_hi, idesc, 1u);
      umma_commit(&bar);
      mbar_wait(&bar, 0);
      *cycles = clock64() - t0;
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tm, 128);
  }
}

int main() {
  float* d_out;
  cudaMalloc(&d_out, 128 * 64 * 4);
  std::vector<float> h(128 * 64);
  const int shifts[] = {0, 1, 3, 8, 9, 13};
  for (int mode = 0; mode < 2; ++mode)
    for (int shift : shifts) {
      cudaMemset(d_out, 0xff, 128 * 64 * 4);
      probe_kernel<<<1, 128>>>(shift, mode, d_out);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) {
        printf("shift %d mode %d: CUDA error %s\n", shift, mode, cudaGetErrorString(e));
        return 1;
      }
      cudaMemcpy(h.data(), d_out, 128 * 64 * 4, cudaMemcpyDeviceToHost);
      int bad = 0;
      double maxerr = 0;
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 64; ++n) {
          double want = 0;
          for (int k = 0; k < 64; ++k)
            want += static_cast<double>(((m + shift) * 3 + k) % 7 - 3) * static_cast<double>((n + 2 * k) % 5 - 2);
          const double err = fabs(want - h[m * 64 + n]);
          if (err > 1e-3) ++bad;
          if (err > maxerr) maxerr = err;
        }
      printf("SW128 row shift %2d  base_offset mode %d: %d / 8192 wrong, max err %.1f\n", shift, mode, bad, maxerr);
    }
  long long* d_cyc;
  cudaMalloc(&d_cyc, 8);
  cudaFuncSetAttribute(time_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 50 * 1024);
  const char* names[] = {"SW128 aligned", "SW128 +1 row", "no-swizzle aligned", "no-swizzle +16 B", "no-swizzle +48 B"};
  for (int N : {64, 128})
    for (int v = 0; v < 5; ++v) {
      long long c = 0;
      for (int rep = 0; rep < 2; ++rep) {
        time_kernel<<<1, 128, 50 * 1024>>>(v, N, 4000, d_cyc);
        cudaDeviceSynchronize();
        cudaMemcpy(&c, d_cyc, 8, cudaMemcpyDeviceToHost);
      }
      printf("M128 N%-3d K16  A %-20s: %.1f cycles / MMA\n", N, names[v], static_cast<double>(c) / 4000.0);
    }
  return 0;
}
