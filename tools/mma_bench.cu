// Micro-benchmark: cycles per tcgen05.mma (M128, K16, bf16) for several shared-memory operand layouts.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mma_bench tools/mma_bench.cu
#include <cstdio>
#include <cstdlib>
#include "../perseus_b200/csrc/common.cuh"
namespace pb200 { void set_error(const char*, ...) {} int cuda_fail(cudaError_t, const char*, const char*, int) { return 2; } }
using namespace pb200;

struct Variant { int n; int a_mode; int a_off; int a_lbo; int a_sbo; int b_mode; const char* name; int vary = 0; };
// mode 0 = interleaved(no swizzle), 1 = SW128

__global__ void __launch_bounds__(128, 1) bench(int N, int a_mode, uint32_t a_off, uint32_t a_lbo, uint32_t a_sbo,
                                                 int b_mode, int iters, long long* out, int vary) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(&slot, 256); tmem_relinquish(); }
  // zero the operands (denormal/NaN garbage could slow nothing, but keep it clean)
  for (uint32_t i = threadIdx.x; i < 100 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw)[i] = 0;
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t(N) >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t a_addr = base + a_off, b_addr = base + 49152;
    uint64_t ad = a_mode ? umma_desc_kmajor(a_addr, 128) : umma_desc_interleaved(a_addr, a_lbo, a_sbo);
    uint64_t bd = b_mode ? umma_desc_kmajor(b_addr, 128) : umma_desc_interleaved(b_addr, 128, 9216);
    // warm
    for (int i = 0; i < 8; ++i) umma_bf16(tm, ad, bd, idesc, 1);
    umma_commit(&bar); mbar_wait(&bar, 0);
    long long t0 = clock64();
    if (vary == 0) {
      for (int i = 0; i < iters; ++i) umma_bf16(tm, ad + (i & 3) * 2 * (a_mode ? 1 : 0), bd, idesc, 1);
    } else {
      // unrolled, compile-time offsets (no issue-side arithmetic): vary bit0 = move A each MMA, bit1 = move B
      const uint64_t sa = (vary & 1) ? 132 : 0, sb = (vary & 2) ? 16 : 0;
      for (int i = 0; i < iters; i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) umma_bf16(tm, ad + sa * j + (j & 1) * 2, bd + sb * j, idesc, 1);
      }
    }
    umma_commit(&bar); mbar_wait(&bar, 1);
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 256); }
}

int main() {
  long long* d; cudaMalloc(&d, 148 * 8);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  Variant vs[] = {
    {64, 1, 0, 0, 0, 1, "N64  A=SW128           B=SW128"},
    {128, 1, 0, 0, 0, 1, "N128 A=SW128           B=SW128"},
    {256, 1, 0, 0, 0, 1, "N256 A=SW128           B=SW128"},
    {64, 0, 0, 4224, 128, 0, "N64  A=nosw aligned    lbo4224 B=nosw"},
    {64, 0, 16, 4224, 128, 0, "N64  A=nosw +16B       lbo4224 B=nosw"},
    {64, 0, 64, 4224, 128, 0, "N64  A=nosw +64B       lbo4224 B=nosw"},
    {64, 0, 0, 16, 128, 0, "N64  A=nosw lbo16 (stem) aligned B=nosw"},
    {64, 0, 32, 16, 128, 0, "N64  A=nosw lbo16 (stem) +32B   B=nosw"},
    {64, 0, 0, 2048, 128, 0, "N64  A=nosw aligned    lbo2048 B=nosw"},
    {64, 0, 0, 128, 256, 0, "N64  A=nosw lbo128 sbo256 (dense core pairs) B=nosw"},
    {64, 0, 0, 4224, 128, 1, "N64  A=nosw aligned    B=SW128"},
    {64, 1, 0, 0, 0, 0, "N64  A=SW128           B=nosw"},
    {128, 0, 0, 4224, 128, 0, "N128 A=nosw aligned    B=nosw"},
    {128, 0, 16, 4224, 128, 0, "N128 A=nosw +16B       B=nosw"},
    {64, 0, 0, 16, 128, 0, "N64  nosw lbo16: A moves", 1},
    {64, 0, 0, 16, 128, 0, "N64  nosw lbo16: B moves", 2},
    {64, 0, 0, 16, 128, 0, "N64  nosw lbo16: A and B move", 3},
    {64, 0, 0, 4224, 128, 0, "N64  nosw lbo4224: A and B move", 3},
    {128, 0, 0, 16, 128, 0, "N128 nosw lbo16: A and B move", 3},
    {64, 1, 0, 0, 0, 1, "N64  SW128: A and B move", 3},
    {32, 1, 0, 0, 0, 1, "N32  A=SW128           B=SW128"},
    {16, 1, 0, 0, 0, 1, "N16  A=SW128           B=SW128"},
  };
  const int iters = 2000;
  for (auto& v : vs) {
    for (int grid : {1, 148}) {
      bench<<<grid, 128, 200 * 1024>>>(v.n, v.a_mode, v.a_off, v.a_lbo, v.a_sbo, v.b_mode, iters, d, v.vary);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("%s: CUDA error %s\n", v.name, cudaGetErrorString(e)); return 1; }
      long long h[148]; cudaMemcpy(h, d, grid * 8, cudaMemcpyDeviceToHost);
      long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
      printf("%-55s grid %3d: %.1f cycles/MMA\n", v.name, grid, double(mx) / iters);
    }
  }
  return 0;
}
