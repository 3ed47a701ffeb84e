// Micro-benchmark: shared-memory fill rate of TMA tile loads vs 1-D bulk copies, per SM, for the operand shapes of
// the conv kernels.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/tma_bench tools/tma_bench.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../perseus_b200/csrc/runtime.cu"
using namespace pb200;

constexpr int S = 6;  // loads in flight
struct Maps { CUtensorMap m; };

// mode 0: 4-D tile {64, w, h, 1} at moving coordinates; mode 1: 2-D tile {64, rows}; mode 2: 1-D bulk copy
__global__ void __launch_bounds__(32, 1) bench(const __grid_constant__ Maps maps, int mode, const uint8_t* src, int bytes,
                                               int iters, int span, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full[S];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* base = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  if (threadIdx.x == 0) {
    for (int i = 0; i < S; ++i) mbar_init(&full[i], 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t0 = clock64();
    for (int i = 0; i < iters + S; ++i) {
      const int s = i % S;
      if (i >= S) mbar_wait(&full[s], ((i / S) - 1) & 1);
      if (i < iters) {
        mbar_expect_tx(&full[s], bytes);
        const int j = (i * 7 + blockIdx.x * 13) % span;  // walk over the tensor so loads are not the same lines
        if (mode == 0) tma_load_4d(base + s * 32768, &maps.m, &full[s], 0, -1 + (j & 1), (j >> 1) % 28, j % 200);
        else if (mode == 1) tma_load_2d(base + s * 32768, &maps.m, &full[s], (j % 30) * 64, (j % 2) * 128);
        else if (mode == 2) bulk_load_1d(base + s * 32768, src + (size_t)j * bytes, bytes, &full[s]);
        else if (mode == 3) {  // two 1-D copies of bytes/2 per iteration on one barrier
          bulk_load_1d(base + s * 32768, src + (size_t)j * bytes, bytes / 2, &full[s]);
          bulk_load_1d(base + s * 32768 + bytes / 2, src + (size_t)j * bytes + bytes / 2, bytes / 2, &full[s]);
        } else {  // mode 4: four copies
          for (int q = 0; q < 4; ++q)
            bulk_load_1d(base + s * 32768 + q * (bytes / 4), src + (size_t)j * bytes + q * (bytes / 4), bytes / 4, &full[s]);
        }
      }
    }
    out[blockIdx.x] = clock64() - t0;
  }
}

static void run(const char* name, Maps& m, int mode, const uint8_t* src, int bytes, int span) {
  long long* d; cudaMalloc(&d, 148 * 8);
  const int iters = 2000;
  for (int grid : {1, 148}) {
    bench<<<grid, 32, 6 * 32768 + 1024>>>(m, mode, src, bytes, iters, span, d);
    cudaError_t e = cudaGetLastError(); if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: CUDA error %s\n", name, cudaGetErrorString(e)); exit(1); }
    long long h[148]; cudaMemcpy(h, d, grid * 8, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("%-62s grid %3d: %7.1f cyc/load  %6.1f B/cyc/SM\n", name, grid, double(mx) / iters, bytes * double(iters) / mx);
  }
  cudaFree(d);
}

int main() {
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * 32768 + 1024);
  uint8_t* buf; const size_t N = 512ull << 20; cudaMalloc(&buf, N); cudaMemset(buf, 0, N);
  Maps m;
  {  // layer2 activation NHWC [B=256][32][32][128] bf16, box {64, 32, 4, 1} -> 128 rows of 128 B (pixel pitch 256 B)
    uint64_t dims[4] = {128, 32, 32, 256}, str[3] = {256, 32 * 256, 32 * 32 * 256}; uint32_t box[4] = {64, 32, 4, 1};
    encode_tma_bf16(&m.m, buf, 4, dims, str, box, 128);
    run("A tile NHWC C=128 32x32: box{64,32,4} 16 KB (128 B rows)", m, 0, buf, 16384, 200);
  }
  {  // channel-blocked [B][C/64][32][32][64]: same box, pixel pitch 128 B (rows of a box row are contiguous)
    uint64_t dims[4] = {64, 32, 32, 512}, str[3] = {128, 32 * 128, 32 * 32 * 128}; uint32_t box[4] = {64, 32, 4, 1};
    encode_tma_bf16(&m.m, buf, 4, dims, str, box, 128);
    run("A tile blocked [C/64][H][W][64]: box{64,32,4} 16 KB", m, 0, buf, 16384, 200);
  }
  {  // layer3 activation NHWC [256][16][16][256], box {64,16,8,1}
    uint64_t dims[4] = {256, 16, 16, 256}, str[3] = {512, 16 * 512, 16 * 16 * 512}; uint32_t box[4] = {64, 16, 8, 1};
    encode_tma_bf16(&m.m, buf, 4, dims, str, box, 128);
    run("A tile NHWC C=256 16x16: box{64,16,8} 16 KB (128 B rows)", m, 0, buf, 16384, 200);
  }
  {  // weights [Cout=256][K=2304] bf16, box {64, 128}: 128 rows of 128 B at pitch 4608 B
    uint64_t dims[2] = {2304, 256}, str[1] = {4608}; uint32_t box[2] = {64, 128};
    encode_tma_bf16(&m.m, buf, 2, dims, str, box, 128);
    run("B tile [256][2304]: box{64,128} 16 KB (128 B rows, pitch 4608)", m, 1, buf, 16384, 60);
  }
  {
    uint64_t dims[2] = {2304, 256}, str[1] = {4608}; uint32_t box[2] = {64, 256};
    encode_tma_bf16(&m.m, buf, 2, dims, str, box, 128);
    run("B tile [256][2304]: box{64,256} 32 KB", m, 1, buf, 32768, 30);
  }
  {
    uint64_t dims[4] = {128, 32, 32, 256}, str[3] = {256, 32 * 256, 32 * 32 * 256}; uint32_t box[4] = {64, 32, 4, 1};
    encode_tma_bf16(&m.m, buf, 4, dims, str, box, 128);
    run("L2-hot A tile NHWC C=128: box{64,32,4} 16 KB", m, 0, buf, 16384, 8);
    uint64_t dims2[2] = {2304, 256}, str2[1] = {4608}; uint32_t box2[2] = {64, 256};
    encode_tma_bf16(&m.m, buf, 2, dims2, str2, box2, 128);
    run("L2-hot B tile box{64,256} 32 KB", m, 1, buf, 32768, 4);
    uint32_t box3[2] = {64, 128};
    encode_tma_bf16(&m.m, buf, 2, dims2, str2, box3, 128);
    run("L2-hot B tile box{64,128} 16 KB", m, 1, buf, 16384, 4);
  }
  run("L2-hot 1-D bulk copy 32 KB", m, 2, buf, 32768, 16);
  run("L2-hot 2 x 16 KB per iteration", m, 3, buf, 32768, 16);
  run("L2-hot 4 x 8 KB per iteration", m, 4, buf, 32768, 16);
  run("L2-hot 4 x 2 KB per iteration", m, 4, buf, 8192, 16);
  run("L2-hot 1-D bulk copy 16 KB", m, 2, buf, 16384, 16);
  run("1-D bulk copy 16 KB", m, 2, buf, 16384, 4000);
  run("1-D bulk copy 32 KB", m, 2, buf, 32768, 4000);
  run("1-D bulk copy 4 KB", m, 2, buf, 4096, 4000);
  return 0;
}
