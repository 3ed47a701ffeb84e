// Host interface of the tcgen05 weight-gradient GEMM (wgrad_umma.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace pb200 {

struct alignas(64) WgradMaps {
  CUtensorMap x[4];  // layer input, the forward's im2col views: [0] stride 1, [hp*2+wp] parity views for stride 2
  CUtensorMap dz;    // gradient of the conv output, [M = B*Ho*Wo][Cout]
};

struct WgradParams {
  int num_items;      // num_pairs * n_tiles * ksplit
  int num_units;      // R*R * Cin/64 (tap, 64-channel block) rows units
  int num_pairs;      // ceil(num_units / 2): two units share one 128-row MMA
  int n_tiles;        // Cout / BLOCK_N
  int ksplit;         // slices of the pixel (K) dimension
  int kblocks_total;  // ceil(M / 128)
  int cblocks;        // Cin / 64
  int S, pad, stride;
  int tiles_per_img, imgs_per_tile, bh;
  int Cin, Cout;
  float* dw;          // fp32 [R*R][Cin][Cout], accumulated with atomics (zeroed by the caller)
};

struct WgradLaunch {
  WgradMaps maps;
  WgradParams p;
  int block_n;
  int grid;
  double flops;
};

int wgrad_umma_init();
// x: NHWC bf16 [B][Hin][Win][Cin]; dz: NHWC bf16 [B][Ho][Wo][Cout]; dw_scratch: fp32 [R*R][Cin][Cout]
int build_wgrad(WgradLaunch* L, const void* x, const void* dz, float* dw_scratch, int B, int Hin, int Win, int Cin,
                int Cout, int R, int stride, int pad);
int launch_wgrad(const WgradLaunch& L, cudaStream_t st);

}  // namespace pb200
