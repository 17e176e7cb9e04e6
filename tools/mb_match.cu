// Dev microbenchmark (not product): latency / throughput of MATCH.ANY on sm_100a.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gpurun_out/mb_match tools/mb_match.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int ILP>
__global__ void k(uint32_t *out, int iters, long long *cyc, uint32_t mask) {
    uint32_t v[ILP];
    const uint32_t lane = threadIdx.x & 31u;
#pragma unroll
    for (int j = 0; j < ILP; ++j) v[j] = (lane * 2654435761u + j * 977u + blockIdx.x) >> 27;
    __syncwarp();
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < ILP; ++j) {
            uint32_t m = __match_any_sync(0xffffffffu, v[j]);
            v[j] = ((v[j] + __popc(m)) * 7u + lane) & mask;   // next key depends on the result
        }
    }
    long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int j = 0; j < ILP; ++j) acc += v[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int ILP>
static void run(int warps_per_block, int blocks, uint32_t mask = 31u) {
    uint32_t *out; long long *cyc, h;
    cudaMalloc(&out, sizeof(uint32_t) * blocks * warps_per_block * 32);
    cudaMalloc(&cyc, sizeof(long long));
    const int iters = 2000;
    k<ILP><<<blocks, warps_per_block * 32>>>(out, iters, cyc, mask);
    k<ILP><<<blocks, warps_per_block * 32>>>(out, iters, cyc, mask);
    cudaDeviceSynchronize();
    cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("mask=%x ILP=%d warps/block=%d blocks=%d: %.1f cycles per match per warp, %.1f cycles per match per SM\n", mask, ILP, warps_per_block,
           blocks, (double)h / iters / ILP, (double)h / iters / ILP / warps_per_block / ((blocks + 147) / 148));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<1>(1, 148); run<2>(1, 148); run<4>(1, 148); run<8>(1, 148);
    run<1>(4, 148); run<4>(4, 148); run<1>(8, 148); run<4>(8, 148); run<4>(16, 148);
    for (uint32_t mk : {0u, 1u, 3u, 7u, 15u, 31u, 0x3fffu}) { run<1>(1, 148, mk); run<4>(6, 148, mk); }
    return 0;
}
