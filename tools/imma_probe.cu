// imma_probe.cu - latency / issue rate of mma.sync.m16n8k32.s32.u8.s8 (IMMA.16832) on this GPU, the tile math of the Q8
// decoder kernel (dec_mega.cu, consume_q8).  nvcc -gencode arch=compute_100a,code=sm_100a -o build/imma_probe tools/imma_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int CHAINS>
__global__ void __launch_bounds__(256) k(int iters, int *out, long long *cyc) {
    int d[CHAINS][4];
    #pragma unroll
    for (int c = 0; c < CHAINS; c++) for (int i = 0; i < 4; i++) d[c][i] = 0;
    uint32_t a0 = threadIdx.x * 0x01010101u, a2 = a0 ^ 0x5a5a5a5au, b0 = threadIdx.x + 7, b1 = b0 * 3;
    const uint32_t z = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        #pragma unroll
        for (int c = 0; c < CHAINS; c++)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(d[c][0]), "+r"(d[c][1]), "+r"(d[c][2]), "+r"(d[c][3]) : "r"(a0), "r"(z), "r"(a2), "r"(z), "r"(b0), "r"(b1));
    }
    const long long t1 = clock64();
    int s = 0;
    #pragma unroll
    for (int c = 0; c < CHAINS; c++) s += d[c][0] + d[c][1] + d[c][2] + d[c][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int CHAINS>
void run(int warps) {
    int *out; long long *cyc;
    cudaMalloc(&out, 148 * 256 * 4); cudaMallocManaged(&cyc, 8);
    const int iters = 4096;
    k<CHAINS><<<148, warps * 32>>>(iters, out, cyc);
    cudaDeviceSynchronize();
    k<CHAINS><<<148, warps * 32>>>(iters, out, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%d warps/SM x %d chains: %.1f cycles per IMMA per warp, %.2f IMMA per cycle per SM  (%s)\n", warps, CHAINS,
           (double)*cyc / (iters * CHAINS), (double)iters * CHAINS * warps / (double)*cyc, cudaGetErrorString(e));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<1>(1); run<1>(4); run<1>(8); run<2>(8); run<4>(8); run<8>(8); run<4>(4); run<8>(1);
    return 0;
}
