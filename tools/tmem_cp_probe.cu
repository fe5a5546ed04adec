// tmem_cp_probe.cu - how does tcgen05.cp.128x256b place the bytes of a NO-SWIZZLE shared-memory
// "matrix" into tensor memory?  Fills 8 KB of smem with word indices, copies 4 KB chunks with two
// descriptor settings, reads the lanes back with tcgen05.ld.32x32b.x8 and prints lane -> word map.
// (Groundwork for dec_mega's TMEM tile ring.)  nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(uint32_t *out, uint32_t lbo, uint32_t sbo) {
    extern __shared__ __align__(1024) uint32_t sm[];
    __shared__ uint32_t tbase;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 4096; i += blockDim.x) sm[i] = i;           // 16 KB of word indices
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tbase)), "r"(32u));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&bar)), "r"(1u));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy smem writes -> async proxy (tcgen05.cp)
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t t0 = tbase;
    if (tid == 0) {
        const uint64_t desc = (uint64_t)((s32(sm) & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
        asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(t0), "l"(desc) : "memory");
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)) : "memory");
    }
    // everyone waits for the copy
    unsigned ok = 0;
    while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(s32(&bar)), "r"(0u) : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (warp < 4) {
        uint32_t r[8];
        const uint32_t taddr = t0 + ((uint32_t)(warp * 32) << 16);
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 8; i++) out[tid * 8 + i] = r[i];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(t0), "r"(32u));
}

int main() {
    uint32_t *d, h[128 * 8];
    cudaMalloc(&d, sizeof(h));
    const uint32_t cfg[3][2] = {{2048, 128}, {128, 2048}, {4096, 256}};
    for (int c = 0; c < 3; c++) {
        cudaMemset(d, 0xff, sizeof(h));
        probe<<<1, 128, 16384>>>(d, cfg[c][0], cfg[c][1]);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("cfg %d: %s\n", c, cudaGetErrorString(e)); return 1; }
        cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        printf("LBO=%u SBO=%u (byte offsets of the 8 columns of lanes 0,1,7,8,9,31,32,127):\n", cfg[c][0], cfg[c][1]);
        const int lanes[8] = {0, 1, 7, 8, 9, 31, 32, 127};
        for (int li = 0; li < 8; li++) {
            printf("  lane %3d:", lanes[li]);
            for (int i = 0; i < 8; i++) printf(" %5u", h[lanes[li] * 8 + i] * 4);
            printf("\n");
        }
    }
    return 0;
}
