// Store-rate probe for peer memory over NVLink (scripts/nvlink_probe.py): how many stores per
// second one B200 can push into another one's memory, by store size and address pattern.
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ uint64_t mix(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// mode 0: random 32-byte rows, one 256-bit store each; 1: random 4-byte words; 2: consecutive 8-byte
// words (coalesced); 3: random 128-byte chunks, 4 lanes x 32 bytes; 4: random 16-byte stores;
// 5: consecutive 32-byte rows (coalesced 256-bit stores)
struct Dsts { unsigned long long *p[16]; int n; };

__global__ void k_probe(Dsts d, uint64_t rows, uint64_t count, int mode)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t h = mix(i);
        unsigned long long *dst = d.p[(h >> 40) % (uint64_t)d.n];      // every store picks one of the destinations
        if (mode == 0) {
            unsigned long long *p = dst + (h % rows) * 4;
            asm volatile("st.relaxed.sys.global.v4.u64 [%0], {%1, %2, %3, %4};" ::"l"(p), "l"(i), "l"(i), "l"(i), "l"(i) : "memory");
        } else if (mode == 1) {
            reinterpret_cast<uint32_t *>(dst)[h % (rows * 8)] = (uint32_t)i;
        } else if (mode == 2) {
            dst[i % (rows * 4)] = i;
        } else if (mode == 3) {
            const uint64_t chunk = mix(i >> 2) % (rows / 4);
            unsigned long long *p = dst + chunk * 16 + (i & 3) * 4;
            asm volatile("st.relaxed.sys.global.v4.u64 [%0], {%1, %2, %3, %4};" ::"l"(p), "l"(i), "l"(i), "l"(i), "l"(i) : "memory");
        } else if (mode == 4) {
            unsigned long long *p = dst + (h % (rows * 2)) * 2;
            asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(i), "l"(i) : "memory");
        } else {
            unsigned long long *p = dst + (i % rows) * 4;
            asm volatile("st.relaxed.sys.global.v4.u64 [%0], {%1, %2, %3, %4};" ::"l"(p), "l"(i), "l"(i), "l"(i), "l"(i) : "memory");
        }
    }
}

extern "C" float probe(void **dsts, int n_dst, uint64_t rows, uint64_t count, int mode, int blocks)
{
    Dsts dst;
    dst.n = n_dst;
    for (int i = 0; i < 16; ++i) dst.p[i] = (unsigned long long *)dsts[i < n_dst ? i : 0];
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k_probe<<<blocks, 256>>>(dst, rows, count / 8, mode);      // warm-up
    cudaEventRecord(e0);
    k_probe<<<blocks, 256>>>(dst, rows, count, mode);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return cudaGetLastError() == cudaSuccess ? ms : -1.0f;
}
