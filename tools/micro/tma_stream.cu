// tma_stream.cu -- how fast can cp.async.bulk (1-D, global -> shared, mbarrier complete_tx) stream HBM on a B200, as a function of
// stage size, ring depth, copies per stage and blocks per SM?  The consumer does nothing: wait for the stage, refill it.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_stream tma_stream.cu && ./tma_stream
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint32_t b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.relaxed.cta.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t b, uint32_t ph) {
    uint32_t ok = 0;
    while (!ok) asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(b), "r"(ph) : "memory");
}
__device__ __forceinline__ void bulk(uint32_t dst, const void* src, uint32_t n, uint32_t b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(n), "r"(b) : "memory");
}
// each block streams a contiguous range of `chunks` stages; a stage is `ncopy` copies of `cbytes` from `ncopy` separate arrays (planes)
__global__ void k_stream(const unsigned char* src, size_t plane_stride, uint32_t chunks_total, uint32_t nstage, uint32_t ncopy, uint32_t cbytes, unsigned long long* sink) {
    extern __shared__ __align__(128) unsigned char sm[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm);
    unsigned char* ring = sm + 128;
    const uint32_t sbytes = ncopy * cbytes;
    const uint32_t c0 = (uint32_t)((uint64_t)chunks_total * blockIdx.x / gridDim.x), c1 = (uint32_t)((uint64_t)chunks_total * (blockIdx.x + 1) / gridDim.x);
    if (threadIdx.x == 0) { for (uint32_t s = 0; s < nstage; ++s) mbar_init(s32(bars + s), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    auto issue = [&](uint32_t c, uint32_t s) {
        mbar_expect(s32(bars + s), sbytes);
        for (uint32_t q = 0; q < ncopy; ++q) bulk(s32(ring + (size_t)s * sbytes + q * cbytes), src + q * plane_stride + (size_t)c * cbytes, cbytes, s32(bars + s));
    };
    unsigned long long acc = 0;
    if (threadIdx.x == 0) for (uint32_t s = 0; s < nstage && c0 + s < c1; ++s) issue(c0 + s, s);
    for (uint32_t c = c0; c < c1; ++c) {
        const uint32_t k = c - c0, s = k % nstage;
        mbar_wait(s32(bars + s), (k / nstage) & 1u);
        acc += *reinterpret_cast<volatile unsigned long long*>(ring + (size_t)s * sbytes + 8 * threadIdx.x);
        __syncthreads();
        if (threadIdx.x == 0 && c + nstage < c1) issue(c + nstage, s);
    }
    if (acc == 0x1234567) sink[0] = acc;
}
// reference: plain 16-byte loads, grid-stride, 8 loads in flight per thread
__global__ void k_ldg(const uint4* src, size_t n, unsigned long long* sink) {
    unsigned long long acc = 0;
    const size_t gs = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 7 * gs < n; i += 8 * gs) {
        uint4 v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = src[i + q * gs];
#pragma unroll
        for (int q = 0; q < 8; ++q) acc += v[q].x ^ v[q].w;
    }
    if (acc == 0x1234567) sink[0] = acc;
}
int main() {
    const size_t total = (size_t)3 << 30;      // 3 GiB: three "planes" of 1 GiB
    unsigned char* buf; unsigned long long* sink;
    cudaMalloc(&buf, total); cudaMalloc(&sink, 8); cudaMemset(buf, 1, total);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaFuncSetAttribute(k_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    {   float best = 1e9;
        for (int rep = 0; rep < 3; ++rep) { cudaEventRecord(e0); k_ldg<<<sms * 8, 512>>>((const uint4*)buf, total / 16, sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best; }
        std::printf("LDG.128 x8 in flight, %d x 8 blocks x 512: %.0f GB/s\n", sms, total / best / 1e6); }
    struct Cfg { uint32_t bps, nstage, ncopy, cbytes, threads; };
    const Cfg cfgs[] = {
        {2, 3, 3, 8192, 512}, {2, 3, 1, 24576, 512}, {2, 3, 7, 4096, 512}, {2, 4, 3, 8192, 512}, {2, 3, 3, 8192, 64}, {2, 4, 1, 24576, 64},
        {1, 6, 3, 8192, 64}, {1, 8, 1, 24576, 64}, {1, 4, 1, 49152, 64}, {4, 3, 1, 16384, 64}, {4, 3, 3, 4096, 64}, {3, 3, 3, 8192, 64}, {2, 2, 3, 8192, 64},
        {8, 3, 1, 8192, 64}, {2, 3, 1, 32768, 64}, {1, 3, 1, 65536, 64},
    };
    for (const Cfg& c : cfgs) {
        const uint32_t sbytes = c.ncopy * c.cbytes; const size_t smem = 128 + (size_t)c.nstage * sbytes;
        const size_t plane = (total / c.ncopy) / c.cbytes * c.cbytes; const uint32_t chunks = (uint32_t)(plane / c.cbytes);      // ncopy planes inside the buffer
        float best = 1e9;
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            k_stream<<<sms * c.bps, c.threads, smem>>>(buf, plane, chunks, c.nstage, c.ncopy, c.cbytes, sink);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
        }
        cudaError_t err = cudaGetLastError();
        std::printf("bulk: %u blocks/SM x %3u thr, %u stages x %u copies x %5u B (%3zu KB smem/block, %3zu KB in flight/SM): %6.0f GB/s %s\n", c.bps, c.threads, c.nstage, c.ncopy, c.cbytes,
                    smem >> 10, (size_t)c.bps * c.nstage * sbytes >> 10, (double)c.ncopy * plane / best / 1e6, err == cudaSuccess ? "" : cudaGetErrorString(err));
    }
    return 0;
}
