// gather_peak.cu - what a B200 delivers for RANDOM 16-byte gathers (one 32-byte sector each) from a buffer far larger than L2:
// the roofline of the steady-state epipolar search, which reads one float4 footprint per step from an image no other
// lane of the warp touches.  Measurement tool, not part of libdvo_cuda.so.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/gather_peak tools/gather_peak.cu && tools/gather_peak [GiB]
// Prints, per (threads per CTA, CTAs per SM, independent loads in flight per thread, ALU work per load), G sectors/s and GB/s (32 B each).
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

// U independent chains per thread; each chain's next address depends on the value loaded (like the search's dependent compare),
// WORK fused multiply-adds between a load and its use stand for the bilinear + SSD arithmetic of one search step.
template <int U, int WORK>
__global__ void k_gather(const float4* __restrict__ buf, uint32_t mask, int iters, float* __restrict__ out)
{
    uint32_t idx[U];
    float acc[U];
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
    for (int u = 0; u < U; u++) { idx[u] = mix(gid * U + u + 1); acc[u] = 0.f; }
    for (int it = 0; it < iters; it++) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = __ldg(buf + (idx[u] & mask));
#pragma unroll
        for (int u = 0; u < U; u++) {
            float t = v[u].x + v[u].y * 0.5f + v[u].z * 0.25f + v[u].w * 0.125f;
#pragma unroll
            for (int w = 0; w < WORK; w++) t = fmaf(t, 1.0001f, 0.5f);
            acc[u] += t;
            idx[u] = mix(idx[u] + (uint32_t)it + (__float_as_uint(t) & 1u)); // the next address needs the loaded value
        }
    }
    float s = 0.f;
#pragma unroll
    for (int u = 0; u < U; u++) s += acc[u];
    if (s == 12345.678f) out[gid] = s;
}

// The search's own access pattern without its arithmetic: every thread walks LINES of `steps` one-pixel steps in a random
// direction through a random 160x120 float4 image (307 KB, like a keyframe's footprint image) out of n_img images, one 16-byte
// load per step whose value the next step waits for (the search compares before it moves on).
__global__ void k_walk(const float4* __restrict__ buf, uint32_t n_img, int walks, int steps, float* __restrict__ out)
{
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t r = mix(gid + 1);
    float acc = 0.f;
    for (int wk = 0; wk < walks; wk++) {
        r = mix(r + wk);
        const float4* img = buf + (size_t)(r % n_img) * (160 * 120);
        const uint32_t r2 = mix(r ^ 0x9e3779b9u);
        float x = 20.f + (float)(r2 & 0x7f) * (120.f / 128.f), y = 20.f + (float)((r2 >> 8) & 0x7f) * (80.f / 128.f);
        float sn, cs;
        __sincosf((float)(r2 >> 16) * (6.2831853f / 65536.f), &sn, &cs);
        for (int k = 0; k < steps; k++) {
            const int xi = (int)x, yi = (int)y;
            if (xi >= 0 && xi < 160 && yi >= 0 && yi < 120) {
                const float4 v = __ldg(img + yi * 160 + xi);
                acc += v.x + v.y + v.z + v.w;
            }
            x += cs + (acc == 1.5f ? 1.f : 0.f); // dependent on the load
            y += sn;
        }
    }
    if (acc == 12345.678f) out[gid] = acc;
}
static void run_walk(const float4* buf, uint32_t n_img, float* out, int threads, int ctas_per_sm, int n_sm, int steps)
{
    const int walks = 2048 / steps, grid = n_sm * ctas_per_sm;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k_walk<<<grid, threads>>>(buf, n_img, 2, steps, out);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    k_walk<<<grid, threads>>>(buf, n_img, walks, steps, out);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double taps = (double)grid * threads * walks * steps;
    printf("line walks of %3d steps through random 307 KB images (%u images, %.1f GiB)  threads %3d  CTAs/SM %2d : %7.2f G taps/s  (%.2f ms)\n", steps, n_img,
           (double)n_img * 160 * 120 * 16 / (1 << 30), threads, ctas_per_sm, taps / ms * 1e-6, ms);
}

// Both 32-byte sectors of a random 64-byte-aligned pair: if the sector rate doubles against k_gather, DRAM is limited by
// ACCESSES per second (64-byte atoms), not by sectors or bytes.
__global__ void k_gather_pair(const float4* __restrict__ buf, uint32_t mask, int iters, float* __restrict__ out)
{
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t idx = mix(gid + 1);
    float acc = 0.f;
    for (int it = 0; it < iters; it++) {
        const uint32_t i0 = (idx & mask) & ~3u;
        const float4 v0 = __ldg(buf + i0), v1 = __ldg(buf + i0 + 2);
        const float t = v0.x + v0.y + v1.z + v1.w;
        acc += t;
        idx = mix(idx + (uint32_t)it + (__float_as_uint(t) & 1u));
    }
    if (acc == 12345.678f) out[gid] = acc;
}
static void run_pair(const float4* buf, uint32_t mask, float* out, int threads, int ctas_per_sm, int n_sm)
{
    const int iters = 2048, grid = n_sm * ctas_per_sm;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k_gather_pair<<<grid, threads>>>(buf, mask, 64, out);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    k_gather_pair<<<grid, threads>>>(buf, mask, iters, out);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double pairs = (double)grid * threads * iters;
    printf("random 64-byte pairs (2 sectors each)  threads %4d  CTAs/SM %2d : %7.2f G pairs/s  %7.2f G sectors/s  %7.1f GB/s  (%.2f ms)\n", threads, ctas_per_sm,
           pairs / ms * 1e-6, 2 * pairs / ms * 1e-6, pairs * 64 / ms * 1e-6, ms);
}

template <int U, int WORK>
static void run(const float4* buf, uint32_t mask, float* out, int threads, int ctas_per_sm, int n_sm)
{
    const int iters = 2048 / U;
    const int grid = n_sm * ctas_per_sm;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k_gather<U, WORK><<<grid, threads>>>(buf, mask, iters / 8, out);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    k_gather<U, WORK><<<grid, threads>>>(buf, mask, iters, out);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double loads = (double)grid * threads * U * iters;
    printf("threads %4d  CTAs/SM %2d  warps/SM %2d  in flight/thread %d  fma/load %3d : %7.2f G sectors/s  %7.1f GB/s  (%.2f ms)\n", threads, ctas_per_sm,
           threads * ctas_per_sm / 32, U, WORK, loads / ms * 1e-6, loads * 32 / ms * 1e-6, ms);
}

int main(int argc, char** argv)
{
    const double gib = argc > 1 ? atof(argv[1]) : 8.0;
    uint32_t n = 1;
    while ((double)n * 2 * 16 <= gib * 1024 * 1024 * 1024) n *= 2; // power of two float4 elements
    if (argc > 2) { // L2 fetch granularity hint: 32, 64 or 128 bytes
        CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(argv[2])));
        size_t gr = 0; CK(cudaDeviceGetLimit(&gr, cudaLimitMaxL2FetchGranularity));
        printf("cudaLimitMaxL2FetchGranularity = %zu\n", gr);
    }
    float4* buf; float* out;
    CK(cudaMalloc(&buf, (size_t)n * 16));
    CK(cudaMemset(buf, 0, (size_t)n * 16));
    CK(cudaMalloc(&out, 64 << 20));
    int n_sm = 0;
    CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, 0));
    printf("random 16-byte gathers from %.1f GiB (%u float4), %d SMs; one 32-byte sector per load\n", (double)n * 16 / (1 << 30), n, n_sm);
    const uint32_t mask = n - 1;
    run<1, 0>(buf, mask, out, 64, 16, n_sm);
    run<1, 0>(buf, mask, out, 64, 32, n_sm);
    run<1, 0>(buf, mask, out, 256, 8, n_sm);
    run<2, 0>(buf, mask, out, 256, 8, n_sm);
    run<4, 0>(buf, mask, out, 256, 8, n_sm);
    run<8, 0>(buf, mask, out, 256, 8, n_sm);
    run<1, 40>(buf, mask, out, 64, 16, n_sm);   // k_map_update's shape: 64-thread CTAs, 16 per SM, ~40 issue slots of arithmetic per step
    run<1, 40>(buf, mask, out, 64, 32, n_sm);
    run<2, 40>(buf, mask, out, 64, 16, n_sm);
    run<4, 40>(buf, mask, out, 64, 16, n_sm);
    run<1, 100>(buf, mask, out, 64, 16, n_sm);
    run_pair(buf, mask, out, 64, 16, n_sm);
    run_pair(buf, mask, out, 256, 8, n_sm);
    {
        const uint32_t n_img = (uint32_t)(((size_t)n * 16) / (160 * 120 * 16));
        for (int steps : {8, 33, 102}) {
            run_walk(buf, n_img, out, 64, 16, n_sm, steps);
            run_walk(buf, n_img, out, 64, 32, n_sm, steps);
        }
        run_walk(buf, 2048, out, 64, 16, n_sm, 33); // 0.6 GiB of images: what an early run (few keyframes per stream) sees
    }
    // footprint sweep at the best configuration: where L2 stops helping
    cudaFree(buf);
    for (double g : {0.0625, 0.25, 1.0, 4.0}) {
        uint32_t m = 1; while ((double)m * 2 * 16 <= g * 1024 * 1024 * 1024) m *= 2;
        CK(cudaMalloc(&buf, (size_t)m * 16)); CK(cudaMemset(buf, 0, (size_t)m * 16));
        printf("footprint %.3f GiB: ", (double)m * 16 / (1 << 30));
        run<4, 0>(buf, m - 1, out, 256, 8, n_sm);
        cudaFree(buf);
    }
    return 0;
}
