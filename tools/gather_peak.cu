// gather_peak.cu -- ceiling for the access pattern of the refinement kernel: how fast can this GPU READ random,
// aligned rows of `row_bytes` bytes from a buffer much larger than L2, with no arithmetic to speak of?
// Variant A: plain LDG.128 (every lane 16 B of the row, `unroll` rows in flight per warp).
// Variant B: cp.async (LDGSTS) into shared memory, 2-deep per warp -- the mechanism the product uses.
// Variant C: TMA bulk copy (cp.async.bulk, SASS UBLKCP) of a whole row into shared memory, completion on one mbarrier per
//            ring slot, 3-deep per warp, issued by lane 0 -- what a TMA version of the product's gather would do per row.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/gather_peak tools/gather_peak.cu
// Run:   tools/gather_peak [row_bytes=4096] [buffer_GB=32] [length multiplier=1]      (prints JSON lines)
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("{\"error\": \"%s\"}\n", cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}

// each warp reads `rows_per_warp` random rows; a lane reads row_bytes / 512 float4 per row
template <int VEC>  // float4 per lane per row
__global__ void __launch_bounds__(256, 2) gather_ldg(const float4* __restrict__ buf, uint32_t n_rows, int rows_per_warp,
                                                     float* __restrict__ sink) {
    const int lane = threadIdx.x & 31;
    const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    float acc = 0.f;
    for (int r = 0; r < rows_per_warp; r += 4) {
        float4 v[4][VEC];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t row = hash32(gw * 9781u + (uint32_t)(r + k) * 6271u + 17u) % n_rows;
            const float4* p = buf + (size_t)row * (32 * VEC) + lane;
#pragma unroll
            for (int j = 0; j < VEC; ++j) v[k][j] = __ldcg(p + 32 * j);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int j = 0; j < VEC; ++j) acc += v[k][j].x + v[k][j].y + v[k][j].z + v[k][j].w;
    }
    if (acc == 1234.5f) sink[0] = acc;
}

template <int VEC>
__global__ void __launch_bounds__(256, 2) gather_cpasync(const float4* __restrict__ buf, uint32_t n_rows, int rows_per_warp,
                                                         float* __restrict__ sink) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    constexpr int ROW = 512 * VEC, DEPTH = 3;
    unsigned char* mine = smem + warp * DEPTH * ROW;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(mine) + lane * 16;
    float acc = 0.f;
    auto issue = [&](int r) {
        const uint32_t row = hash32(gw * 9781u + (uint32_t)r * 6271u + 17u) % n_rows;
        const float4* p = buf + (size_t)row * (32 * VEC) + lane;
        const uint32_t d = base + (r % DEPTH) * ROW;
#pragma unroll
        for (int j = 0; j < VEC; ++j)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d + j * 512), "l"(p + 32 * j) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int r = 0; r < DEPTH - 1 && r < rows_per_warp; ++r) issue(r);
    for (int r = 0; r < rows_per_warp; ++r) {
        if (r + DEPTH - 1 < rows_per_warp) issue(r + DEPTH - 1); else asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group %0;" ::"n"(DEPTH - 1) : "memory");
        __syncwarp();
        const float4* s = reinterpret_cast<const float4*>(mine + (r % DEPTH) * ROW) + lane;
#pragma unroll
        for (int j = 0; j < VEC; ++j) { const float4 v = s[32 * j]; acc += v.x + v.y + v.z + v.w; }
        __syncwarp();
    }
    if (acc == 1234.5f) sink[0] = acc;
}

template <int VEC>
__global__ void __launch_bounds__(256, 2) gather_bulk(const float4* __restrict__ buf, uint32_t n_rows, int rows_per_warp,
                                                      float* __restrict__ sink) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) unsigned long long bars[8][4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    constexpr int ROW = 512 * VEC, DEPTH = 3;
    unsigned char* mine = smem + warp * DEPTH * ROW;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(mine);
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&bars[warp][0]);
    if (lane == 0) {
        for (int d = 0; d < DEPTH; ++d) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8 * d) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    float acc = 0.f;
    auto issue = [&](int r) {
        if (lane == 0) {
            const uint32_t row = hash32(gw * 9781u + (uint32_t)r * 6271u + 17u) % n_rows;
            const float4* p = buf + (size_t)row * (32 * VEC);
            const uint32_t d = base + (r % DEPTH) * ROW, bar = bar0 + 8 * (r % DEPTH);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the slot's last reads (generic proxy) before the bulk write
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(ROW) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(d), "l"(p),
                         "r"(ROW), "r"(bar)
                         : "memory");
        }
    };
    for (int r = 0; r < DEPTH - 1 && r < rows_per_warp; ++r) issue(r);
    for (int r = 0; r < rows_per_warp; ++r) {
        if (r + DEPTH - 1 < rows_per_warp) issue(r + DEPTH - 1);
        const uint32_t bar = bar0 + 8 * (r % DEPTH), parity = (uint32_t)(r / DEPTH) & 1u;
        uint32_t done = 0;
        while (!done) {
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                         : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        }
        const float4* s = reinterpret_cast<const float4*>(mine + (r % DEPTH) * ROW) + lane;
#pragma unroll
        for (int j = 0; j < VEC; ++j) { const float4 v = s[32 * j]; acc += v.x + v.y + v.z + v.w; }
        __syncwarp();
    }
    if (acc == 1234.5f) sink[0] = acc;
}

template <int VEC>
int run(const float4* buf, uint32_t n_rows, float* sink, int sms, size_t row_bytes, int mult) {
    const int grid = sms * 2, warps = grid * 8, rows_per_warp = mult * 4096 * 4 / VEC;
    const double bytes = (double)warps * rows_per_warp * row_bytes;
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    for (int variant = 0; variant < 3; ++variant) {
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
            CK(cudaEventRecord(a));
            if (variant == 0) gather_ldg<VEC><<<grid, 256>>>(buf, n_rows, rows_per_warp, sink);
            else if (variant == 1) {
                CK(cudaFuncSetAttribute(gather_cpasync<VEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 3 * 512 * VEC));
                gather_cpasync<VEC><<<grid, 256, 8 * 3 * 512 * VEC>>>(buf, n_rows, rows_per_warp, sink);
            } else {
                CK(cudaFuncSetAttribute(gather_bulk<VEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 3 * 512 * VEC));
                gather_bulk<VEC><<<grid, 256, 8 * 3 * 512 * VEC>>>(buf, n_rows, rows_per_warp, sink);
            }
            CK(cudaEventRecord(b));
            CK(cudaEventSynchronize(b));
            float ms; CK(cudaEventElapsedTime(&ms, a, b));
            if (rep > 0 && ms < best) best = ms;
        }
        printf("{\"variant\": \"%s\", \"row_bytes\": %zu, \"gb_read\": %.1f, \"ms\": %.3f, \"gbs\": %.1f}\n",
               variant == 0 ? "ldg128" : (variant == 1 ? "cp.async" : "cp.async.bulk (TMA, mbarrier per slot)"), row_bytes, bytes / 1e9, best, bytes / best / 1e6);
    }
    return 0;
}

int main(int argc, char** argv) {
    const size_t row_bytes = argc > 1 ? (size_t)atol(argv[1]) : 4096;
    const double gb = argc > 2 ? atof(argv[2]) : 32.0;
    const int mult = argc > 3 ? atoi(argv[3]) : 1;  // 1: ~3 ms per launch, 16: ~45 ms (sustained clocks)
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const size_t total = (size_t)(gb * 1e9) / row_bytes * row_bytes;
    float4* buf; float* sink;
    CK(cudaMalloc(&buf, total)); CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(buf, 0, total));
    const uint32_t n_rows = (uint32_t)(total / row_bytes);
    switch (row_bytes) {
        case 512: return run<1>(buf, n_rows, sink, sms, row_bytes, mult);
        case 1024: return run<2>(buf, n_rows, sink, sms, row_bytes, mult);
        case 2048: return run<4>(buf, n_rows, sink, sms, row_bytes, mult);
        case 4096: return run<8>(buf, n_rows, sink, sms, row_bytes, mult);
        default: printf("{\"error\": \"row_bytes must be 512, 1024, 2048 or 4096\"}\n"); return 1;
    }
}
