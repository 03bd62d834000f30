// Micro-benchmark: how fast can one B200 gather "32 blocks x W bytes" tiles (block stride 32 KiB), the
// access pattern of the lane-per-block coders?  Variants: TMA 2D boxes, LDGSTS (cp.async 16 B), both
// double-buffered per warp; consumption = each lane XORs its row from shared memory.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gather_bw gather_bw.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load(uint32_t dst, const CUtensorMap* map, int32_t x, int32_t y, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(x), "r"(y), "r"(bar) : "memory");
}
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ uint4 lds128(uint32_t a) { uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v; }
__device__ __forceinline__ void cp16(uint32_t s, const void* g) { asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(g) : "memory"); }

constexpr uint32_t B = 8192; // symbols per block

// ROWB = bytes per row and tile; STAGES buffers
template <int ROWB, int STAGES>
__global__ void __launch_bounds__(1024, 1) k_tma(const __grid_constant__ CUtensorMap tmap, uint32_t n_tasks, uint32_t* out)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t warps = blockDim.x >> 5, lane = threadIdx.x & 31;
    const uint32_t warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    constexpr uint32_t TILE = 32 * ROWB;
    const uint32_t tiles0 = smem_u32(smem) + warp * STAGES * TILE;
    const uint32_t bars = smem_u32(smem) + warps * STAGES * TILE + warp * 8 * STAGES;
    if (lane == 0) {
        for (int s = 0; s < STAGES; s++) mbar_init(bars + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t acc = 0, ph = 0;
    const uint32_t tiles = B * 4 / ROWB;
    for (uint32_t task = blockIdx.x * warps + warp; task < n_tasks; task += gridDim.x * warps) {
        auto issue = [&](uint32_t t) {
            if (elect_one()) {
                const uint32_t s = t % STAGES;
                mbar_expect_tx(bars + 8 * s, TILE);
                tma_load(tiles0 + s * TILE, &tmap, int32_t(B - (ROWB / 4) * (t + 1)), int32_t(task * 32), bars + 8 * s);
            }
        };
        for (uint32_t t = 0; t < STAGES - 1; t++) issue(t);
        for (uint32_t t = 0; t < tiles; t++) {
            __syncwarp();
            if (t + STAGES - 1 < tiles) issue(t + STAGES - 1);
            const uint32_t s = t % STAGES;
            mbar_wait(bars + 8 * s, (ph >> s) & 1);
            ph ^= 1u << s;
            const uint32_t row = tiles0 + s * TILE + lane * ROWB;
#pragma unroll
            for (int j = 0; j < ROWB / 16; j++) { const uint4 v = lds128(row + j * 16); acc ^= v.x ^ v.y ^ v.z ^ v.w; }
        }
    }
    if (acc == 0x12345678) out[0] = acc;
}

template <int ROWB>
__global__ void __launch_bounds__(1024, 1) k_ldgsts(const uint32_t* in, uint32_t n_tasks, uint32_t* out)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t warps = blockDim.x >> 5, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr uint32_t ROWS = ROWB + 16; // padded row
    constexpr uint32_t TILE = 32 * ROWS;
    const uint32_t tiles0 = smem_u32(smem) + warp * 2 * TILE;
    uint32_t acc = 0;
    const uint32_t tiles = B * 4 / ROWB;
    constexpr int LPR = ROWB / 16;     // lanes per row
    constexpr int RPI = 32 / LPR;      // rows per instruction
    for (uint32_t task = blockIdx.x * warps + warp; task < n_tasks; task += gridDim.x * warps) {
        const uint8_t* base = reinterpret_cast<const uint8_t*>(in) + uint64_t(task) * 32 * B * 4;
        auto issue = [&](uint32_t t, uint32_t b) {
#pragma unroll
            for (int it = 0; it < 32 / RPI; it++) {
                const uint32_t row = it * RPI + lane / LPR, ch = lane % LPR;
                cp16(tiles0 + b * TILE + row * ROWS + ch * 16, base + uint64_t(row) * B * 4 + (B * 4 - ROWB * (t + 1)) + ch * 16);
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        issue(0, 0);
        for (uint32_t t = 0; t < tiles; t++) {
            __syncwarp();
            if (t + 1 < tiles) issue(t + 1, (t + 1) & 1); else asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group 1;" ::: "memory");
            __syncwarp();
            const uint32_t row = tiles0 + (t & 1) * TILE + lane * ROWS;
#pragma unroll
            for (int j = 0; j < ROWB / 16; j++) { const uint4 v = lds128(row + j * 16); acc ^= v.x ^ v.y ^ v.z ^ v.w; }
        }
    }
    if (acc == 0x12345678) out[0] = acc;
}

typedef CUresult (*enc_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <class F> float time_it(F f)
{
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f(); CK(cudaDeviceSynchronize());
    float best = 1e9;
    for (int i = 0; i < 3; i++) { cudaEventRecord(a); f(); cudaEventRecord(b); CK(cudaEventSynchronize(b)); float ms; cudaEventElapsedTime(&ms, a, b); best = ms < best ? ms : best; }
    return best;
}

int main()
{
    const uint64_t n = 1ull << 30;
    uint32_t *d_in, *d_out;
    CK(cudaMalloc(&d_in, n * 4)); CK(cudaMalloc(&d_out, 64));
    CK(cudaMemset(d_in, 1, n * 4));
    const uint32_t n_blocks = uint32_t(n / B), n_tasks = n_blocks / 32;
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    enc_fn encode = (enc_fn)p;
    auto make_map = [&](uint32_t rowb, CUtensorMapSwizzle sw) {
        CUtensorMap m; const cuuint64_t dims[2] = { B, n_blocks }; const cuuint64_t strides[1] = { B * 4ull };
        const cuuint32_t box[2] = { rowb / 4, 32 }; const cuuint32_t es[2] = { 1, 1 };
        CUresult rc = encode(&m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, d_in, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (rc != CUDA_SUCCESS) { printf("tensor map failed %d\n", int(rc)); exit(1); }
        return m;
    };
    const double gb = n * 4 / 1e9;
    for (int warps : { 7, 14, 28 }) {
#define RUN_TMA(ROWB, ST, SW) { auto m = make_map(ROWB, SW); size_t sm = size_t(warps) * ST * 32 * ROWB + warps * 8 * ST + 64; \
        if (sm <= 227 * 1024) { CK(cudaFuncSetAttribute(k_tma<ROWB, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sm))); \
        float ms = time_it([&] { k_tma<ROWB, ST><<<148, warps * 32, sm>>>(m, n_tasks, d_out); }); \
        printf("tma    rowb=%3d stages=%d warps=%2d : %.3f ms  %.0f GB/s\n", ROWB, ST, warps, ms, gb / ms * 1e3); } }
        RUN_TMA(32, 2, CU_TENSOR_MAP_SWIZZLE_32B) RUN_TMA(64, 2, CU_TENSOR_MAP_SWIZZLE_64B) RUN_TMA(128, 2, CU_TENSOR_MAP_SWIZZLE_128B)
        RUN_TMA(256, 2, CU_TENSOR_MAP_SWIZZLE_NONE) RUN_TMA(64, 3, CU_TENSOR_MAP_SWIZZLE_64B) RUN_TMA(64, 4, CU_TENSOR_MAP_SWIZZLE_64B) RUN_TMA(128, 4, CU_TENSOR_MAP_SWIZZLE_128B)
#define RUN_LDG(ROWB) { size_t sm = size_t(warps) * 2 * 32 * (ROWB + 16) + 64; if (sm <= 227 * 1024) { \
        CK(cudaFuncSetAttribute(k_ldgsts<ROWB>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sm))); \
        float ms = time_it([&] { k_ldgsts<ROWB><<<148, warps * 32, sm>>>(d_in, n_tasks, d_out); }); \
        printf("ldgsts rowb=%3d          warps=%2d : %.3f ms  %.0f GB/s\n", ROWB, warps, ms, gb / ms * 1e3); } }
        RUN_LDG(32) RUN_LDG(64) RUN_LDG(128) RUN_LDG(256)
    }
    return 0;
}
