// wprobe.cu — write-only HBM bandwidth probe for B200 (measurement aid, not product code).
//
// The rasteriser only WRITES (49 152 B per frame, nothing read back), while MEASURED_PEAKS.json's hbm_gbs is a copy
// (read + write).  This probe measures what a pure write stream can reach on this GPU with the store mechanisms the
// rasteriser could use, over the same footprint as bench.py (a 3 x 201 MB ring, each pass writes one 201 MB slot, so
// every pass evicts the previous one from the 126 MB L2):
//   stg     : st.global.cs.v4 from registers, grid-stride, G CTAs x 256 threads
//   bulk    : cp.async.bulk shared->global (TMA engine) of CHUNK bytes from a constant shared buffer, one CTA per SM,
//             D stores in flight per CTA before waiting for the engine to have read the buffer
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/wprobe.bin tools/wprobe.cu
// Run  : tools/wprobe.bin > gpurun_out/wprobe.json
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void __launch_bounds__(256) stg_kernel(uint4 *dst, size_t vecs, uint32_t seed)
{
    const uint4 v = make_uint4(seed, seed + 1, seed + 2, seed + 3);
    for (size_t i = blockIdx.x * 256ull + threadIdx.x; i < vecs; i += gridDim.x * 256ull) __stcs(dst + i, v);
}

// each CTA writes contiguous CHUNK-byte pieces: piece p goes to CTA p % gridDim.x
template <int kChunk>
__global__ void __launch_bounds__(128) bulk_kernel(uint8_t *dst, size_t bytes, int depth, uint32_t seed)
{
    extern __shared__ __align__(128) uint8_t smem[];
    for (int i = threadIdx.x; i < kChunk / 16; i += blockDim.x) reinterpret_cast<uint4 *>(smem)[i] = make_uint4(seed, i, seed, i);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x != 0) return;
    const size_t pieces = bytes / kChunk;
    const uint32_t src = static_cast<uint32_t>(__cvta_generic_to_shared(smem));
    int inflight = 0;
    for (size_t p = blockIdx.x; p < pieces; p += gridDim.x) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(dst + p * kChunk), "r"(src), "r"(kChunk) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        if (++inflight >= depth) { // keep at most `depth` groups pending (reads of the shared buffer)
            switch (depth) {
            case 1: asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); break;
            case 2: asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); break;
            case 4: asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory"); break;
            default: asm volatile("cp.async.bulk.wait_group.read 7;" ::: "memory"); break;
            }
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <typename F>
static void timeit(const char *name, const char *detail, size_t bytes_per_pass, F launch, bool last)
{
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    for (int i = 0; i < 6; ++i) launch(i);
    CK(cudaDeviceSynchronize());
    // (1) back-to-back passes over the ring, one event pair around 60 passes: steady state like bench.py
    const int passes = 60;
    CK(cudaEventRecord(a));
    for (int i = 0; i < passes; ++i) launch(i);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, a, b));
    const double steady = bytes_per_pass * double(passes) / (ms * 1e-3) / 1e9;
    // (2) per-launch timing, best and median
    std::vector<float> t;
    for (int i = 0; i < 30; ++i) {
        CK(cudaEventRecord(a));
        launch(i);
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        CK(cudaEventElapsedTime(&ms, a, b));
        t.push_back(ms);
    }
    std::sort(t.begin(), t.end());
    printf("  {\"variant\": \"%s\", \"detail\": \"%s\", \"steady_GBs\": %.1f, \"best_GBs\": %.1f, \"median_GBs\": %.1f}%s\n", name, detail, steady,
           bytes_per_pass / (t.front() * 1e-3) / 1e9, bytes_per_pass / (t[t.size() / 2] * 1e-3) / 1e9, last ? "" : ",");
    CK(cudaEventDestroy(a));
    CK(cudaEventDestroy(b));
}

int main()
{
    const size_t slot = 4096ull * 49152; // 201 MB: one bench.py step
    uint8_t *ring = nullptr;
    CK(cudaMalloc(&ring, 3 * slot));
    cudaDeviceProp prop{};
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"slot_bytes\": %zu, \"results\": [\n", prop.name, sms, slot);
    char detail[128];
    for (int per_sm : {2, 4, 8, 16}) {
        snprintf(detail, sizeof detail, "%d CTAs/SM x 256 thr, st.global.cs.v4", per_sm);
        timeit("stg", detail, slot, [&](int i) { stg_kernel<<<sms * per_sm, 256>>>(reinterpret_cast<uint4 *>(ring + (i % 3) * slot), slot / 16, i); }, false);
    }
    CK(cudaFuncSetAttribute(bulk_kernel<16384>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384));
    CK(cudaFuncSetAttribute(bulk_kernel<49152>, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152));
    for (int ctas : {1, 2, 3, 4}) {
        for (int depth : {1, 2, 4, 8}) {
            snprintf(detail, sizeof detail, "%d CTAs/SM, 16 KB per cp.async.bulk, %d in flight per CTA", ctas, depth);
            timeit("bulk16k", detail, slot, [&](int i) { bulk_kernel<16384><<<sms * ctas, 128, 16384>>>(ring + (i % 3) * slot, slot, depth, i); }, false);
        }
    }
    for (int ctas : {1, 2, 3, 4}) {
        for (int depth : {1, 2, 4}) {
            snprintf(detail, sizeof detail, "%d CTAs/SM, 48 KB per cp.async.bulk, %d in flight per CTA", ctas, depth);
            timeit("bulk48k", detail, slot, [&](int i) { bulk_kernel<49152><<<sms * ctas, 128, 49152>>>(ring + (i % 3) * slot, slot, depth, i); }, ctas == 4 && depth == 4);
        }
    }
    printf("]}\n");
    CK(cudaFree(ring));
    return 0;
}
