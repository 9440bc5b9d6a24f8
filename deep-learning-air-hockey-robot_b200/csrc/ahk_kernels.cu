// ahk_kernels.cu — the stand-alone kernels around the fused step kernel (ahk_step.cu).
//
//   env_kernel     K1+K4 alone: per environment (one thread each, SoA state, coalesced): auto-reset, programmed AI,
//                  physics tick, reward / terminal / labels, running statistics, and a 16-byte draw record per frame
//                  to paint (ahk_env.cuh).  Used by the split path (AHK_PATH=split), where a raster kernel follows.
//   observe_kernel materialises the reference-format observation (processor.py:52-64) from the ring on request.
#include "ahk_env.cuh"

namespace ahk {

template <bool kStep>
__global__ void __launch_bounds__(kEnvThreads) env_kernel(EnvArgs a)
{
    const int i = blockIdx.x * kEnvThreads + threadIdx.x;
    StatDelta d;
    if (i < a.n) {
        uint4 rec[3];
        env_update<kStep>(a, i, rec, d);
        uint4 *dl = reinterpret_cast<uint4 *>(a.drawlist);
        const size_t n = static_cast<size_t>(a.n);
        dl[i] = rec[0];
        dl[n + i] = rec[1];
        dl[2 * n + i] = rec[2];
    }
    accumulate_stats<kStep>(a.stats, d);
}

// ------------------------------------------------------------------------------------------------ observation
// One thread per 16 pixels of one output plane position. processor.py:52 ('rgb'), :53-64 ('gray-diff').
__global__ void __launch_bounds__(256) observe_kernel(ObserveArgs a)
{
    const size_t n = static_cast<size_t>(a.n);
    if (a.format == 2) { // gray-diff: float64 [n][1][128][128]
        const size_t total = n * 1024; // vectors per plane
        for (size_t idx = blockIdx.x * 256ull + threadIdx.x; idx < total; idx += gridDim.x * 256ull) {
            const size_t env = idx >> 10, v = idx & 1023;
            uint4 q[3][3]; // [age][channel]
#pragma unroll
            for (int age = 0; age < 3; ++age) {
                const uint4 *f = reinterpret_cast<const uint4 *>(stack_frame(a.frames, a.warm, a.meta, a.depth, n, a.head, env, age));
#pragma unroll
                for (int c = 0; c < 3; ++c) q[age][c] = f[c * 1024 + v];
            }
            double *out = reinterpret_cast<double *>(a.out) + (env * 1024 + v) * 16;
#pragma unroll
            for (int p = 0; p < 16; ++p) {
                double ch[3];
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    auto byte = [&](const uint4 &u) { return (reinterpret_cast<const uint32_t *>(&u)[p >> 2] >> (8 * (p & 3))) & 255u; };
                    const uint32_t d = (byte(q[2][c]) - byte(q[1][c]) - byte(q[0][c])) & 255u; // uint8 wrap-around, processor.py:56
                    ch[c] = static_cast<double>(d);
                }
                const double gray = __dadd_rn(__dadd_rn(__dmul_rn(0.2989, ch[0]), __dmul_rn(0.5870, ch[1])), __dmul_rn(0.1140, ch[2]));
                out[p] = __ddiv_rn(__dsub_rn(gray, 128.0), 128.0);
            }
        }
        return;
    }
    // stacked 9-channel forms, newest first: uint8 or (x-128)/128 float32
    const size_t total = n * 3 * kFrameVecs;
    for (size_t idx = blockIdx.x * 256ull + threadIdx.x; idx < total; idx += gridDim.x * 256ull) {
        const size_t env = idx / (3 * kFrameVecs);
        const size_t rem = idx - env * (3 * kFrameVecs);
        const int age = static_cast<int>(rem / kFrameVecs);
        const size_t v = rem - static_cast<size_t>(age) * kFrameVecs;
        const uint4 q = *reinterpret_cast<const uint4 *>(stack_frame(a.frames, a.warm, a.meta, a.depth, n, a.head, env, age) + v * 16);
        if (a.format == 0) {
            reinterpret_cast<uint4 *>(a.out)[idx] = q;
        } else {
            float4 *out = reinterpret_cast<float4 *>(a.out) + idx * 4;
            const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                float4 f;
                f.x = __fdiv_rn(__fsub_rn(static_cast<float>(w[g] & 255u), 128.0f), 128.0f);
                f.y = __fdiv_rn(__fsub_rn(static_cast<float>((w[g] >> 8) & 255u), 128.0f), 128.0f);
                f.z = __fdiv_rn(__fsub_rn(static_cast<float>((w[g] >> 16) & 255u), 128.0f), 128.0f);
                f.w = __fdiv_rn(__fsub_rn(static_cast<float>(w[g] >> 24), 128.0f), 128.0f);
                out[g] = f;
            }
        }
    }
}

__global__ void set_pending_kernel(uint32_t *state, const uint8_t *pending, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t *m = state + static_cast<size_t>(15) * n + i;
    *m = (*m & ~(1u << 21)) | (pending[i] ? 1u << 21 : 0u);
}

// ------------------------------------------------------------------------------------------------ launchers
cudaError_t launch_env(const EnvArgs &a, bool step, cudaStream_t stream)
{
    const int blocks = (a.n + kEnvThreads - 1) / kEnvThreads;
    if (step) env_kernel<true><<<blocks, kEnvThreads, 0, stream>>>(a);
    else env_kernel<false><<<blocks, kEnvThreads, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_observe(const ObserveArgs &a, int grid, cudaStream_t stream)
{
    observe_kernel<<<grid, 256, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_set_pending(uint32_t *state, const uint8_t *pending, int n, cudaStream_t stream)
{
    set_pending_kernel<<<(n + 255) / 256, 256, 0, stream>>>(state, pending, n);
    return cudaGetLastError();
}

} // namespace ahk
