// prologue_probe.cu — where the fused kernel's prologue goes: one warp per CTA runs env_update<true> (csrc/ahk_env.cuh)
// on a real SoA state with clock64 stamps after the state load, the tick and the stores; launched twice (cold / warm L2).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/prologue_probe.bin tools/prologue_probe.cu
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#include "../deep-learning-air-hockey-robot_b200/csrc/ahk_env.cuh"
namespace ahk { void set_error_text(const char *) {} }
using namespace ahk;
__global__ void __launch_bounds__(1024, 1) probe(EnvArgs a, long long *out, int per_cta)
{
    extern __shared__ uint8_t smem[];
    const long long t0 = clock64();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int i = blockIdx.x * per_cta + warp;
    long long st[3] = {0, 0, 0};
    uint4 rec[3];
    StatDelta d;
    if (lane == 0 && warp < per_cta && i < a.n) env_update<true>(a, i, rec, d, st);
    __syncthreads();
    const long long t3 = clock64();
    if (lane == 0 && warp < per_cta && i < a.n) {
        long long *o = out + static_cast<size_t>(i) * 5;
        o[0] = st[0] - t0; o[1] = st[1] - t0; o[2] = st[2] - t0; o[3] = t3 - t0; o[4] = rec[0].x;
    }
}
int main()
{
    const int n = 4096, per_cta = 28, grid = (n + per_cta - 1) / per_cta;
    uint32_t *state; float *reward; uint8_t *b8; uint64_t *stats; long long *out;
    cudaMalloc(&state, 16 * n * 4); cudaMalloc(&reward, n * 4); cudaMalloc(&b8, 4 * n); cudaMalloc(&stats, 256); cudaMalloc(&out, n * 5 * 8);
    std::vector<uint32_t> h(16 * n);
    auto f = [](float v) { uint32_t u; memcpy(&u, &v, 4); return u; };
    for (int i = 0; i < n; ++i) {
        float v[12] = {100.f + (i * 37 % 250), 200.f + (i * 91 % 300), (i % 7) - 3.f, (i % 5) - 2.f, 225, 600, 0, 0, 225, 100, 0, 0};
        for (int w = 0; w < 12; ++w) h[w * n + i] = f(v[w]);
        h[12 * n + i] = 1; h[13 * n + i] = 1; h[14 * n + i] = 0; h[15 * n + i] = 7u << 18;
    }
    cudaMemcpy(state, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    EnvArgs a{};
    a.state = state; a.reward = reward; a.done = b8; a.robot_label = b8 + n; a.human_label = b8 + 2 * n; a.info = b8 + 3 * n;
    a.stats = nullptr; a.key = make_uint2(1, 2); a.n = n; a.auto_reset = 1; a.max_episode_steps = 200;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int rep = 0; rep < 4; ++rep) {
        probe<<<grid, 1024, 198 * 1024>>>(a, out, per_cta);
        cudaDeviceSynchronize();
        std::vector<long long> o(n * 5);
        cudaMemcpy(o.data(), out, o.size() * 8, cudaMemcpyDeviceToHost);
        double m[4] = {0, 0, 0, 0};
        for (int i = 0; i < n; ++i) for (int k = 0; k < 4; ++k) m[k] += o[i * 5 + k];
        printf("launch %d: cycles from CTA start: state loaded %.0f, tick done %.0f, stores issued %.0f, after barrier %.0f  (%s)\n", rep,
               m[0] / n, m[1] / n, m[2] / n, m[3] / n, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
