// tick_probe.cu — how long is one game tick (csrc/ahk_sim.cuh sim_tick) as a dependent chain on one warp, with a cold
// and with a warm instruction cache?  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/tick_probe.bin tools/tick_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../deep-learning-air-hockey-robot_b200/csrc/ahk_sim.cuh"
using namespace ahk;
__global__ void probe(float *state, long long *cycles, int iters, int lanes)
{
    const int lane = threadIdx.x;
    Sim s;
    const float *p = state + (blockIdx.x * 32 + lane) * 12;
    s.puck = Body{p[0], p[1], p[2], p[3]};
    s.bot = Body{p[4], p[5], p[6], p[7]};
    s.top = Body{p[8], p[9], p[10], p[11]};
    s.serve_count = 1;
    uint2 key = make_uint2(1, 2);
    for (int k = 0; k < iters; ++k) {
        const long long t0 = clock64();
        if (lane < lanes) {
            Tick t;
            sim_tick(s, kActNone, kActNone, kUseAll, blockIdx.x * 32 + lane, key, t);
            if (t.info == 0xFFFFFFFFu) state[0] = 1.0f;
        }
        __syncwarp();
        const long long t1 = clock64();
        if (lane == 0) cycles[blockIdx.x * iters + k] = t1 - t0;
    }
    if (lane < lanes) state[(blockIdx.x * 32 + lane) * 12] = s.puck.x;
}
int main()
{
    const int blocks = 148, iters = 8;
    float *st; long long *cy;
    cudaMalloc(&st, blocks * 32 * 12 * 4); cudaMalloc(&cy, blocks * iters * 8);
    float h[148 * 32 * 12];
    for (int i = 0; i < blocks * 32; ++i) {
        float *p = h + i * 12;
        p[0] = 100.f + (i * 37 % 250); p[1] = 200.f + (i * 91 % 300); p[2] = (i % 7) - 3.f; p[3] = (i % 5) - 2.f;
        p[4] = 225; p[5] = 600; p[6] = 0; p[7] = 0; p[8] = 225; p[9] = 100; p[10] = 0; p[11] = 0;
    }
    for (int lanes : {1, 32}) {
        for (int rep = 0; rep < 2; ++rep) {
            cudaMemcpy(st, h, sizeof h, cudaMemcpyHostToDevice);
            probe<<<blocks, 32>>>(st, cy, iters, lanes);
            cudaDeviceSynchronize();
            long long c[148 * 8];
            cudaMemcpy(c, cy, sizeof c, cudaMemcpyDeviceToHost);
            printf("lanes %2d launch %d: cycles per tick, mean over blocks:", lanes, rep);
            for (int k = 0; k < iters; ++k) { double m = 0; for (int b = 0; b < blocks; ++b) m += c[b * iters + k]; printf(" %.0f", m / blocks); }
            printf("\n");
        }
    }
    return 0;
}
