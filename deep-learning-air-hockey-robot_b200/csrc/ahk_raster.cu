// ahk_raster.cu — K2, the rasteriser: draw-record -> uint8 CHW 3x128x128 frame in the HBM ring, written exactly once.
//
// Replaces the Pygame blits of the missing simulator plus DataProcessor.process_observation's resize / transpose /
// stack (gym-air-hockey/gym_air_hockey/processor.py:31-49).  NEAREST resize of a native 450x700 drawing == evaluating
// the drawing at the 128x128 sample points (SPEC.md §8), so nothing is ever drawn at native resolution.
//
// One persistent CTA per SM = 3 groups of 256 threads.  Shared memory: the table image at the sample points (48 KB,
// one bulk async load) shared by the groups, and one 48 KB staging frame per group.  Per frame, per group:
//   phase 0  128 threads: per output row, where the rail line and the arm fall (exact inverse of the sample map);
//   phase 1  all threads: staging = table, with line rows and arm spans substituted by byte masks — work is per
//            128-bit vector, no per-pixel loop;
//   phase 2  a few hundred candidate pixels (bounding boxes of puck and mallets): top-most layer wins, byte stores;
//   phase 3  one thread hands the staged frame to the TMA engine: cp.async.bulk shared -> global (UBLKCP), three
//            16 KB planes; the next frame's phase 0 overlaps the transfer, phase 1 waits for the engine to have READ
//            the staging buffer (bulk_group .read), not for the write to land.
// Frames of `random` episodes are Philox noise and go straight to global memory with 128-bit stores.
#include "ahk_kernels.cuh"
#include "ahk_sim.cuh"

namespace ahk {

namespace {

constexpr uint32_t kNone = 0xFF000000u; // "lower layer shows through"; colours are r | g<<8 | b<<16
__device__ __forceinline__ constexpr uint32_t rgb(uint32_t r, uint32_t g, uint32_t b) { return r | g << 8 | b << 16; }

constexpr uint32_t kLineColour = rgb(235, 235, 225);
constexpr uint32_t kArmEdge = rgb(160, 110, 95), kArmInner = rgb(200, 150, 130);

struct Scene {
    int ipx, ipy, ibx, iby, itx, ity;
    uint32_t use;
};

// native sample column / row of output pixel j / i: Pillow NEAREST, int((j + 0.5) * 450 / 128) (SPEC.md §8)
__device__ __forceinline__ int sample_x(int j) { return ((2 * j + 1) * 225) >> 7; }
__device__ __forceinline__ int sample_y(int i) { return ((2 * i + 1) * 175) >> 6; }

// exact inverses: the output columns j with x0 <= sample_x(j) <= x1 are lo..hi (empty when lo > hi); rows likewise
__device__ __forceinline__ void columns_within(int x0, int x1, int &lo, int &hi)
{
    lo = x0 <= 0 ? 0 : ((128 * x0 + 224) / 225) >> 1;
    hi = x1 < 0 ? -1 : min(127, (((128 * (x1 + 1) + 224) / 225) - 2) >> 1);
}
__device__ __forceinline__ void rows_within(int y0, int y1, int &lo, int &hi)
{
    lo = y0 <= 0 ? 0 : ((64 * y0 + 174) / 175) >> 1;
    hi = y1 < 0 ? -1 : min(127, (((64 * (y1 + 1) + 174) / 175) - 2) >> 1);
}

__device__ __forceinline__ void arm_span(const Scene &sc, int y, int &xc, int &hw)
{ // SPEC.md §7 item 6, valid for y <= ity + 10
    const int t = sc.ity + 10 - y;
    xc = sc.itx + ((t * 5) >> 5);
    hw = 26 + ((t * 3) >> 4);
}

__device__ __forceinline__ uint32_t mallet_colour(int dx, int dy, uint32_t shine, uint32_t knob, uint32_t rim, uint32_t body)
{ // SPEC.md §7 items 4, 5
    const int d2 = dx * dx + dy * dy;
    if (d2 > 576) return kNone;
    if ((dx + 3) * (dx + 3) + (dy + 3) * (dy + 3) <= 9) return shine;
    if (d2 <= 100) return knob;
    return d2 > 400 ? rim : body;
}

// colour of native pixel (x, y) from the layers above the rail line, top-most first; kNone = none of them
__device__ __forceinline__ uint32_t shade_bodies(const Scene &sc, int x, int y)
{
    if ((sc.use & kUseArm) && y <= sc.ity + 10) {
        int xc, hw;
        arm_span(sc, y, xc, hw);
        const int a = abs(x - xc);
        if (a <= hw) return a > hw - 4 ? kArmEdge : kArmInner;
    }
    if (sc.use & kUseTop) {
        const uint32_t c = mallet_colour(x - sc.itx, y - sc.ity, rgb(225, 215, 215), rgb(110, 70, 70), rgb(35, 20, 20), rgb(70, 40, 40));
        if (c != kNone) return c;
    }
    {
        const uint32_t c = mallet_colour(x - sc.ibx, y - sc.iby, rgb(220, 220, 230), rgb(90, 90, 100), rgb(20, 20, 24), rgb(40, 40, 48));
        if (c != kNone) return c;
    }
    if (sc.use & kUsePuck) {
        const int dx = x - sc.ipx, dy = y - sc.ipy, d2 = dx * dx + dy * dy;
        if (d2 <= 324) {
            if ((dx + 5) * (dx + 5) + (dy + 5) * (dy + 5) <= 16) return rgb(240, 120, 130);
            return d2 > 196 ? rgb(150, 20, 30) : rgb(215, 40, 50);
        }
    }
    return kNone;
}

// 16-bit pixel mask of columns lo..hi (absolute) inside the vector that starts at column base
__device__ __forceinline__ uint32_t span_mask(int lo, int hi, int base)
{
    const int a = max(lo - base, 0), b = min(hi - base, 15);
    return a <= b ? ((0xFFFFu >> (15 - b)) & (0xFFFFu << a) & 0xFFFFu) : 0u;
}

// 4 pixel bits -> 4 byte lanes of 0xFF
__device__ __forceinline__ uint32_t byte_lanes(uint32_t nibble) { return ((nibble * 0x00204081u) & 0x01010101u) * 0xFFu; }

__device__ __forceinline__ uint32_t splat(uint32_t colour, int channel) { return ((colour >> (8 * channel)) & 0xFFu) * 0x01010101u; }

__device__ __forceinline__ void bulk_store(void *global_dst, const void *shared_src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 :: "l"(global_dst), "r"(static_cast<uint32_t>(__cvta_generic_to_shared(shared_src))), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// named barrier 1 + group over the kRasterThreads threads of one group (barrier 0 stays __syncthreads)
__device__ __forceinline__ void group_sync(int group) { asm volatile("bar.sync %0, %1;" :: "r"(group + 1), "n"(kRasterThreads) : "memory"); }

__device__ __forceinline__ void mbarrier_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred done;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 done, [%0], %1;\n\t"
        "@!done bra WAIT_%=;\n\t}"
        :: "r"(bar), "r"(parity) : "memory");
}

} // namespace

// One CTA per SM, kRasterGroups independent groups of kRasterThreads threads.  The groups share the table image in
// shared memory; each owns a staging frame, a record batch and a named barrier, and walks its own subset of frames — so
// while one group waits for the TMA engine to drain its staging buffer, the others compose.
__global__ void __launch_bounds__(kRasterGroups * kRasterThreads, 1) raster_staged_kernel(RasterArgs a)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x % kRasterThreads; // thread within the group
    const int grp = threadIdx.x / kRasterThreads;
    const uint4 *s_table = reinterpret_cast<const uint4 *>(smem);                          // [3][128][8] vectors
    uint4 *s_stage = reinterpret_cast<uint4 *>(smem + (1 + grp) * kFrameBytes);             // the frame being composed
    uint8_t *tail = smem + (1 + kRasterGroups) * kFrameBytes + grp * kRasterGroupTail;
    uint2 *s_row = reinterpret_cast<uint2 *>(tail);                                         // [128] per-row spans
    uint4 *s_recs = reinterpret_cast<uint4 *>(tail + kDim * 8);                             // [kRasterThreads] valid records
    uint32_t *s_items = reinterpret_cast<uint32_t *>(tail + kDim * 8 + kRasterThreads * 16); // their item ids
    int *s_count = reinterpret_cast<int *>(tail + kDim * 8 + kRasterThreads * 20);
    __shared__ __align__(8) uint64_t s_table_ready;

    // the table: one bulk async copy global -> shared (TMA), completion on an mbarrier
    const uint32_t bar = static_cast<uint32_t>(__cvta_generic_to_shared(&s_table_ready));
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(kFrameBytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))), "l"(a.table), "r"(kFrameBytes), "r"(bar) : "memory");
    }
    bool table_ready = false;

    const size_t n = static_cast<size_t>(a.n);
    const long long items = 3ll * a.n;
    const long long lanes = static_cast<long long>(gridDim.x) * kRasterGroups; // groups in the whole grid
    bool store_in_flight = false;

    // This group owns items first, first + lanes, ...  Most of the 3n records are empty (only envs that were reset
    // paint the two older slots), and a record read costs a trip to L2: fetch a batch of kRasterThreads records with
    // one load per thread, keep the valid ones in shared memory, then paint from there.
    for (long long batch = static_cast<long long>(blockIdx.x) * kRasterGroups + grp; batch < items; batch += lanes * kRasterThreads) {
        group_sync(grp); // previous batch fully consumed
        if (tid == 0) *s_count = 0;
        group_sync(grp);
        {
            const long long item = batch + static_cast<long long>(tid) * lanes;
            if (item < items) {
                const uint4 rec = __ldg(reinterpret_cast<const uint4 *>(a.drawlist) + item);
                if (rec.w & kDrawValid) {
                    const int at = atomicAdd(s_count, 1);
                    s_recs[at] = rec;
                    s_items[at] = static_cast<uint32_t>(item);
                }
            }
        }
        group_sync(grp);
        const int count = *s_count;

        for (int q = 0; q < count; ++q) {
            const uint4 rec = s_recs[q];
            const uint32_t item = s_items[q];
            const int back = item >= 2u * a.n ? 2 : (item >= static_cast<uint32_t>(a.n) ? 1 : 0); // 0 newest
            const int env = static_cast<int>(item - static_cast<uint32_t>(back) * a.n);
            const int slot = (a.head + 3 - back) % 3;
            uint8_t *frame = a.frames + (static_cast<size_t>(slot) * n + env) * kFrameBytes;

            if (rec.w & kDrawNoise) { // SPEC.md §7 last paragraph
                const uint32_t env_id = a.env_id_base + static_cast<uint32_t>(env);
#pragma unroll 1
                for (int v = tid; v < kFrameVecs; v += kRasterThreads) {
                    uint4 r = philox(static_cast<uint32_t>(v), rec.x, env_id, 2u, a.key);
                    r.x = __vsubus4(r.x, 0x01010101u); // per byte b -> (b * 255) >> 8 == max(b - 1, 0)
                    r.y = __vsubus4(r.y, 0x01010101u);
                    r.z = __vsubus4(r.z, 0x01010101u);
                    r.w = __vsubus4(r.w, 0x01010101u);
                    __stcs(reinterpret_cast<uint4 *>(frame) + v, r);
                }
                continue;
            }

            Scene sc;
            sc.ipx = static_cast<int>(rec.x & 0xFFFFu); sc.ipy = static_cast<int>(rec.x >> 16);
            sc.ibx = static_cast<int>(rec.y & 0xFFFFu); sc.iby = static_cast<int>(rec.y >> 16);
            sc.itx = static_cast<int>(rec.z & 0xFFFFu); sc.ity = static_cast<int>(rec.z >> 16);
            sc.use = rec.w & 7u;

            // ---- phase 0: per-row spans (does not touch the staging buffer, so it overlaps the previous bulk store)
            if (tid < kDim) {
                const int y = sample_y(tid);
                uint32_t line = (y >= sc.iby - 3 && y <= sc.iby + 2) ? 1u : 0u;
                uint32_t spans = 0u; // bytes: outer lo+1, outer hi+1, inner lo+1, inner hi+1; 0 = no arm on this row
                if ((sc.use & kUseArm) && y <= sc.ity + 10) {
                    int xc, hw, olo, ohi, ilo, ihi;
                    arm_span(sc, y, xc, hw);
                    columns_within(xc - hw, xc + hw, olo, ohi);
                    columns_within(xc - hw + 4, xc + hw - 4, ilo, ihi);
                    if (olo <= ohi) spans = static_cast<uint32_t>(olo + 1) | static_cast<uint32_t>(ohi + 1) << 8 |
                                            static_cast<uint32_t>(ilo + 1) << 16 | static_cast<uint32_t>(ihi + 1) << 24;
                }
                s_row[tid] = make_uint2(spans, line);
            }
            if (store_in_flight && tid == 0) bulk_wait_read(); // the TMA engine has finished reading the staging buffer
            if (!table_ready) {
                mbarrier_wait(bar, 0u);
                table_ready = true;
            }
            group_sync(grp);

            // ---- phase 1: staging = table + rail line + arm, one (row, 16-pixel vector) position for all 3 channels
#pragma unroll 1
            for (int pos = tid; pos < kDim * 8; pos += kRasterThreads) {
                const int i = pos >> 3, k = pos & 7;
                uint4 out[3] = {s_table[pos], s_table[1024 + pos], s_table[2048 + pos]};
                const uint2 info = s_row[i];
                if (info.y) {
#pragma unroll
                    for (int c = 0; c < 3; ++c) { const uint32_t s = splat(kLineColour, c); out[c] = make_uint4(s, s, s, s); }
                }
                if (info.x) {
                    const int olo = static_cast<int>(info.x & 255u) - 1, ohi = static_cast<int>((info.x >> 8) & 255u) - 1;
                    const int ilo = static_cast<int>((info.x >> 16) & 255u) - 1, ihi = static_cast<int>(info.x >> 24) - 1;
                    const uint32_t outer = span_mask(olo, ohi, 16 * k);
                    if (outer) {
                        const uint32_t inner = span_mask(ilo, ihi, 16 * k) & outer, edge = outer & ~inner;
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            const uint32_t me = byte_lanes((edge >> (4 * g)) & 15u), mi = byte_lanes((inner >> (4 * g)) & 15u);
                            if (me | mi) {
#pragma unroll
                                for (int c = 0; c < 3; ++c) {
                                    uint32_t *w = reinterpret_cast<uint32_t *>(&out[c]) + g;
                                    *w = (*w & ~(me | mi)) | (splat(kArmEdge, c) & me) | (splat(kArmInner, c) & mi);
                                }
                            }
                        }
                    }
                }
                s_stage[pos] = out[0];
                s_stage[1024 + pos] = out[1];
                s_stage[2048 + pos] = out[2];
            }
            group_sync(grp);

            // ---- phase 2: puck and mallets — candidate pixels of three bounding boxes, 16 columns per box row
            {
                int r_lo[3], r_n[3], c_lo[3];
                const int cx[3] = {sc.ipx, sc.ibx, sc.itx}, cy[3] = {sc.ipy, sc.iby, sc.ity}, rad[3] = {18, 24, 24};
                const bool on[3] = {(sc.use & kUsePuck) != 0u, true, (sc.use & kUseTop) != 0u};
                int total = 0;
#pragma unroll
                for (int o = 0; o < 3; ++o) {
                    int lo, hi, clo, chi;
                    rows_within(cy[o] - rad[o], cy[o] + rad[o], lo, hi);
                    columns_within(cx[o] - rad[o], cx[o] + rad[o], clo, chi);
                    r_lo[o] = lo;
                    c_lo[o] = clo;
                    r_n[o] = (on[o] && hi >= lo && chi >= clo) ? hi - lo + 1 : 0;
                    total += r_n[o] * 16;
                }
                uint8_t *stage = reinterpret_cast<uint8_t *>(s_stage);
#pragma unroll 1
                for (int idx = tid; idx < total; idx += kRasterThreads) {
                    int local = idx, o = 0;
                    if (local >= r_n[0] * 16) { local -= r_n[0] * 16; o = 1; }
                    if (o == 1 && local >= r_n[1] * 16) { local -= r_n[1] * 16; o = 2; }
                    const int i = (o == 0 ? r_lo[0] : (o == 1 ? r_lo[1] : r_lo[2])) + (local >> 4);
                    const int j = (o == 0 ? c_lo[0] : (o == 1 ? c_lo[1] : c_lo[2])) + (local & 15);
                    if (j > 127) continue;
                    const uint32_t c = shade_bodies(sc, sample_x(j), sample_y(i));
                    if (c != kNone) {
                        const int at = i * kDim + j;
                        stage[at] = static_cast<uint8_t>(c);
                        stage[kDim * kDim + at] = static_cast<uint8_t>(c >> 8);
                        stage[2 * kDim * kDim + at] = static_cast<uint8_t>(c >> 16);
                    }
                }
            }
            fence_async_smem(); // generic-proxy writes to shared memory -> visible to the async proxy (TMA)
            group_sync(grp);

            // ---- phase 3: hand the frame to the TMA engine
            if (tid == 0) {
#pragma unroll
                for (int c = 0; c < 3; ++c)
                    bulk_store(frame + c * kDim * kDim, reinterpret_cast<const uint8_t *>(s_stage) + c * kDim * kDim, kDim * kDim);
                bulk_commit();
            }
            store_in_flight = true;
        }
    }
    if (store_in_flight && tid == 0) bulk_wait_all();
    if (!table_ready) mbarrier_wait(bar, 0u); // never exit with the table copy still in flight
}

cudaError_t raster_staged_prepare()
{
    return cudaFuncSetAttribute(raster_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRasterStagedSmem);
}

cudaError_t launch_raster_staged(const RasterArgs &a, int grid, cudaStream_t stream)
{
    raster_staged_kernel<<<grid, kRasterGroups * kRasterThreads, kRasterStagedSmem, stream>>>(a);
    return cudaGetLastError();
}

} // namespace ahk
