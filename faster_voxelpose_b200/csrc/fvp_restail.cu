// The tail of P2PNet in one kernel (reference lib/models/cnns_2d.py: the second convolution, skip addition and ReLU of
// skip_res1, Res2DBlock.forward :72-76; `x = decoder_upsample1(x) + skip_x1`, :179-180; `output_layer`, :212 / :221):
//
//     out[n][co][y][x] = bo[co] + sum_c wo[co][c] * ( relu( conv3x3(h, w2)[c] + b2[c] + xres[c] ) + up[c] )          (y, x, n implied)
//
// h = the block's first convolution (ReLU done), xres = the block's input, up = the last up-sampling as the 1x1 convolution to
// 4 x 32 channels whose pixel shuffle is taken here.  Before, this was cuDNN's bare convolution (0.20 ms on the 960 planes of
// the bench) and fvp_pointwise_planar_res (0.42 ms); 503 MB of convolution output went to HBM and came back.
//
//   * ONE RASTER FOR THE WHOLE BATCH.  The images are laid end to end as a raster of pitch W + 1 with H + 1 rows per image: the
//     one extra column is the right padding of a row and the left padding of the next, the one extra row the bottom padding of
//     an image and the top padding of the next.  A tile is 256 consecutive raster positions wherever they fall (96.9 % of the
//     positions are pixels at 64 x 64; tiles of whole rows would waste 18-31 %), staged with a halo of W + 2 positions either
//     side as [channel / 4][position][4 floats] -- the K-major no-swizzle operand layout whose descriptor may start at any row,
//     so tap (ky, kx) is a start address shifted by ky * (W + 1) + kx (csrc/fvp_res2d.cu).
//   * 36 tcgen05.mma.ws (M = 32 filters, N = 256 positions, K = 8 channels; TF32 in, fp32 accumulate in TMEM) per tile.
//   * Roles around mbarriers: warps 0-7 drain one of two accumulators (thread = output channel x 32 positions: xres and up are
//     fetched before the wait on the tensor core, so their latency hides behind it), warp 8 issues from one lane, warps 9-11
//     stage rasters into three slots.  The raster slot of a finished tile is where its 32-channel result is turned around for
//     the 1x1 layer (thread = position, fp32 FMAs, the weights as kernel parameters): nothing but `out` is written.
#include <algorithm>
#include <cstdint>

#include "fvp_common.cuh"

namespace fvp {

constexpr int RT_C = 32;                     // channels of the block
constexpr int RT_DRAIN_WARPS = 8, RT_LOAD_WARPS = 3, RT_THREADS = 32 * (RT_DRAIN_WARPS + 1 + RT_LOAD_WARPS);
constexpr int RT_SLOTS = 3;
constexpr int RT_MAX_COUT = 20;
constexpr int RT_STAGE_PITCH = 36;           // floats per staged position: 16-byte rows, conflict-free written by channel and read by position

struct ResTailParams {
    const float *h;        // [N][H][W][32]   relu(conv1(x) + b1)
    const float *xres;     // [N][H][W][32]   the block's input
    const float *up;       // [N][H/2][W/2][4][32]   un-shuffled up-sampling, (a, b) = (y & 1, x & 1) major
    const float4 *w2;      // [9][8][32] float4: tap, channel / 4, output channel; TF32-rounded
    float *out;            // [N][cout][H][W]
    int N, H, W, PW, NPX, cout;
    long ntiles, qout;
    long long *prof;       // null, or 12 counters: cycles per phase of CTA 0's roles (fvp_res_tail_profile)
    float b2[RT_C];
    float wo[RT_MAX_COUT * RT_C];
    float bo[RT_MAX_COUT];
};

__device__ __forceinline__ unsigned rt_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void rt_wait(unsigned bar, unsigned parity) {
    unsigned done = 0;
    while (!done)
        // the time hint (ns) bounds how long the hardware may park the thread per attempt: fewer spin instructions beside the
        // warps that have work
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar), "r"(parity), "r"(1000000u)
                     : "memory");
}
__device__ __forceinline__ void rt_arrive(unsigned bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void rt_commit(unsigned bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

constexpr unsigned RT_DESC_HI = (128u >> 4) | (1u << 14);   // stride byte offset 128, descriptor version 1
constexpr unsigned RT_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((256u >> 3) << 17) | ((32u >> 4) << 24);   // f32 += tf32 x tf32, M 32, N 256

__device__ __forceinline__ void rt_mma(unsigned d_tmem, unsigned a_lo, unsigned b_lo, unsigned accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 da, {%1, %5};\n\tmov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.ws.cta_group::1.kind::tf32.collector::b0::discard [%0], da, db, %3, p;\n\t}" ::"r"(d_tmem),
        "r"(a_lo), "r"(b_lo), "r"(RT_IDESC), "r"(accumulate), "r"(RT_DESC_HI)
        : "memory");
}

__device__ __forceinline__ void rt_tmem_ld32(unsigned addr, float (&v)[32]) {
    unsigned u[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,"
        "%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]), "=r"(u[8]), "=r"(u[9]),
          "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15]), "=r"(u[16]), "=r"(u[17]), "=r"(u[18]),
          "=r"(u[19]), "=r"(u[20]), "=r"(u[21]), "=r"(u[22]), "=r"(u[23]), "=r"(u[24]), "=r"(u[25]), "=r"(u[26]), "=r"(u[27]),
          "=r"(u[28]), "=r"(u[29]), "=r"(u[30]), "=r"(u[31])
        : "r"(addr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(u[j]);
}

// The 32 consecutive raster positions a drain thread owns (tile, TMEM lane quadrant, column half): at most one row change
// (PW > 32), at element `split`; element split - 1 is then the padding column.
struct RtGroup {
    unsigned ma, mb;       // bit j: element j is a pixel, in the row of the first position / in the next row
    int na, ya, nb, yb;    // image and row of either raster row (y == H: the padding row between two images)
    int c0, split;         // column of the first position, PW - c0
    int pixa, pixb;        // pixel index (n H + y) W + x of element j: pixa + j / pixb + j (virtual where masked)
};

__device__ __forceinline__ RtGroup rt_group(const ResTailParams &p, int tile, int quad, int half) {
    RtGroup g;
    const int q0 = p.PW + tile * 256 + 64 * quad + 32 * half;
    const int g0 = q0 / p.PW - 1;                        // >= 0: image row counter with the padding rows in
    g.c0 = q0 - (g0 + 1) * p.PW; g.split = p.PW - g.c0;
    g.na = g0 / (p.H + 1); g.ya = g0 - g.na * (p.H + 1);
    g.nb = g.ya == p.H ? g.na + 1 : g.na; g.yb = g.ya == p.H ? 0 : g.ya + 1;
    g.ma = 0u; g.mb = 0u;
    if (g.ya < p.H && g.na < p.N) {
        const int hi = min(min(p.W - g.c0, g.split), 32);
        if (hi > 0) g.ma = hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u);
    }
    if (g.split < 32 && g.yb < p.H && g.nb < p.N) {
        const int hi = min(g.split + p.W, 32);
        g.mb = (hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << g.split) - 1u);
    }
    g.pixa = (g.na * p.H + g.ya) * p.W + g.c0; g.pixb = (g.nb * p.H + g.yb) * p.W - g.split;
    return g;
}

// xres and the un-shuffled up-sampling -- pixel (n, y, x) -> ((n H/2 + y/2) W/2 + x/2) * 4 + (y & 1) * 2 + (x & 1) -- of channel
// `lane` at the group's 32 positions.  Tiles away from both ends of the tensors take the lean form: no predicates, every
// element loads from the row it would belong to (the padding column with the second row), masked elements read neighbouring
// pixels that exist; the addresses are a base pointer per (row, parity of j) plus a compile-time offset.
__device__ __forceinline__ void rt_fetch(const ResTailParams &p, const RtGroup &g, long tile, int lane, float (&r)[32], float (&u)[32]) {
    const int W2 = p.W >> 1, H2 = p.H >> 1;
    const int rowa = ((g.na * H2 + (g.ya >> 1)) * W2) * 4 + (g.ya & 1) * 2, rowb = ((g.nb * H2 + (g.yb >> 1)) * W2) * 4 + (g.yb & 1) * 2;
    if (tile >= 1 && tile + 3 <= p.ntiles) {
        const bool oka = g.ya < p.H, okb = g.yb < p.H;                  // one of them is a row of an image
        const float *ra = p.xres + (long)g.pixa * RT_C + lane, *rb = p.xres + (long)g.pixb * RT_C + lane;
        if (!oka) ra = rb + g.split * RT_C;
        if (!okb) rb = ra - 32 * RT_C;
        // element j of row A is x = c0 + j: from the even pixel at or below c0 that is j + par on, i.e. entry
        // (j >> 1) * 4 + (j & 1) + {par, 3 par} for even / odd j; row B the same from x = j - split
        const int para = g.c0 & 1, parb = g.split & 1;
        const int ubasea = rowa + ((g.c0 - para) >> 1) * 4, ubaseb = rowb - (g.split + parb) * 2;
        const float *uae = p.up + (long)(ubasea + para) * RT_C + lane, *uao = p.up + (long)(ubasea + 3 * para) * RT_C + lane;
        const float *ube = p.up + (long)(ubaseb + parb) * RT_C + lane, *ubo = p.up + (long)(ubaseb + 3 * parb) * RT_C + lane;
        if (!oka) { uae = ube + 2 * g.split * RT_C; uao = ubo + 2 * g.split * RT_C; }
        if (!okb) { ube = uae - 64 * RT_C; ubo = uao - 64 * RT_C; }
        const int sa = g.split - 1;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const bool second = j >= sa;
            const float *pr = (second ? rb : ra) + j * RT_C;
            const float *pv = ((j & 1) ? (second ? ubo : uao) : (second ? ube : uae)) + ((j >> 1) * 4 + (j & 1)) * RT_C;
            FVP_ASSERT(pr >= p.xres && pr < p.xres + (size_t)p.N * p.H * p.W * RT_C);     // masked elements read pixels that exist
            FVP_ASSERT(pv >= p.up && pv < p.up + (size_t)p.N * p.H * p.W * RT_C);
            FVP_ASSERT(!(((g.ma | g.mb) >> j) & 1u) || pr == p.xres + (size_t)((((g.ma >> j) & 1u) ? g.pixa : g.pixb) + j) * RT_C + lane);
            r[j] = __ldg(pr);
            u[j] = __ldg(pv);
        }
        return;
    }
    const float *ra = p.xres + (size_t)g.pixa * RT_C + lane, *rb = p.xres + (size_t)g.pixb * RT_C + lane;
    const float *pu = p.up + lane;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const bool va = (g.ma >> j) & 1u, vb = (g.mb >> j) & 1u;
        const int x = va ? g.c0 + j : j - g.split;
        const int upix = (va ? rowa : rowb) + (x >> 1) * 4 + (x & 1);
        r[j] = 0.0f; u[j] = 0.0f;
        if (va || vb) {
            r[j] = __ldg((va ? ra : rb) + j * RT_C);
            u[j] = __ldg(pu + (size_t)upix * RT_C);
        }
    }
}

template <int COUT>
__global__ void __launch_bounds__(RT_THREADS, 1)
res_tail_kernel(const __grid_constant__ ResTailParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // [0..2] raster slot staged, [3..5] raster slot free (its tile drained), [6,7] accumulator complete, [8,9] accumulator read
    __shared__ __align__(8) unsigned long long s_bar[10];
    __shared__ unsigned s_tmem;
    // the 1x1 layer's filters [channel][output]: kernel parameters reach an FFMA through one uniform-register load each, which two
    // warps per scheduler cannot hide; a broadcast LDS.128 brings four
    __shared__ __align__(16) float s_wo[RT_C][COUT];
    float4 *sX = reinterpret_cast<float4 *>(smem_raw);            // [RT_SLOTS][8][NPX]
    float4 *sW = sX + (size_t)RT_SLOTS * 8 * p.NPX;               // [9][8][32]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int PW = p.PW;
    const unsigned bar0 = rt_u32(&s_bar[0]);
    auto BAR = [&](int i) { return bar0 + 8u * (unsigned)i; };

    // everything the tensor core may read is finite from the start
    for (int e = tid; e < RT_SLOTS * 8 * p.NPX; e += RT_THREADS) sX[e] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int e = tid; e < 9 * 8 * 32; e += RT_THREADS) sW[e] = __ldg(p.w2 + e);
    for (int e = tid; e < RT_C * COUT; e += RT_THREADS) s_wo[e / COUT][e % COUT] = p.wo[(e % COUT) * RT_C + e / COUT];
    if (tid == 0) {
        for (int s = 0; s < RT_SLOTS; ++s) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(BAR(0 + s)), "r"(RT_LOAD_WARPS * 32));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(BAR(3 + s)), "r"(RT_DRAIN_WARPS * 32));
        }
        for (int a = 0; a < 2; ++a) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(BAR(6 + a)));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(BAR(8 + a)), "r"(RT_DRAIN_WARPS * 32));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(rt_u32(&s_tmem)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tm = s_tmem;
    const unsigned aX = rt_u32(sX), aW = rt_u32(sW);
    const unsigned slot_bytes = (unsigned)(8 * p.NPX * 16);
    const int halo = PW + 1;                      // raster entries in front of a tile's first position

    if (warp < RT_DRAIN_WARPS) {
        // ------------------------------------------------------------------------------------------------ drain
        const int quad = warp & 3, half = warp >> 2;      // TMEM lanes 32 * quad .., columns 32 * half ..
        const float b2 = p.b2[lane];
        const size_t plane = (size_t)p.H * p.W;
        float r[32], u[32];                               // xres and up of this thread's channel at its 32 positions
        RtGroup gr;
        if ((long)blockIdx.x < p.ntiles) {
            gr = rt_group(p, (int)blockIdx.x, quad, half);
            rt_fetch(p, gr, (long)blockIdx.x, lane, r, u);
        }
        long it = 0;
        for (long tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
            const int s = (int)(it % RT_SLOTS), a = (int)(it & 1);
            const bool prof = p.prof && blockIdx.x == 0 && tid == 0;
            long long t0 = 0, t1 = 0, t2 = 0, t3 = 0;
            if (prof) t0 = clock64();
            rt_wait(BAR(6 + a), (unsigned)((it >> 1) & 1));
            if (prof) t1 = clock64();
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            float v[32];
            rt_tmem_ld32(tm + ((unsigned)(quad * 32) << 16) + (unsigned)(a * 64 + half * 32), v);
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            rt_arrive(BAR(8 + a));                        // the accumulator may be overwritten
            // ---- turn the 32 x 32 block around in the raster slot this tile no longer needs: [position][channel]
            float *stage = reinterpret_cast<float *>(smem_raw + (size_t)s * slot_bytes) + warp * (32 * RT_STAGE_PITCH);
#pragma unroll
            for (int j = 0; j < 32; ++j) stage[j * RT_STAGE_PITCH + lane] = fmaxf(v[j] + b2 + r[j], 0.0f) + u[j];
            __syncwarp();
            if (prof) t2 = clock64();
            // ---- this thread's output pixel (thread = position `lane` from here on), then the next tile's operands: their
            //      latency hides behind the 1x1 layer
            const bool mine_a = (gr.ma >> lane) & 1u, mine_b = (gr.mb >> lane) & 1u;
            float *po = nullptr;
            if (mine_a || mine_b) {
                const int n = mine_a ? gr.na : gr.nb;
                po = p.out + ((size_t)n * p.cout - n) * plane + (size_t)((mine_a ? gr.pixa : gr.pixb) + lane);   // (n H + y) W + x
                FVP_ASSERT(n >= 0 && n < p.N && (size_t)((mine_a ? gr.pixa : gr.pixb) + lane) >= (size_t)n * plane &&
                           (size_t)((mine_a ? gr.pixa : gr.pixb) + lane) < (size_t)(n + 1) * plane);
            }
            if (tile + gridDim.x < p.ntiles) {
                gr = rt_group(p, (int)(tile + gridDim.x), quad, half);
                rt_fetch(p, gr, tile + gridDim.x, lane, r, u);
            }
            if (prof) t3 = clock64();
            // ---- 1x1 output layer
            float2 acc[COUT / 2];
#pragma unroll
            for (int co = 0; co < COUT / 2; ++co) acc[co] = make_float2(p.bo[2 * co], p.bo[2 * co + 1]);
            const float4 *mine = reinterpret_cast<const float4 *>(stage + lane * RT_STAGE_PITCH);
#pragma unroll
            for (int c4 = 0; c4 < RT_C / 4; ++c4) {
                const float4 t4 = mine[c4];
                const float t[4] = {t4.x, t4.y, t4.z, t4.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float2 tt = make_float2(t[i], t[i]);
#pragma unroll
                    for (int o4 = 0; o4 < COUT / 4; ++o4) {
                        const float4 w = *reinterpret_cast<const float4 *>(&s_wo[4 * c4 + i][4 * o4]);
                        acc[2 * o4] = __ffma2_rn(make_float2(w.x, w.y), tt, acc[2 * o4]);
                        acc[2 * o4 + 1] = __ffma2_rn(make_float2(w.z, w.w), tt, acc[2 * o4 + 1]);
                    }
                }
            }
            rt_arrive(BAR(3 + s));                        // the raster slot may be refilled
            if (po) {
#pragma unroll
                for (int co = 0; co < COUT; ++co)
                    if (co < p.cout) po[(size_t)co * plane] = (co & 1) ? acc[co >> 1].y : acc[co >> 1].x;
            }
            if (prof) {
                const long long t4 = clock64();
                p.prof[0] += t1 - t0; p.prof[1] += t2 - t1; p.prof[2] += t3 - t2; p.prof[3] += t4 - t3; p.prof[11] += 1;
            }
        }
    } else if (warp == RT_DRAIN_WARPS) {
        // ------------------------------------------------------------------------------------------------ MMA issue
        if (lane == 0) {
            long it = 0;
            const unsigned aw0 = (aW >> 4) & 0x3fffu;
            for (long tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
                const int s = (int)(it % RT_SLOTS), a = (int)(it & 1);
                const bool prof = p.prof && blockIdx.x == 0;
                long long t0 = 0, t1 = 0, t2 = 0;
                if (prof) t0 = clock64();
                rt_wait(BAR(0 + s), (unsigned)((it / RT_SLOTS) & 1));           // raster staged
                if (prof) t1 = clock64();
                rt_wait(BAR(8 + a), (unsigned)(((it >> 1) & 1) ^ 1));           // accumulator read (free at first)
                if (prof) t2 = clock64();
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the loaders' generic-proxy writes -> the tensor core
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const unsigned bx = (((aX + s * slot_bytes) >> 4) & 0x3fffu) | ((unsigned)p.NPX << 16);   // LBO = NPX * 16 bytes
                const unsigned d = tm + (unsigned)(a * 64);
#pragma unroll 1
                for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        const unsigned aw = (aw0 + (unsigned)((ky * 3 + kx) * 8 * 32)) | (32u << 16);    // LBO = 32 rows
                        const unsigned bt = bx + (unsigned)(ky * PW + kx);
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            rt_mma(d, aw + (unsigned)(ks * 64), bt + (unsigned)(ks * 2 * p.NPX), (ky | kx | ks) != 0);
                    }
                }
                rt_commit(BAR(6 + a));
                if (prof) { p.prof[4] += t1 - t0; p.prof[5] += t2 - t1; p.prof[6] += clock64() - t2; }
            }
        }
    } else {
        // ------------------------------------------------------------------------------------------------ raster loader
        const int lt = tid - 32 * (RT_DRAIN_WARPS + 1);          // 0 .. 95
        const int k = lt & 7;                                    // channel chunk (plane) of this thread
        const int npos = 256 + 2 * halo;
        long it = 0;
        for (long tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
            const int s = (int)(it % RT_SLOTS);
            const bool prof = p.prof && blockIdx.x == 0 && lt == 0;
            long long t0 = 0, t1 = 0, t2 = 0;
            if (prof) t0 = clock64();
            rt_wait(BAR(3 + s), (unsigned)(((it / RT_SLOTS) & 1) ^ 1));         // slot free (free at first)
            if (prof) t1 = clock64();
            const int qs = PW + (int)tile * 256 - halo;                         // raster position of entry 0 (>= -1)
            const unsigned dst0 = aX + s * slot_bytes + (unsigned)(k * p.NPX * 16);
            // entry e0 is decoded once -- (image, row with the padding row in, column); + PW keeps the division unsigned -- and
            // then walked: 12 entries on, next row when the column passes the pitch, next image when the row passes H
            constexpr int STEP = (RT_LOAD_WARPS * 32) >> 3;
            const int e0 = lt >> 3;
            const unsigned qq = (unsigned)(qs + PW + e0);
            const int g = (int)(qq / (unsigned)PW) - 2;          // rows since the first image row (< 0: above it)
            int c = (int)(qq % (unsigned)PW);
            int n = g >= 0 ? g / (p.H + 1) : 0, y = g >= 0 ? g - n * (p.H + 1) : g;
            bool rowok = y >= 0 && y < p.H && n < p.N;
            const float *row = p.h + ((long)(n * p.H + y) * p.W) * RT_C + k * 4;      // virtual where !rowok
            // (LDG.128 + STS.128 through registers would cost a third of the shared-memory wavefronts -- a 16-byte cp.async that
            // scatters a pixel's chunks over the eight planes is written sector by sector as the data returns, ncu: 24
            // wavefronts per warp instruction where 4 would do -- but exposes the load latency once per tile: 0.70 ms against 0.44.)
            unsigned dst = dst0 + (unsigned)(e0 * 16);
#pragma unroll 4
            for (int e = e0; e < npos; e += STEP) {
                const bool ok = rowok && c < p.W;
                FVP_ASSERT(!ok || ((long)(n * p.H + y) * p.W + c >= 0 && (long)(n * p.H + y) * p.W + c < (long)p.N * p.H * p.W));
                FVP_ASSERT(dst + 16 <= aX + (unsigned)(s + 1) * slot_bytes && e < p.NPX);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(ok ? row + c * RT_C : p.h), "r"(ok ? 16 : 0));
                c += STEP; dst += STEP * 16;
                if (c >= PW) {                                   // PW > STEP: one row at most
                    c -= PW;
                    if (++y > p.H) { y = 0; ++n; }               // row H is the padding between two images
                    rowok = y >= 0 && y < p.H && n < p.N;
                    row = p.h + ((long)(n * p.H + y) * p.W) * RT_C + k * 4;
                }
            }
            if (prof) t2 = clock64();
            // arrives on the barrier when this thread's copies have landed; the loader goes on to the next slot meanwhile
            asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(BAR(0 + s)) : "memory");
            if (prof) { p.prof[7] += t1 - t0; p.prof[8] += t2 - t1; p.prof[9] += clock64() - t2; }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tm));
}

}  // namespace fvp

using namespace fvp;

static bool res_tail_plan(int H, int W, int cout, ResTailParams &p, size_t &smem) {
    if (W < 32 || H < 2 || (W & 1) || (H & 1) || cout < 1 || cout > RT_MAX_COUT) return false;
    p.PW = W + 1;
    const int npos = 256 + 2 * (p.PW + 1);
    p.NPX = ((npos + 6) / 8) * 8 + 1;          // = 1 mod 8: the eight channel planes start in different bank groups
    smem = (size_t)RT_SLOTS * 8 * p.NPX * 16 + (size_t)9 * 8 * 32 * 16;
    // the slot also stages 8 warps x 32 positions x 33 floats for the 1x1 layer
    return smem <= 224 * 1024 && (size_t)8 * p.NPX * 16 >= (size_t)RT_DRAIN_WARPS * 32 * RT_STAGE_PITCH * 4;
}

static long long *g_res_tail_prof = nullptr;

extern "C" int fvp_res_tail_profile(long long *counters) {
    g_res_tail_prof = counters;
    return FVP_OK;
}

extern "C" int fvp_res_tail_supported(int H, int W, int cout) {
    ResTailParams p;
    size_t smem;
    return res_tail_plan(H, W, cout, p, smem) ? 1 : 0;
}

extern "C" int fvp_res_tail_tf32(const float *h, const float *xres, const float *up, int N, int H, int W, const float *w2packed,
                                 const float *b2, const float *wout, const float *bout, int cout, float *out, void *stream) {
    FVP_REQUIRE(h && xres && up && w2packed && b2 && wout && bout && out, "fvp_res_tail_tf32: null pointer");
    FVP_REQUIRE(N >= 0 && H >= 1 && W >= 1, "fvp_res_tail_tf32: bad shape N=%d H=%d W=%d", N, H, W);
    FVP_REQUIRE(((uintptr_t)h & 15) == 0 && ((uintptr_t)w2packed & 15) == 0, "fvp_res_tail_tf32: misaligned pointer");
    ResTailParams q;
    size_t smem = 0;
    FVP_REQUIRE(res_tail_plan(H, W, cout, q, smem),
                "fvp_res_tail_tf32: H=%d W=%d cout=%d outside the kernel's tiles (even sizes, W >= 32, cout <= %d: fvp_res_tail_supported)",
                H, W, cout, RT_MAX_COUT);
    FVP_REQUIRE((long)(N + 1) * (H + 1) * (W + 1) < (1L << 31) - 1024, "fvp_res_tail_tf32: N * (H + 1) * (W + 1) must stay below 2^31");
    if (N == 0) return FVP_OK;
    q.h = h; q.xres = xres; q.up = up; q.w2 = reinterpret_cast<const float4 *>(w2packed); q.out = out;
    q.N = N; q.H = H; q.W = W; q.cout = cout;
    q.prof = g_res_tail_prof;
    q.qout = (long)N * (H + 1) * q.PW;
    q.ntiles = (q.qout + 255) / 256;
    for (int i = 0; i < RT_C; ++i) q.b2[i] = b2[i];
    for (int i = 0; i < RT_MAX_COUT * RT_C; ++i) q.wo[i] = i < cout * RT_C ? wout[i] : 0.0f;
    for (int i = 0; i < RT_MAX_COUT; ++i) q.bo[i] = i < cout ? bout[i] : 0.0f;
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = (int)std::min<long>(q.ntiles, (long)sms);
    if (cout <= 16) {
        FVP_CUDA(cudaFuncSetAttribute(res_tail_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        res_tail_kernel<16><<<grid, RT_THREADS, smem, (cudaStream_t)stream>>>(q);
    } else {
        FVP_CUDA(cudaFuncSetAttribute(res_tail_kernel<RT_MAX_COUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        res_tail_kernel<RT_MAX_COUT><<<grid, RT_THREADS, smem, (cudaStream_t)stream>>>(q);
    }
    return check_launch("res_tail_kernel");
}
