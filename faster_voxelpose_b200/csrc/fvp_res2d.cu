// Res2DBlock of P2PNet / CenterNet at 32 output channels (reference lib/models/cnns_2d.py:39-76:
// out = ReLU(BN(conv3x3(ReLU(BN(conv3x3(x))))) + skip(x)), skip = identity or BN(conv1x1)) as ONE kernel on the
// 5th-generation tensor cores: both convolutions are tcgen05 TF32 implicit GEMMs with their accumulators in tensor
// memory, and the activation between them never leaves the SM.  BatchNorm is folded into the filters by the caller.
//
// Formulation (every property used here is checked on the device by tools/umma_probe.cu):
//   * A tile is TH output rows x the whole image width.  Its input (2-pixel halo) is staged in shared memory as a pixel
//     RASTER of pitch PW = W + 4, one plane per group of four channels: [channel / 4][position][4 floats].  That is the
//     canonical K-major NO-SWIZZLE operand layout (rows 16 bytes apart, SBO = 128 B, LBO = plane size), and the one
//     layout whose descriptor may start at ANY row -- so tap (dy, dx) of a 3x3 stencil is the same raster read through a
//     start address shifted by (dy-1) * PW + (dx-1) positions.  No im2col, no copies per tap.
//   * tcgen05.mma.ws ("weight stationary", M = 32): A = the 32 filters of one tap and eight input channels, B = 256
//     consecutive raster positions, D[cout][position] accumulates over 9 taps x Cin/8 steps in TMEM at lane
//     cout + 32 * (position / 64), column position % 64.  With N = Cout = 32 the ordinary M = 128 form spends 65.8
//     cycles per 128 positions whatever N is; this form spends 78.5 cycles per 256 (measured) -- the operand fetch from
//     shared memory (9 KB per instruction) is what bounds it, at 43 % of the TF32 peak.
//   * The first convolution is evaluated on the raster positions of rows -1 .. TH (the halo the second one needs); its
//     epilogue (tcgen05.ld, + bias, ReLU, zero outside the image = the zero padding of the second convolution, round to
//     TF32) writes the result back to shared memory in the same raster layout -- over the input when Cin == 32, whose
//     job is done once the first accumulators are complete.  A thread of the epilogue is one output channel (its TMEM
//     lane) and walks 32 positions; a warp therefore stores whole 128-byte pixel records to global memory.
//   * A 1x1 skip convolution is Cin/8 more MMAs per tile of the second accumulator, issued in front of the second
//     convolution; an identity skip is added in the last epilogue from global memory (the tile was read moments ago: L2).
//   * Roles: warp 8 issues every MMA from one lane (two 32-bit adds build the next pair of descriptors) and commits one
//     mbarrier per accumulator tile; warps 0-7 drain tiles as they complete (tile t while the tensor core is on t+1);
//     warps 9-10 stage the next tile's raster with cp.async once the skip conv has read this one, and their copies arrive
//     on an mbarrier that only the issuing thread waits for: it queues the FIRST convolution of tile i + 1 behind the
//     SECOND one of tile i, so the tensor core works through epilogue 2 (15.3 k cycles per tile, 13.7 k of them MMAs).
//     With Cin == 32 the intermediate raster takes the place of the input raster, so the first epilogue has to wait for
//     all of the first convolution, the next raster can only be staged under the last epilogue (by the draining warps) and
//     nothing of the next tile can be queued early -- which is why that variant loses to cuDNN and is not routed.
// TF32: operands are read as fp32 words of which the tensor core uses the upper 19 bits (filters are rounded to TF32
// by the caller, the intermediate activation here); accumulation is fp32.  Used only where torch itself would run TF32
// convolutions (torch.backends.cudnn.allow_tf32, its default and what the reference runs with).
#include <algorithm>
#include <cstdint>

#include "fvp_common.cuh"

namespace fvp {

constexpr int R2_EPI_WARPS = 8, R2_LOAD_WARPS = 2, R2_THREADS = 32 * (R2_EPI_WARPS + 1 + R2_LOAD_WARPS), R2_COUT = 32, R2_ACC2_COL = 256;   // warp 8 issues the MMAs, warps 9-10 stage rasters (PIPE)

struct Res2dParams {
    const float *x;           // [N][H][W][CIN]
    float *out;               // [N][H][W][32]
    const float4 *w1, *w2, *ws;  // packed filters [tap][CIN/4][32] float4, [tap][8][32], [CIN/4][32] (ws null = identity)
    const float *b1, *b2;
    long long *prof;          // null, or 8 counters: cycles CTA 0 spends per phase (fvp_res2d_profile)
    int N, H, W, TH, PW, NPXX, NPXM, nt1, nt2, tiles_y;
    long ntiles;
};

__device__ __forceinline__ unsigned r2_smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

// K-major, no swizzle: start address, leading (between the two 16-byte K chunks) and stride (between 8-row groups) byte
// offsets, all in 16-byte units; descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t r2_desc(unsigned saddr, unsigned lbo_bytes) {
    return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) | ((uint64_t)(128u >> 4) << 32) |
           ((uint64_t)1 << 46);
}

// The same descriptor as two words: everything but the start address is loop invariant, and the start address (16-byte
// units, < 2^14 for any shared-memory address) can be advanced by plain 32-bit adds on the low word.
constexpr unsigned R2_DESC_HI = (128u >> 4) | (1u << 14);
__device__ __forceinline__ unsigned r2_desc_lo(unsigned saddr, unsigned lbo_bytes) { return ((saddr >> 4) & 0x3fffu) | ((lbo_bytes >> 4) << 16); }

// D (fp32) = A (tf32, K-major) x B (tf32, K-major), M = 32, N = 256
constexpr unsigned R2_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((256u >> 3) << 17) | ((32u >> 4) << 24);

__device__ __forceinline__ void r2_mma(unsigned d_tmem, uint64_t a_desc, uint64_t b_desc, unsigned accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.ws.cta_group::1.kind::tf32.collector::b0::discard [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(R2_IDESC), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ void r2_mma_lo(unsigned d_tmem, unsigned a_lo, unsigned b_lo, unsigned accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 da, {%1, %5};\n\tmov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.ws.cta_group::1.kind::tf32.collector::b0::discard [%0], da, db, %3, p;\n\t}" ::"r"(d_tmem),
        "r"(a_lo), "r"(b_lo), "r"(R2_IDESC), "r"(accumulate), "r"(R2_DESC_HI)
        : "memory");
}

// the 9 taps x KS k-steps of one 3x3 convolution into one accumulator tile: 2 adds per instruction
template <int KS>
__device__ __forceinline__ void r2_conv_tile(unsigned d_tmem, unsigned a_lo, unsigned b_lo, int PW, unsigned plane2, bool first_accumulates) {
    // a_lo: filters of tap 0; consecutive (tap, k-step) blocks are 2 planes x 32 rows = 64 units apart.  b_lo: raster position
    // of the tile's first output; tap (dy, dx) starts (dy-1) * PW + (dx-1) positions away, k-step ks 2 planes further.
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
        const unsigned bt = b_lo + (unsigned)((tap / 3 - 1) * PW + (tap % 3 - 1));
#pragma unroll
        for (int ks = 0; ks < KS; ++ks)
            r2_mma_lo(d_tmem, a_lo + (unsigned)((tap * KS + ks) * 64), bt + ks * plane2, (first_accumulates || (tap | ks) != 0) ? 1u : 0u);
    }
}

__device__ __forceinline__ void r2_commit(unsigned bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void r2_wait(unsigned bar, unsigned parity) {
    unsigned done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar), "r"(parity)
                     : "memory");
}

// 32 lanes x 32 columns of TMEM: thread = lane, v[j] = column col0 + j
__device__ __forceinline__ void r2_tmem_ld32(unsigned addr, float (&v)[32]) {
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

__device__ __forceinline__ float r2_tf32(float x) {
    unsigned r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

// bits [lo, hi) of a 32-bit word (empty when hi <= lo)
__device__ __forceinline__ unsigned r2_bits(int lo, int hi) {
    lo = max(lo, 0); hi = min(hi, 32);
    if (hi <= lo) return 0u;
    return (hi == 32 ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);
}

// A block of 32 consecutive raster positions q0 .. q0+31 (pitch PW >= 32: at most one row change, at element `split`):
// bit j of the result is set when position q0 + j is a pixel of the image columns [0, W) -- raster columns [2, W+2) --
// in a raster row for which the caller says the image row exists (okA: the row of q0, okB: the next one).
__device__ __forceinline__ unsigned r2_block_mask(int c0, int split, int W, bool okA, bool okB) {
    unsigned m = 0u;
    if (okA) m |= r2_bits(2 - c0, min(W + 2 - c0, split));
    if (okB) m |= r2_bits(split + 2, split + W + 2);
    return m;
}

template <int CIN>
__global__ void __launch_bounds__(R2_THREADS, 1)
res2d_block_kernel(const __grid_constant__ Res2dParams p) {
    constexpr int PLX = CIN / 4, PLM = R2_COUT / 4;
    constexpr bool OVERLAY = CIN == 32;   // the intermediate activation takes the place of the input tile
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // [0..3]: accumulator tile t of conv 1 complete, [4..7]: of conv 2, [8]: the skip conv has read the input raster,
    // [9]: the next input raster has landed (one asynchronous arrival per staging thread)
    __shared__ __align__(8) unsigned long long s_bar[10];
    __shared__ unsigned s_tmem;
    float4 *sX = reinterpret_cast<float4 *>(smem_raw);                  // [PLX][NPXX]
    float4 *sM = OVERLAY ? sX : sX + (size_t)PLX * p.NPXX;               // [PLM][NPXM]
    float4 *sW1 = sM + (size_t)PLM * p.NPXM;                             // [9][PLX][32]
    float4 *sW2 = sW1 + 9 * PLX * R2_COUT;                               // [9][PLM][32]
    float4 *sWs = sW2 + 9 * PLM * R2_COUT;                               // [PLX][32]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool has_skip = p.ws != nullptr;
    const int PW = p.PW, TH = p.TH;

    for (int e = tid; e < 9 * PLX * R2_COUT; e += R2_THREADS) sW1[e] = __ldg(p.w1 + e);
    for (int e = tid; e < 9 * PLM * R2_COUT; e += R2_THREADS) sW2[e] = __ldg(p.w2 + e);
    if (has_skip)
        for (int e = tid; e < PLX * R2_COUT; e += R2_THREADS) sWs[e] = __ldg(p.ws + e);
    if (tid == 0) {
#pragma unroll
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(r2_smem_u32(&s_bar[9])), "r"(R2_LOAD_WARPS * 32));
#pragma unroll
        for (int b = 0; b < 9; ++b) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(r2_smem_u32(&s_bar[b])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(r2_smem_u32(&s_tmem)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tm = s_tmem;
    const unsigned bar0 = r2_smem_u32(&s_bar[0]);
    const unsigned aX = r2_smem_u32(sX), aM = r2_smem_u32(sM), aW1 = r2_smem_u32(sW1), aW2 = r2_smem_u32(sW2), aWs = r2_smem_u32(sWs);
    const float bias1 = __ldg(p.b1 + lane), bias2 = __ldg(p.b2 + lane);   // an epilogue thread is one output channel
    const int quad = warp & 3, half = (warp >> 2) & 1;
    const bool issuer = tid == R2_EPI_WARPS * 32, drains = warp < R2_EPI_WARPS;

    // Stage the input raster of a tile: position q = index - 1 is image pixel (y0 - 2 + q / PW, q % PW - 2), zeros outside.
    // A warp takes whole raster rows (no division); lanes run over (pixel, 16-byte chunk), coalesced in global memory.
    // Rows r0, r0 + rstep, ... of the raster are this warp's: the eight draining warps share a raster (r0 = warp, rstep = 8), or
    // the two loader warps do (PIPE, inside the loop).
    auto fill = [&](long tile, int r0, int rstep) {
        const int n = (int)(tile / p.tiles_y), y0 = (int)(tile - (long)n * p.tiles_y) * TH;
        const float *img = p.x + (size_t)n * p.H * p.W * CIN;
        for (int r = r0; r < TH + 4; r += rstep) {
            const int y = y0 - 2 + r;
            const bool yok = y >= 0 && y < p.H;
            const float *row = img + (size_t)(yok ? y : 0) * p.W * CIN;
            const unsigned dst0 = aX + (unsigned)((r * PW + 1) * 16);
#pragma unroll 3
            for (int e = lane; e < PW * PLX; e += 32) {   // unrolled: a cp.async holds its address registers until it is read
                const int c = e / PLX, k = e - c * PLX, xx = c - 2;   // PLX is 4 or 8
                const bool ok = yok && xx >= 0 && xx < p.W;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst0 + (unsigned)((k * p.NPXX + c) * 16)),
                             "l"(row + (size_t)(ok ? xx : 0) * CIN + k * 4), "r"(ok ? 16 : 0));
            }
        }
        if (tid < 2 * PLX) {   // the entry before the raster and the one after it (never written by the copies: once is enough
            const int k = tid >> 1, idx = (tid & 1) ? (TH + 4) * PW + 1 : 0;   // where the raster is not overlaid)
            sX[(size_t)k * p.NPXX + idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    const bool loader = warp > R2_EPI_WARPS;

    // issued by one thread: the 1x1 skip conv into the second accumulators, and the first convolution (one commit per
    // accumulator tile)
    auto issue_skip = [&]() {
        const unsigned bx = r2_desc_lo(aX, p.NPXX * 16), plane2 = 2u * p.NPXX, as = r2_desc_lo(aWs, R2_COUT * 16);
        for (int t = 0; t < p.nt2; ++t)
#pragma unroll
            for (int ks = 0; ks < CIN / 8; ++ks)
                r2_mma_lo(tm + R2_ACC2_COL + t * 64, as + ks * 64, bx + (unsigned)(2 * PW + 256 * t + 1) + ks * plane2, ks != 0);
    };
    auto issue_conv1 = [&]() {
        const unsigned bx = r2_desc_lo(aX, p.NPXX * 16), plane2 = 2u * p.NPXX, a1 = r2_desc_lo(aW1, R2_COUT * 16);
        for (int t = 0; t < p.nt1; ++t) {
            r2_conv_tile<CIN / 8>(tm + t * 64, a1, bx + (unsigned)(PW + 256 * t + 1), PW, plane2, false);
            r2_commit(bar0 + 8 * t);
        }
    };
    // PIPE (the intermediate activation has its own buffer, the skip is a 1x1 conv): the first convolution of tile i + 1 is
    // issued behind the second one of tile i, so the tensor core works through epilogue 2 instead of waiting for it; the
    // skip conv, which writes the accumulators epilogue 2 is reading, moves in front of the second convolution of its own tile.
    constexpr bool PIPE = !OVERLAY;

    long long pt[7] = {0, 0, 0, 0, 0, 0, 0};
    unsigned parity = 0;
    if ((long)blockIdx.x < p.ntiles && drains) fill(blockIdx.x, warp, R2_EPI_WARPS);
    if (PIPE) {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (issuer && (long)blockIdx.x < p.ntiles) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            issue_conv1();
        }
    }
    for (long tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, parity ^= 1u) {
        const int n = (int)(tile / p.tiles_y), y0 = (int)(tile - (long)n * p.tiles_y) * TH;
        const long next = tile + gridDim.x;
        if (p.prof && tid == 0) pt[0] = clock64();
        if (!PIPE) {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the tensor core
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();   // the raster is staged; the previous tile's accumulators have been read
        }
        if (p.prof && tid == 0) pt[1] = clock64();

        if (!PIPE && issuer) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (has_skip) issue_skip();
            issue_conv1();
        }
        if (p.prof && tid == 0) pt[2] = clock64();

        // ---- epilogue 1: + bias, ReLU, zero outside the image, TF32 -> the raster of the second convolution.  With the
        // intermediate activation in its own buffer a tile is drained while the tensor core works on the next one.
        if (drains) {
            if (OVERLAY) r2_wait(bar0 + 8 * (p.nt1 - 1), parity);
            float *mid = reinterpret_cast<float *>(sM) + (size_t)(lane >> 2) * p.NPXM * 4 + (lane & 3);
            const int qend = (TH + 3) * PW;
            for (int t = 0; t < p.nt1; ++t) {
                if (!OVERLAY) r2_wait(bar0 + 8 * t, parity);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (p.prof && tid == 0 && t == 0) pt[3] = clock64();
                const int q0 = PW + 256 * t + 64 * quad + 32 * half;
                if (q0 >= qend) continue;   // warp uniform
                float v[32];
                r2_tmem_ld32(tm + ((unsigned)(quad * 32) << 16) + t * 64 + half * 32, v);
                const int r0 = q0 / PW, c0 = q0 - r0 * PW, split = PW - c0;
                const unsigned m = r2_block_mask(c0, split, p.W, (unsigned)(y0 - 2 + r0) < (unsigned)p.H,
                                                 (unsigned)(y0 - 1 + r0) < (unsigned)p.H);
                float *dst = mid + (size_t)(q0 + 1) * 4;
                const int lim = qend - q0;
                if (lim >= 32) {   // all 32 positions belong to the raster: store, then clear the few that lie outside the image
#pragma unroll
                    for (int j = 0; j < 32; ++j) dst[j * 4] = r2_tf32(fmaxf(v[j] + bias1, 0.0f));
                    for (unsigned z = ~m; z != 0u; z &= z - 1u) dst[(__ffs(z) - 1) * 4] = 0.0f;
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < lim) dst[j * 4] = (m >> j) & 1u ? r2_tf32(fmaxf(v[j] + bias1, 0.0f)) : 0.0f;
                }
            }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (p.prof && tid == 0) pt[4] = clock64();

        // ---- second convolution
        if (issuer) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (PIPE) {   // the second accumulators are free now: epilogue 2 of the previous tile ended before the barrier above
                issue_skip();
                r2_commit(bar0 + 8 * 8);
            }
            const unsigned bm = r2_desc_lo(aM, p.NPXM * 16), a2 = r2_desc_lo(aW2, R2_COUT * 16);
            for (int t = 0; t < p.nt2; ++t) {
                r2_conv_tile<R2_COUT / 8>(tm + R2_ACC2_COL + t * 64, a2, bm + (unsigned)(2 * PW + 256 * t + 1), PW, 2u * p.NPXM, has_skip);
                r2_commit(bar0 + 8 * (4 + t));
            }
        }
        if (PIPE && next < p.ntiles) {
            // the next tile's raster, as soon as the skip conv has read this one (the first convolution has: epilogue 1 waited
            // for it): two loader warps stage it, their copies arrive on a barrier of their own, and the issuing thread queues the
            // first convolution of the next tile behind the second one of this tile once they have landed.  The draining warps
            // go straight on to epilogue 2.
            if (loader) {
                r2_wait(bar0 + 8 * 8, parity);
                fill(next, warp - (R2_EPI_WARPS + 1), R2_LOAD_WARPS);
                asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar0 + 8 * 9) : "memory");
            } else if (issuer) {
                r2_wait(bar0 + 8 * 9, parity);
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the copies -> visible to the tensor core
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                issue_conv1();
            }
        }
        if (p.prof && tid == 0) pt[5] = clock64();

        // ---- epilogue 2: + bias (+ identity skip), ReLU -> global memory, one 128-byte pixel record per warp store;
        // a tile is drained while the tensor core works on the next one
        if (drains) {
            float *out_img = p.out + (size_t)n * p.H * p.W * R2_COUT + lane;
            const float *idn_img = p.x + (size_t)n * p.H * p.W * CIN + lane;   // identity skip: CIN == 32
            for (int t = 0; t < p.nt2; ++t) {
                r2_wait(bar0 + 8 * (4 + t), parity);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (p.prof && tid == 0 && t == 0) pt[6] = clock64();
                if (OVERLAY && t == p.nt2 - 1 && next < p.ntiles) fill(next, warp, R2_EPI_WARPS);   // the raster has been read for the last time
                const int q0 = 2 * PW + 256 * t + 64 * quad + 32 * half;
                const int r0 = q0 / PW, c0 = q0 - r0 * PW, split = PW - c0;
                const unsigned m = r2_block_mask(c0, split, p.W, r0 < TH + 2 && y0 + r0 - 2 < p.H, r0 + 1 < TH + 2 && y0 + r0 - 1 < p.H);
                if (m == 0u) continue;   // warp uniform
                float v[32];
                r2_tmem_ld32(tm + ((unsigned)(quad * 32) << 16) + R2_ACC2_COL + t * 64 + half * 32, v);
                // pixel index of element j: pa + j in the row of q0, pa - 4 + j in the next one (the raster is 4 wider)
                const long pa = (long)(y0 + r0 - 2) * p.W + (c0 - 2);
                const unsigned ma = m & r2_bits(0, split), mb = m & ~r2_bits(0, split);
                if (!has_skip) {
                    const float *ia = idn_img + pa * R2_COUT, *ib = ia - 4 * R2_COUT;
                    float w[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        w[j] = 0.0f;
                        if ((ma >> j) & 1u) w[j] = __ldg(ia + j * R2_COUT);
                        if ((mb >> j) & 1u) w[j] = __ldg(ib + j * R2_COUT);
                    }
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] += w[j];
                }
                float *oa = out_img + pa * R2_COUT, *ob = oa - 4 * R2_COUT;
                if (ma == 0xffffffffu) {   // one image row, every position a pixel (about half of the blocks)
#pragma unroll
                    for (int j = 0; j < 32; ++j) oa[j * R2_COUT] = fmaxf(v[j] + bias2, 0.0f);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float val = fmaxf(v[j] + bias2, 0.0f);
                        if ((ma >> j) & 1u) oa[j * R2_COUT] = val;
                        if ((mb >> j) & 1u) ob[j * R2_COUT] = val;
                    }
                }
            }
        }
        if (p.prof && tid == 0 && blockIdx.x == 0) {
            // cycles of CTA 0 per phase: wait for the staged tile, issue 1, until the first tile of conv 1 is complete,
            // epilogue 1 (+ staging the next tile), issue 2, until the first tile of conv 2 is complete, epilogue 2; tiles
            const long long now = clock64();
            p.prof[0] += pt[1] - pt[0]; p.prof[1] += pt[2] - pt[1]; p.prof[2] += pt[3] - pt[2]; p.prof[3] += pt[4] - pt[3];
            p.prof[4] += pt[5] - pt[4]; p.prof[5] += pt[6] - pt[5]; p.prof[6] += now - pt[6]; p.prof[7] += 1;
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm));
}

}  // namespace fvp

using namespace fvp;

// Tile height and buffer sizes for an image width; false when no tile fits (the caller stays on its library path)
static bool res2d_plan(int cin, int W, int H, Res2dParams &p, size_t &smem) {
    const int PW = W + 4, plx = cin / 4;
    if (PW < 32) return false;   // an epilogue block of 32 positions changes raster row at most once
    for (int TH : {8, 6, 4, 2}) {
        if (TH > 8 || (TH > H && TH != 2)) continue;
        const int nt1 = ceil_div((TH + 2) * PW, 256), nt2 = ceil_div(TH * PW, 256);
        if (nt1 > 4 || nt2 > 4) continue;                       // 4 + 4 accumulator tiles of 64 TMEM columns
        auto odd8 = [](int v) { return ((v + 6) / 8) * 8 + 1; };   // == 1 (mod 8): the planes start 4 banks apart
        const int npxx = odd8((TH + 4) * PW + 2), npxm = cin == 32 ? npxx : odd8((TH + 3) * PW + 2);
        const size_t act = (cin == 32 ? 8 * (size_t)npxx : (size_t)plx * npxx + 8 * (size_t)npxm) * 16;
        const size_t wts = (size_t)(9 * plx + 9 * 8 + plx) * R2_COUT * 16;
        // the last accumulator tile reads up to 3 * PW + 256 * nt + 2 positions of the last plane: must stay inside the allocation
        const size_t over = (size_t)(3 * PW + 256 * std::max(nt1, nt2) + 8) * 16;
        const size_t total = act + std::max(wts, over);
        if (total > 225 * 1024) continue;
        p.TH = TH; p.PW = PW; p.NPXX = npxx; p.NPXM = npxm; p.nt1 = nt1; p.nt2 = nt2;
        smem = total;
        return true;
    }
    return false;
}

static long long *g_res2d_prof = nullptr;

extern "C" int fvp_res2d_profile(long long *counters) {
    g_res2d_prof = counters;
    return FVP_OK;
}

extern "C" int fvp_res2d_supported(int cin, int H, int W) {
    Res2dParams p;
    size_t smem;
    return (cin == 16 || cin == 32) && H >= 1 && W >= 28 && res2d_plan(cin, W, H, p, smem) ? 1 : 0;
}

extern "C" int fvp_res2d_block_tf32(const float *x, int N, int H, int W, int cin, const float *w1, const float *b1,
                                    const float *w2, const float *b2, const float *wskip, float *out, void *stream) {
    FVP_REQUIRE(x && w1 && b1 && w2 && b2 && out, "fvp_res2d_block_tf32: null pointer");
    FVP_REQUIRE(cin == 16 || cin == 32, "fvp_res2d_block_tf32: 16 or 32 input channels (got %d); output channels are 32", cin);
    FVP_REQUIRE(cin == 32 || wskip != nullptr, "fvp_res2d_block_tf32: an identity skip needs 32 input channels");
    FVP_REQUIRE(N >= 0 && H >= 1 && W >= 28, "fvp_res2d_block_tf32: bad shape N=%d H=%d W=%d (W >= 28)", N, H, W);
    FVP_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)w1 & 15) == 0 && ((uintptr_t)w2 & 15) == 0 && ((uintptr_t)wskip & 15) == 0 &&
                    ((uintptr_t)out & 15) == 0,
                "fvp_res2d_block_tf32: pointers must be 16-byte aligned");
    if (N == 0) return FVP_OK;
    Res2dParams p;
    size_t smem = 0;
    FVP_REQUIRE(res2d_plan(cin, W, H, p, smem), "fvp_res2d_block_tf32: image width %d does not fit a tile (fvp_res2d_supported)", W);
    p.x = x; p.out = out;
    p.w1 = reinterpret_cast<const float4 *>(w1); p.w2 = reinterpret_cast<const float4 *>(w2);
    p.ws = reinterpret_cast<const float4 *>(wskip);
    p.b1 = b1; p.b2 = b2;
    p.prof = g_res2d_prof;
    p.N = N; p.H = H; p.W = W;
    p.tiles_y = ceil_div(H, p.TH);
    p.ntiles = (long)N * p.tiles_y;
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = (int)std::min<long>(p.ntiles, (long)sms);
    if (cin == 16) {
        FVP_CUDA(cudaFuncSetAttribute(res2d_block_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        res2d_block_kernel<16><<<grid, R2_THREADS, smem, (cudaStream_t)stream>>>(p);
    } else {
        FVP_CUDA(cudaFuncSetAttribute(res2d_block_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        res2d_block_kernel<32><<<grid, R2_THREADS, smem, (cudaStream_t)stream>>>(p);
    }
    return check_launch("res2d_block_kernel");
}
