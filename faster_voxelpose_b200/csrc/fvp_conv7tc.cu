// Front layer of P2PNet / CenterNet (reference lib/models/cnns_2d.py:190-192: conv 7x7, 16 output channels, BatchNorm
// folded, ReLU) on the tcgen05 tensor cores.  The mma.sync kernel of fvp_conv7.cu ends up bound by the legacy tensor pipe
// (~12 cycles per HMMA.1688.TF32 per SM sub-partition, 0.52 ms for the 960 planes of the bench); this one uses the raster /
// shifted-descriptor formulation of fvp_res2d.cu (a tap = a shifted start address over one staged pixel raster, filters as
// the A operand of tcgen05.mma.ws M = 32, 256 raster positions as B) -- with one twist, because 16 filters would leave half
// of the 32 accumulator rows idle:
//   * TAP PAIRS.  Rows 0-15 of A hold the filters of tap (ky, kx), rows 16-31 those of tap (ky, kx+1), for kx = 0, 2, 4 (and
//     zeros beside kx = 6).  Both halves see the same pixels, B = raster shifted for (ky, kx); the upper half therefore
//     accumulates, at position q, what tap (ky, kx+1) contributes to the output at q - 1.  All pairs have the same offset, so
//     after 28 x Cin/8 instructions (instead of 49 x Cin/8):   out[q] = lo[q] + hi[q + 1].
//   * The epilogue forms that sum from two TMEM rows: lane L < 16 holds lo (filter L), lane L + 16 holds hi, a shuffle brings
//     hi[q + 1] over; a thread walks the 64 positions of its lane quadrant, and the one position per quadrant whose right
//     neighbour lives in the next quadrant (another warp) gets it through a 16-float exchange slot in shared memory.
//   * Roles around mbarriers: warps 0-7 drain (two accumulator sets: the tensor core works on tile i+1 meanwhile), warp 8
//     issues the MMAs from one lane, warps 9-10 stage the next raster (cp.async, two buffers).  The filters (28 pair-taps x
//     Cin/4 planes x 32 rows x 16 B) stay resident.
#include <algorithm>
#include <cstdint>

#include "fvp_common.cuh"

namespace fvp {

constexpr int T7_DRAIN_WARPS = 8, T7_XLOAD_WARPS = 2, T7_THREADS = 32 * (T7_DRAIN_WARPS + 1 + T7_XLOAD_WARPS);
constexpr int T7_PAD = 4;          // raster entries in front of position 0 (taps reach 3 positions back)
constexpr int T7_MAX_NT = 4;       // accumulator tiles (256 positions, 64 TMEM columns) per raster tile

struct Conv7tcParams {
    const float *x;        // [N][H][W][CIN]
    float *out;            // [N][H][W][16]
    const float4 *w;       // [28][CIN/4][32] float4: pair-taps, TF32-rounded; CIN == 20: followed by 32 float4 of zeros
    const float *bias;     // [16]
    int N, H, W, TH, PW, NPX, nt, tiles_y, relu;
    long ntiles;
};

__device__ __forceinline__ unsigned t7_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void t7_wait(unsigned bar, unsigned parity) {
    unsigned done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar), "r"(parity)
                     : "memory");
}
__device__ __forceinline__ void t7_arrive(unsigned bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void t7_commit(unsigned bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

constexpr unsigned T7_DESC_HI = (128u >> 4) | (1u << 14);   // stride byte offset 128, descriptor version 1
constexpr unsigned T7_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((256u >> 3) << 17) | ((32u >> 4) << 24);   // f32 += tf32 x tf32, M 32, N 256

__device__ __forceinline__ void t7_mma(unsigned d_tmem, unsigned a_lo, unsigned b_lo, unsigned accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 da, {%1, %5};\n\tmov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.ws.cta_group::1.kind::tf32.collector::b0::discard [%0], da, db, %3, p;\n\t}" ::"r"(d_tmem),
        "r"(a_lo), "r"(b_lo), "r"(T7_IDESC), "r"(accumulate), "r"(T7_DESC_HI)
        : "memory");
}

__device__ __forceinline__ void t7_tmem_ld32(unsigned addr, float (&v)[32]) {
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

__device__ __forceinline__ float t7_tmem_ld1(unsigned addr) {
    unsigned u;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(u) : "r"(addr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    return __uint_as_float(u);
}

// bits [lo, hi) of a 32-bit word (empty when hi <= lo)
__device__ __forceinline__ unsigned t7_bits(int lo, int hi) {
    lo = max(lo, 0); hi = min(hi, 32);
    if (hi <= lo) return 0u;
    return (hi == 32 ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);
}

template <int CIN>
__global__ void __launch_bounds__(T7_THREADS, 1)
conv7x7_tc_kernel(const __grid_constant__ Conv7tcParams p) {
    constexpr int KS = (CIN + 7) / 8;          // k-steps per pair-tap
    constexpr int PLX = CIN / 4;               // raster planes that exist; CIN == 20: the "sixth" plane of the last k-step is whatever
                                               // follows in shared memory (finite: initialised below) times a plane of zero filters
    constexpr int PLW = CIN / 4;               // filter planes stored per pair-tap; CIN == 20: one shared plane of zeros follows the filters
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // [0,1] raster staged, [2,3] accumulator set complete (= raster free), [4,5] accumulator set drained
    __shared__ __align__(8) unsigned long long s_bar[6];
    __shared__ unsigned s_tmem;
    __shared__ float s_xchg[2][T7_MAX_NT][4][16];
    float4 *sX = reinterpret_cast<float4 *>(smem_raw);            // [2][PLX][NPX]
    float4 *sW = sX + (size_t)2 * PLX * p.NPX;                    // [28][PLW][32]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int PW = p.PW, TH = p.TH, NT = p.nt;
    const unsigned bar0 = t7_u32(&s_bar[0]);
    auto BAR = [&](int i) { return bar0 + 8u * (unsigned)i; };

    // everything the tensor core may read is finite from the start (over-reads of the rasters are multiplied by zeros)
    for (int e = tid; e < 2 * PLX * p.NPX; e += T7_THREADS) sX[e] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int e = tid; e < 28 * PLW * 32 + (CIN == 20 ? 32 : 0); e += T7_THREADS) sW[e] = __ldg(p.w + e);
    if (tid == 0) {
        for (int b = 0; b < 2; ++b) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(BAR(0 + b)), "r"(T7_XLOAD_WARPS * 32));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(BAR(2 + b)));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(BAR(4 + b)), "r"(T7_DRAIN_WARPS * 32));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(t7_u32(&s_tmem)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tm = s_tmem;
    const unsigned aX = t7_u32(sX), aW = t7_u32(sW);
    const unsigned xbuf = (unsigned)(PLX * p.NPX * 16);

    if (warp < T7_DRAIN_WARPS) {
        // ------------------------------------------------------------------------------------------------ drain
        const int quad = warp & 3, set = warp >> 2;       // TMEM lanes 32 * quad ..; accumulator tiles t with t % 2 == set
        const int cout = lane & 15;
        const float bias = __ldg(p.bias + cout);
        long it = 0;
        for (long tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
            const int b = (int)(it & 1);
            const int n = (int)(tile / p.tiles_y), y0 = (int)(tile - (long)n * p.tiles_y) * TH;
            t7_wait(BAR(2 + b), (unsigned)((it >> 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const unsigned acc = tm + ((unsigned)(quad * 32) << 16) + (unsigned)(b * 64 * T7_MAX_NT);
            // pass 1: the first position of every quadrant, upper rows, for the quadrant before it
            for (int t = set; t < NT; t += 2) {
                const float first = t7_tmem_ld1(acc + t * 64);
                if (lane >= 16) s_xchg[b][t][quad][lane - 16] = first;
            }
            asm volatile("bar.sync 1, %0;" ::"n"(T7_DRAIN_WARPS * 32) : "memory");
            // pass 2
            float *out_img = p.out + (size_t)n * p.H * p.W * 16 + cout;
            for (int t = set; t < NT; t += 2) {
                float nextq = 0.0f;   // hi[first position of the next quadrant]
                if (quad < 3) nextq = s_xchg[b][t][quad + 1][cout];
                else if (t + 1 < NT) nextq = s_xchg[b][t + 1][0][cout];
#pragma unroll 1
                for (int h = 0; h < 2; ++h) {
                    float v[32];
                    t7_tmem_ld32(acc + t * 64 + h * 32, v);
                    // hi[q + 1] of the last position of this half: first position of the other half / of the next quadrant
                    float edge = nextq;
                    if (h == 0) edge = __shfl_down_sync(0xffffffffu, t7_tmem_ld1(acc + t * 64 + 32), 16);
                    // 32 consecutive raster positions from q0: at most one row change (PW >= 64), at element `split`.  Bit j of
                    // ma / mb: element j is a pixel of the image in the row of q0 / in the next row (raster columns [3, W+3),
                    // rows [3, TH+3) of the tile that exist in the image); only lanes 0-15 (the lo rows) store.
                    const int q0 = 3 * PW + 256 * t + 64 * quad + 32 * h;
                    const int r0 = q0 / PW, c0 = q0 - r0 * PW, split = PW - c0;
                    unsigned ma = 0u, mb = 0u;
                    if (lane < 16) {
                        if (r0 < TH + 3 && y0 + r0 - 3 < p.H) ma = t7_bits(3 - c0, min(p.W + 3 - c0, split));
                        if (r0 + 1 < TH + 3 && y0 + r0 - 2 < p.H) mb = t7_bits(split + 3, split + p.W + 3);
                    }
                    // pixel index of element j: pa + j in the row of q0, pa - 6 + j in the next one (the raster is 6 wider)
                    float *oa = out_img + ((long)(y0 + r0 - 3) * p.W + (c0 - 3)) * 16, *ob = oa - 6 * 16;
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float hi = j < 31 ? __shfl_down_sync(0xffffffffu, v[j < 31 ? j + 1 : 31], 16) : edge;
                        float val = v[j] + hi + bias;
                        if (p.relu) val = fmaxf(val, 0.0f);
                        FVP_ASSERT(!((ma >> j) & 1u) || (oa + j * 16 >= out_img && oa + j * 16 < out_img + (size_t)p.H * p.W * 16));
                        FVP_ASSERT(!((mb >> j) & 1u) || (ob + j * 16 >= out_img && ob + j * 16 < out_img + (size_t)p.H * p.W * 16));
                        if ((ma >> j) & 1u) oa[j * 16] = val;
                        if ((mb >> j) & 1u) ob[j * 16] = val;
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            t7_arrive(BAR(4 + b));
        }
    } else if (warp == T7_DRAIN_WARPS) {
        // ------------------------------------------------------------------------------------------------ MMA issue
        if (lane == 0) {
            long it = 0;
            const unsigned zero_plane = ((aW >> 4) & 0x3fffu) + (unsigned)(28 * PLW * 32);   // CIN == 20: 32 rows of zeros
            for (long tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
                const int b = (int)(it & 1);
                t7_wait(BAR(0 + b), (unsigned)((it >> 1) & 1));         // raster staged
                t7_wait(BAR(4 + b), (unsigned)(((it >> 1) & 1) ^ 1));   // accumulator set drained (free at first)
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const unsigned bx = (((aX + b * xbuf) >> 4) & 0x3fffu) | ((unsigned)p.NPX << 16);   // LBO = NPX * 16 bytes
                for (int t = 0; t < NT; ++t) {
                    const unsigned d = tm + (unsigned)(b * 64 * T7_MAX_NT + t * 64);
#pragma unroll 1
                    for (int ky = 0; ky < 7; ++ky) {
#pragma unroll
                        for (int pr = 0; pr < 4; ++pr) {
                            const unsigned aw = ((aW >> 4) & 0x3fffu) + (unsigned)(((ky * 4 + pr) * PLW) * 32);
                            const unsigned bt = bx + (unsigned)(3 * PW + 256 * t + (ky - 3) * PW + (2 * pr - 3) + T7_PAD);
#pragma unroll
                            for (int ks = 0; ks < KS; ++ks) {
                                // leading byte offset of A: 32 rows to the next plane -- or, for the half-filled last k-step of
                                // 20 channels, the distance to the plane of zeros behind the filters
                                const unsigned a_start = aw + (unsigned)(ks * 64);
                                const unsigned lbo = (CIN == 20 && ks == KS - 1) ? zero_plane - a_start : 32u;
                                t7_mma(d, a_start | (lbo << 16), bt + (unsigned)(ks * 2 * p.NPX), (ky | pr | ks) != 0);
                            }
                        }
                    }
                }
                t7_commit(BAR(2 + b));           // raster free, accumulator set complete
            }
        }
    } else {
        // ------------------------------------------------------------------------------------------------ raster loader
        const int lw = warp - (T7_DRAIN_WARPS + 1);
        long it = 0;
        for (long tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
            const int b = (int)(it & 1);
            const int n = (int)(tile / p.tiles_y), y0 = (int)(tile - (long)n * p.tiles_y) * TH;
            t7_wait(BAR(2 + b), (unsigned)(((it >> 1) & 1) ^ 1));       // raster buffer free (free at first)
            const float *img = p.x + (size_t)n * p.H * p.W * CIN;
            const unsigned base = aX + b * xbuf;
            for (int r = lw; r < TH + 6; r += T7_XLOAD_WARPS) {
                const int y = y0 - 3 + r;
                const bool yok = y >= 0 && y < p.H;
                const float *row = img + (size_t)(yok ? y : 0) * p.W * CIN;
                const unsigned dst0 = base + (unsigned)((r * PW + T7_PAD) * 16);
#pragma unroll 4
                for (int e = lane; e < PW * PLX; e += 32) {
                    const int c = e / PLX, k = e - c * PLX, xx = c - 3;
                    const bool ok = yok && xx >= 0 && xx < p.W;
                    FVP_ASSERT(r * PW + T7_PAD + c < p.NPX && k < PLX);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst0 + (unsigned)((k * p.NPX + c) * 16)),
                                 "l"(row + (size_t)(ok ? xx : 0) * CIN + k * 4), "r"(ok ? 16 : 0));
                }
            }
            asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the tensor core
            t7_arrive(BAR(0 + b));
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm));
}

}  // namespace fvp

using namespace fvp;

static bool conv7tc_plan(int cin, int H, int W, Conv7tcParams &p, size_t &smem) {
    const int PW = W + 6, plx = cin / 4, plw = cin / 4;
    if (PW < 64 || H < 1) return false;   // a quadrant of 64 positions changes raster row at most once
    auto odd8 = [](int v) { return ((v + 6) / 8) * 8 + 1; };
    const size_t wts = (size_t)28 * plw * 32 * 16 + (cin == 20 ? 512 : 0);
    int best = 0;
    double best_eff = 0.0;
    for (int TH = 1; TH <= std::min(H, 32); ++TH) {
        const int nt = (TH * PW + 255) / 256;
        if (nt > T7_MAX_NT) break;
        const int npx = odd8((TH + 6) * PW + 2 * T7_PAD);
        if ((size_t)2 * plx * npx * 16 + wts > 223 * 1024) break;
        const int tiles_y = (H + TH - 1) / TH;
        const double eff = (double)H * PW / ((double)tiles_y * nt * 256);   // useful positions per issued position
        if (eff > best_eff + 1e-9) { best_eff = eff; best = TH; }
    }
    if (best == 0) return false;
    p.TH = best; p.PW = PW; p.nt = (best * PW + 255) / 256; p.NPX = odd8((best + 6) * PW + 2 * T7_PAD);
    smem = (size_t)2 * plx * p.NPX * 16 + wts;
    return true;
}

extern "C" int fvp_conv7x7_tc_supported(int cin, int H, int W) {
    Conv7tcParams p;
    size_t smem;
    return (cin == 16 || cin == 20) && conv7tc_plan(cin, H, W, p, smem) ? 1 : 0;
}

extern "C" int fvp_conv7x7_tc_tf32(const float *in, const float *wpairs, const float *bias, int N, int H, int W, int cin, int relu,
                                   float *out, void *stream) {
    FVP_REQUIRE(in && wpairs && bias && out, "fvp_conv7x7_tc_tf32: null pointer");
    FVP_REQUIRE(cin == 16 || cin == 20, "fvp_conv7x7_tc_tf32: 16 or 20 input channels (got %d)", cin);
    FVP_REQUIRE(N >= 0 && H >= 1 && W >= 1, "fvp_conv7x7_tc_tf32: bad shape N=%d H=%d W=%d", N, H, W);
    FVP_REQUIRE(((uintptr_t)in & 15) == 0 && ((uintptr_t)wpairs & 15) == 0 && ((uintptr_t)out & 3) == 0,
                "fvp_conv7x7_tc_tf32: misaligned pointer");
    if (N == 0) return FVP_OK;
    Conv7tcParams p;
    size_t smem = 0;
    FVP_REQUIRE(conv7tc_plan(cin, H, W, p, smem), "fvp_conv7x7_tc_tf32: image width %d does not fit a tile (fvp_conv7x7_tc_supported)", W);
    p.x = in; p.out = out; p.w = reinterpret_cast<const float4 *>(wpairs); p.bias = bias;
    p.N = N; p.H = H; p.W = W; p.relu = relu ? 1 : 0;
    p.tiles_y = ceil_div(H, p.TH);
    p.ntiles = (long)N * p.tiles_y;
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = (int)std::min<long>(p.ntiles, (long)sms);
    if (cin == 16) {
        FVP_CUDA(cudaFuncSetAttribute(conv7x7_tc_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        conv7x7_tc_kernel<16><<<grid, T7_THREADS, smem, (cudaStream_t)stream>>>(p);
    } else {
        FVP_CUDA(cudaFuncSetAttribute(conv7x7_tc_kernel<20>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        conv7x7_tc_kernel<20><<<grid, T7_THREADS, smem, (cudaStream_t)stream>>>(p);
    }
    return check_launch("conv7x7_tc_kernel");
}
