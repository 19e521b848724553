// Front layer of P2PNet / CenterNet (reference lib/models/cnns_2d.py:190-192,
// Basic2DBlock(cin, 16, 7): conv 7x7 + BatchNorm + ReLU) as a TF32 tensor-core implicit GEMM.
// On the 960 x 64 x 64 x 16 planes of the Panoptic step cuDNN's best engine for this shape (16 -> 16
// channels, 7x7) is a 1.19 ms kernel plus a 0.13 ms layout conversion; the arithmetic is 98.6 GFLOP and
// the traffic 0.5 GB.  This kernel is used when torch allows TF32 convolutions (its default, and what the
// reference runs with); with TF32 disabled the caller stays on cuDNN's fp32 path.
//
// Layout: activations channels-last [N][H][W][16] fp32, BatchNorm folded into the filters by the caller.
// CTA = 16 warps (one per SM: 173 KB of shared memory), persistent over tiles of 16 rows x 64 columns; a warp owns
// four rows x 16 columns of the tile (one m16 pixel group per row), both n8 halves of the 16 output channels -> 32
// accumulators per thread; an A fragment loaded from a tile row is used for all the output rows it feeds.  A tile is
// loaded in a few % of the time its 49 taps take, so one buffer suffices.
// Shared memory: the input tile with its 3-pixel halo, 20 floats per pixel (80-byte pitch: the eight pixel
// rows of an ldmatrix 8x(4 x tf32) fragment fall in distinct banks), rounded to TF32 once after the
// cp.async fill; and the filters pre-arranged by the host as mma B fragments [tap][k-step][lane][4].
// Inner loop per (kx, 8-channel k-step): seven LDS.128 (the B fragments of the seven ky), ten ldmatrix.x4 (A, one per tile
// row of the group), 56 mma.m16n8k8.tf32.  (First form: warp = one output row, 28 ldmatrix for the same 56 MMAs; ncu showed
// the shared-memory pipe at 86 % -- 0.63 ms for the 960 planes of the bench.)
#include <algorithm>
#include <cstdint>

#include "fvp_common.cuh"

namespace fvp {

constexpr int C7_TH = 16, C7_TW = 64, C7_HALO = 3, C7_PITCH = 20;
constexpr int C7_THREADS = 32 * C7_TH;
constexpr int C7_TROWS = C7_TH + 2 * C7_HALO, C7_TCOLS = C7_TW + 2 * C7_HALO;  // 22 x 70 pixels
constexpr int C7_WFRAG = 49 * 2 * 32 * 4;                                      // floats: channels 0..15, two k8 steps per tap
constexpr int C7_WFRAG4 = 49 * 32 * 2;                                         // floats: channels 16..19 (CIN == 20), one k4 step per tap
constexpr int C7_TILE = C7_TROWS * C7_TCOLS * C7_PITCH;                        // floats
constexpr size_t C7_SMEM = sizeof(float) * (C7_WFRAG + C7_TILE);
constexpr size_t C7_SMEM20 = sizeof(float) * (C7_WFRAG + C7_WFRAG4 + C7_TILE);

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ float to_tf32(float x) {
    unsigned r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

// CIN == 20 (17-joint rigs: the planes arrive padded to 20 channels): the same kernel with a third, k4 step per tap for
// channels 16..19 (mma.m16n8k4, ldmatrix.x2); the 80-byte pixel pitch of the tile already has room for them.
template <int CIN>
__global__ void __launch_bounds__(C7_THREADS, 1)
conv7x7_c16_tf32_kernel(const float *__restrict__ in, const float4 *__restrict__ wfrag, const float *__restrict__ bias,
                        float *__restrict__ out, int N, int H, int W, int relu) {
    static_assert(CIN == 16 || CIN == 20, "16 or 20 input channels");
    constexpr int CHUNKS = CIN / 4;
    extern __shared__ __align__(16) float smem[];
    float4 *s_w = reinterpret_cast<float4 *>(smem);  // [49][2][32]
    float2 *s_w4 = reinterpret_cast<float2 *>(smem + C7_WFRAG);                  // [49][32] (CIN == 20)
    float *s_t = smem + C7_WFRAG + (CIN == 20 ? C7_WFRAG4 : 0);                  // [22][70][20]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;

    for (int e = tid; e < (C7_WFRAG + (CIN == 20 ? C7_WFRAG4 : 0)) / 4; e += C7_THREADS) s_w[e] = __ldg(wfrag + e);
    float bz[2][2];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
        bz[nt][0] = __ldg(bias + nt * 8 + 2 * t);
        bz[nt][1] = __ldg(bias + nt * 8 + 2 * t + 1);
    }

    const int tiles_x = (W + C7_TW - 1) / C7_TW, tiles_y = (H + C7_TH - 1) / C7_TH;
    const long ntiles = (long)N * tiles_y * tiles_x;
    // ldmatrix: lane supplies the row address of matrix (lane >> 3), row (lane & 7):
    //   matrices 0/1 = pixels 0-7 / 8-15 at channels +0..3, matrices 2/3 = the same pixels at channels +4..7
    const int lm_pix = (lane & 7) + ((lane >> 3) & 1) * 8, lm_ch = (lane >> 4) * 4;

    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int tx = (int)(tile % tiles_x);
        const long r = tile / tiles_x;
        const int ty = (int)(r % tiles_y), n = (int)(r / tiles_y);
        const int y0 = ty * C7_TH, x0 = tx * C7_TW;
        __syncthreads();  // previous tile fully consumed (and the filters are in place on the first pass)
        // fill: 16-byte chunks, zero outside the image (cp.async src-size 0)
        for (int e = tid; e < C7_TROWS * C7_TCOLS * CHUNKS; e += C7_THREADS) {
            const int pix = e / CHUNKS, chunk = e - pix * CHUNKS, tr = pix / C7_TCOLS, tc = pix - tr * C7_TCOLS;
            const int y = y0 + tr - C7_HALO, x = x0 + tc - C7_HALO;
            const bool ok = y >= 0 && y < H && x >= 0 && x < W;
            const float *src = in + (((size_t)n * H + (ok ? y : 0)) * W + (ok ? x : 0)) * CIN + chunk * 4;
            const unsigned dst = smem_u32(s_t + pix * C7_PITCH + chunk * 4);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(ok ? 16 : 0));
        }
        asm volatile("cp.async.commit_group;\n cp.async.wait_group 0;" ::: "memory");
        // round to TF32 once (each thread its own chunks), as cuDNN's TF32 engines do on load
        for (int e = tid; e < C7_TROWS * C7_TCOLS * CHUNKS; e += C7_THREADS) {
            float4 *p = reinterpret_cast<float4 *>(s_t + (e / CHUNKS) * C7_PITCH + (e % CHUNKS) * 4);
            float4 v = *p;
            v.x = to_tf32(v.x); v.y = to_tf32(v.y); v.z = to_tf32(v.z); v.w = to_tf32(v.w);
            *p = v;
        }
        __syncthreads();

        // Warp = (group of four output rows rg, strip of 16 columns st).  An A fragment (one tile row, 16 pixels shifted by kx,
        // eight channels) is loaded ONCE and used for every output row it contributes to: tile row rg*4 + ii feeds output row
        // r = ii - ky of the group for each ky with 0 <= r < 4, so ten fragment loads serve the 28 (row, ky) pairs of a
        // (kx, k-step) instead of 28 -- the kernel is bound by shared-memory operand fetch, and this is 2.8 x less of it.
        // The seven B fragments of the (kx, k-step) wait in registers.
        const int rg = warp >> 2, st = warp & 3;
        const bool strip_live = x0 + st * 16 < W;
        float acc[4][2][4];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                acc[r][nt][0] = acc[r][nt][2] = bz[nt][0];
                acc[r][nt][1] = acc[r][nt][3] = bz[nt][1];
            }
        if (strip_live) {
            const unsigned a_base = smem_u32(s_t + ((rg * 4) * C7_TCOLS + st * 16 + lm_pix) * C7_PITCH + lm_ch);
#pragma unroll 1
            for (int kx = 0; kx < 7; ++kx) {
#pragma unroll
                for (int ks = 0; ks < 2; ++ks) {
                    float4 b[7];
#pragma unroll
                    for (int ky = 0; ky < 7; ++ky) b[ky] = s_w[((ky * 7 + kx) * 2 + ks) * 32 + lane];
#pragma unroll
                    for (int ii = 0; ii < 10; ++ii) {
                        unsigned a[4];
                        asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                                     : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3])
                                     : "r"(a_base + (unsigned)(((ii * C7_TCOLS + kx) * C7_PITCH + ks * 8) * 4)));
#pragma unroll
                        for (int ky = 0; ky < 7; ++ky) {
                            const int r = ii - ky;
                            if (r < 0 || r > 3) continue;   // compile time
                            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                                         : "+f"(acc[r][0][0]), "+f"(acc[r][0][1]), "+f"(acc[r][0][2]), "+f"(acc[r][0][3])
                                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(__float_as_uint(b[ky].x)), "r"(__float_as_uint(b[ky].y)));
                            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                                         : "+f"(acc[r][1][0]), "+f"(acc[r][1][1]), "+f"(acc[r][1][2]), "+f"(acc[r][1][3])
                                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(__float_as_uint(b[ky].z)), "r"(__float_as_uint(b[ky].w)));
                        }
                    }
                }
                if (CIN == 20) {   // channels 16..19: A = two 8 x 4 matrices (pixels 0-7 / 8-15), B = one register per n8 tile
                    float2 b4[7];
#pragma unroll
                    for (int ky = 0; ky < 7; ++ky) b4[ky] = s_w4[(ky * 7 + kx) * 32 + lane];
                    const unsigned a4 = a_base - (unsigned)(lm_ch * 4) + 16 * 4;   // same pixel rows, channel 16
#pragma unroll
                    for (int ii = 0; ii < 10; ++ii) {
                        unsigned a[2];
                        asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];"
                                     : "=r"(a[0]), "=r"(a[1])
                                     : "r"(a4 + (unsigned)(((ii * C7_TCOLS + kx) * C7_PITCH) * 4)));
#pragma unroll
                        for (int ky = 0; ky < 7; ++ky) {
                            const int r = ii - ky;
                            if (r < 0 || r > 3) continue;
                            asm volatile("mma.sync.aligned.m16n8k4.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                                         : "+f"(acc[r][0][0]), "+f"(acc[r][0][1]), "+f"(acc[r][0][2]), "+f"(acc[r][0][3])
                                         : "r"(a[0]), "r"(a[1]), "r"(__float_as_uint(b4[ky].x)));
                            asm volatile("mma.sync.aligned.m16n8k4.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                                         : "+f"(acc[r][1][0]), "+f"(acc[r][1][1]), "+f"(acc[r][1][2]), "+f"(acc[r][1][3])
                                         : "r"(a[0]), "r"(a[1]), "r"(__float_as_uint(b4[ky].y)));
                        }
                    }
                }
            }
            // epilogue: c0,c1 = pixel g, channels 2t,2t+1; c2,c3 = pixel g+8
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int y = y0 + rg * 4 + r;
                if (y >= H) continue;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const int x = x0 + st * 16 + g + half * 8;
                    if (x >= W) continue;
                    float *dst = out + (((size_t)n * H + y) * W + x) * 16 + 2 * t;
#pragma unroll
                    for (int nt = 0; nt < 2; ++nt) {
                        float v0 = acc[r][nt][half * 2], v1 = acc[r][nt][half * 2 + 1];
                        if (relu) { v0 = fmaxf(v0, 0.0f); v1 = fmaxf(v1, 0.0f); }
                        *reinterpret_cast<float2 *>(dst + nt * 8) = make_float2(v0, v1);
                    }
                }
            }
        }
    }
}

}  // namespace fvp

using namespace fvp;

extern "C" int fvp_conv7x7_c16_tf32(const float *in, const float *wfrag, const float *bias, int N, int H, int W, int relu,
                                    float *out, void *stream) {
    FVP_REQUIRE(in && wfrag && bias && out, "fvp_conv7x7_c16_tf32: null pointer");
    FVP_REQUIRE(N >= 0 && H >= 1 && W >= 1, "fvp_conv7x7_c16_tf32: bad shape N=%d H=%d W=%d", N, H, W);
    FVP_REQUIRE(((uintptr_t)in & 15) == 0 && ((uintptr_t)wfrag & 15) == 0 && ((uintptr_t)out & 7) == 0,
                "fvp_conv7x7_c16_tf32: misaligned pointer");
    if (N == 0) return FVP_OK;
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long ntiles = (long)N * ((H + C7_TH - 1) / C7_TH) * ((W + C7_TW - 1) / C7_TW);
    const int grid = (int)std::min<long>(ntiles, (long)sms);
    FVP_CUDA(cudaFuncSetAttribute(conv7x7_c16_tf32_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C7_SMEM));
    conv7x7_c16_tf32_kernel<16><<<grid, C7_THREADS, C7_SMEM, (cudaStream_t)stream>>>(in, reinterpret_cast<const float4 *>(wfrag), bias,
                                                                              out, N, H, W, relu);
    return check_launch("conv7x7_c16_tf32_kernel");
}

extern "C" int fvp_conv7x7_c20_tf32(const float *in, const float *wfrag, const float *bias, int N, int H, int W, int relu,
                                    float *out, void *stream) {
    FVP_REQUIRE(in && wfrag && bias && out, "fvp_conv7x7_c20_tf32: null pointer");
    FVP_REQUIRE(N >= 0 && H >= 1 && W >= 1, "fvp_conv7x7_c20_tf32: bad shape N=%d H=%d W=%d", N, H, W);
    FVP_REQUIRE(((uintptr_t)in & 15) == 0 && ((uintptr_t)wfrag & 15) == 0 && ((uintptr_t)out & 7) == 0,
                "fvp_conv7x7_c20_tf32: misaligned pointer");
    if (N == 0) return FVP_OK;
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long ntiles = (long)N * ((H + C7_TH - 1) / C7_TH) * ((W + C7_TW - 1) / C7_TW);
    const int grid = (int)std::min<long>(ntiles, (long)sms);
    FVP_CUDA(cudaFuncSetAttribute(conv7x7_c16_tf32_kernel<20>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C7_SMEM20));
    conv7x7_c16_tf32_kernel<20><<<grid, C7_THREADS, C7_SMEM20, (cudaStream_t)stream>>>(in, reinterpret_cast<const float4 *>(wfrag), bias,
                                                                                out, N, H, W, relu);
    return check_launch("conv7x7_c20_tf32_kernel");
}
