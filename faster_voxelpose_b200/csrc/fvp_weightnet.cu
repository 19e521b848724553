// WeightNet front end as one kernel (reference lib/models/weight_net.py:36-88):
//   conv3x3(1 -> F, pad 1) + eval-mode BatchNorm (folded on the host) + MaxPool2d(2) + ReLU +
//   AdaptiveAvgPool2d(1)            per joint feature map [n][n] -> F numbers.
// The PyTorch path writes the [M][F][n][n] activation (7.5 GB for the Panoptic batch) and reads it back
// twice; here every map is staged once in shared memory and only [M][F] leaves the SM.
// ReLU and max-pool commute (both monotone), so relu(max(conv) + b) is evaluated per pooled cell.
#include <cstdint>

#include "fvp_common.cuh"

namespace fvp {

constexpr int WN_MAX_F = 64;

struct WeightNetParams {
    float w[WN_MAX_F * 9];  // folded conv weight [F][3][3]
    float b[WN_MAX_F];      // folded bias
};

template <int F>
__global__ void __launch_bounds__(256)
weightnet_features_kernel(const float *__restrict__ maps, int n, const __grid_constant__ WeightNetParams prm,
                          float *__restrict__ out) {
    extern __shared__ float s_map[];  // [(n + 2)][pitch], zero halo
    __shared__ float s_part[8][F];
    const int pitch = n + 3;          // odd pitch for even n: the 2-row strides of the pooled cells spread over banks
    const int m = blockIdx.x, tid = threadIdx.x;
    const float *src = maps + (size_t)m * n * n;
    for (int e = tid; e < (n + 2) * pitch; e += 256) s_map[e] = 0.0f;
    __syncthreads();
    for (int e = tid; e < n * n; e += 256) {
        const int r = e / n, c = e - r * n;
        s_map[(r + 1) * pitch + c + 1] = __ldg(src + e);
    }
    __syncthreads();

    float sum[F];
#pragma unroll
    for (int f = 0; f < F; ++f) sum[f] = 0.0f;
    const int np = n >> 1;  // pooled cells per side (MaxPool2d(2) floors)
    for (int cell = tid; cell < np * np; cell += 256) {
        const int pr = cell / np, pc = cell - pr * np;
        // input patch rows 2pr-1..2pr+2, cols 2pc-1..2pc+2 (halo-shifted by +1), held as the three horizontal
        // pairs of every row: one two-wide FMA then feeds the left and right conv output of a pooled cell
        float2 pp[4][3];
#pragma unroll
        for (int dy = 0; dy < 4; ++dy) {
            const float *row = s_map + (2 * pr + dy) * pitch + 2 * pc;
            const float a = row[0], b = row[1], c = row[2], d = row[3];
            pp[dy][0] = make_float2(a, b); pp[dy][1] = make_float2(b, c); pp[dy][2] = make_float2(c, d);
        }
#pragma unroll
        for (int f = 0; f < F; ++f) {
            float2 top = make_float2(0.f, 0.f), bot = make_float2(0.f, 0.f);  // (o00, o01), (o10, o11)
#pragma unroll
            for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) {
                    const float w = prm.w[f * 9 + ky * 3 + kx];
                    const float2 ww = make_float2(w, w);
                    top = __ffma2_rn(ww, pp[ky][kx], top);
                    bot = __ffma2_rn(ww, pp[ky + 1][kx], bot);
                }
            const float mx = fmaxf(fmaxf(top.x, top.y), fmaxf(bot.x, bot.y)) + prm.b[f];
            sum[f] += fmaxf(mx, 0.0f);
        }
    }
    // block reduction of the F sums
    const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
    for (int f = 0; f < F; ++f) {
        float v = sum[f];
#pragma unroll
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_part[warp][f] = v;
    }
    __syncthreads();
    if (tid < F) {
        float v = 0.0f;
#pragma unroll
        for (int w = 0; w < 8; ++w) v += s_part[w][tid];
        out[(size_t)m * F + tid] = v / (float)(np * np);
    }
}

// ------------------------------------------------------------------------------------------------
// The same front end on TF32 tensor cores (used where torch itself runs TF32 convolutions -- its default, and
// what the reference runs with): an implicit GEMM with M = pixels, N = features, K = taps.  One m16n8k8 mma per
// feature octet covers taps 0..7 of 16 pixels; the ninth tap is one fp32 FMA per accumulator.
//   m-tiles = two per iteration over a block of 8 x by 4 y pixels: fragment rows g / g+8 of tile A are the top-left /
//             top-right pixel of pooled cell g, those of tile B the bottom pair, so the 2x2 max-pool is three
//             in-thread max operations per feature column -- no shuffles
//   A frag  = plain LDS.32 from the zero-haloed map (pitch = 9 mod 32 floats spreads the four map rows an
//             instruction touches over the banks), rounded to TF32 once when the map is staged
//   B frag  = the folded filters, 2 registers per feature octet, loaded once per thread
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float wn_tf32(float x) {
    unsigned r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

template <int F>
__global__ void __launch_bounds__(256)
weightnet_features_tf32_kernel(const float *__restrict__ maps, int n, int pitch,
                               const __grid_constant__ WeightNetParams prm, float *__restrict__ out) {
    constexpr int NT = F / 8;
    extern __shared__ float s_map[];  // [(n + 2)][pitch], zero halo, TF32-rounded
    __shared__ float s_part[8][F];
    __shared__ float s_wb[F * 10];    // [F][9] filters + [F] bias (kernel parameters are not lane-indexable)
    const int m = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const float *src = maps + (size_t)m * n * n;
    for (int e = tid; e < (n + 2) * pitch; e += 256) s_map[e] = 0.0f;
    for (int e = tid; e < F * 10; e += 256) s_wb[e] = e < F * 9 ? prm.w[e] : prm.b[e - F * 9];
    __syncthreads();
    for (int e = tid; e < n * n / 4; e += 256) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(src) + e);
        const int r = (4 * e) / n, c = 4 * e - r * n;
        float *d = s_map + (r + 1) * pitch + c + 1;
        d[0] = wn_tf32(v.x); d[1] = wn_tf32(v.y); d[2] = wn_tf32(v.z); d[3] = wn_tf32(v.w);
    }
    // B fragments: b0 = W[tap t][feature nt*8 + g], b1 = W[tap t + 4][feature nt*8 + g]; ninth tap and bias for the
    // two features this thread accumulates per octet (columns 2t, 2t + 1 of the C fragment)
    unsigned b0[NT], b1[NT];
    float w8[NT][2], bias[NT][2], sum[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        b0[nt] = __float_as_uint(wn_tf32(s_wb[(nt * 8 + g) * 9 + t]));
        b1[nt] = __float_as_uint(wn_tf32(s_wb[(nt * 8 + g) * 9 + t + 4]));
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            w8[nt][h] = wn_tf32(s_wb[(nt * 8 + 2 * t + h) * 9 + 8]);
            bias[nt][h] = s_wb[F * 9 + nt * 8 + 2 * t + h];
            sum[nt][h] = 0.0f;
        }
    }
    __syncthreads();
    // tap k of the A fragment: k = t (a0, a1) and t + 4 (a2, a3); offset of tap (ky, kx) in the haloed map
    const int o_lo = (t / 3) * pitch + (t % 3), o_hi = ((t + 4) / 3) * pitch + ((t + 4) % 3), o_8 = 2 * pitch + 2;
    // One iteration = a block of 8 x by 4 y output pixels = 8 pooled cells, one per g: cell (cy, cx) = (g / 4, g % 4).
    // Two m-tiles: tile A holds the cell's top pixels (fragment row g = left, g + 8 = right), tile B the bottom ones,
    // so every thread owns all four conv outputs of its cell for its two feature columns.
    const int cell_off = (2 * (g >> 2)) * pitch + 2 * (g & 3);
    const int blocks_x = n >> 3, blocks_y = n >> 2;
    // warp w takes blocks w, w + 8, ... in row-major order; (by, bx) advance without a division
    int by = warp / blocks_x, bx = warp - by * blocks_x;
    for (; by < blocks_y; bx += 8) {
        while (bx >= blocks_x) { bx -= blocks_x; ++by; }
        if (by >= blocks_y) break;
        const float *pa = s_map + (4 * by) * pitch + 8 * bx + cell_off;  // top-left pixel of the cell (haloed origin)
        const float *pb = pa + pitch;
        const unsigned a0 = __float_as_uint(pa[o_lo]), a1 = __float_as_uint(pa[o_lo + 1]);
        const unsigned a2 = __float_as_uint(pa[o_hi]), a3 = __float_as_uint(pa[o_hi + 1]);
        const unsigned e0 = __float_as_uint(pb[o_lo]), e1 = __float_as_uint(pb[o_lo + 1]);
        const unsigned e2 = __float_as_uint(pb[o_hi]), e3 = __float_as_uint(pb[o_hi + 1]);
        const float xa0 = pa[o_8], xa1 = pa[o_8 + 1], xb0 = pb[o_8], xb1 = pb[o_8 + 1];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            float c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f, d0 = 0.f, d1 = 0.f, d2 = 0.f, d3 = 0.f;
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c0), "+f"(c1), "+f"(c2), "+f"(c3)
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0[nt]), "r"(b1[nt]));
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(d0), "+f"(d1), "+f"(d2), "+f"(d3)
                         : "r"(e0), "r"(e1), "r"(e2), "r"(e3), "r"(b0[nt]), "r"(b1[nt]));
            c0 = fmaf(w8[nt][0], xa0, c0); c1 = fmaf(w8[nt][1], xa0, c1);
            c2 = fmaf(w8[nt][0], xa1, c2); c3 = fmaf(w8[nt][1], xa1, c3);
            d0 = fmaf(w8[nt][0], xb0, d0); d1 = fmaf(w8[nt][1], xb0, d1);
            d2 = fmaf(w8[nt][0], xb1, d2); d3 = fmaf(w8[nt][1], xb1, d3);
            const float m0 = fmaxf(fmaxf(c0, c2), fmaxf(d0, d2)), m1 = fmaxf(fmaxf(c1, c3), fmaxf(d1, d3));
            sum[nt][0] += fmaxf(m0 + bias[nt][0], 0.0f);
            sum[nt][1] += fmaxf(m1 + bias[nt][1], 0.0f);
        }
    }
    // reduce over the eight g of a warp (lanes with equal t), then over the warps
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float v = sum[nt][h];
            v += __shfl_xor_sync(0xffffffffu, v, 4);
            v += __shfl_xor_sync(0xffffffffu, v, 8);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            if (g == 0) s_part[warp][nt * 8 + 2 * t + h] = v;
        }
    __syncthreads();
    if (tid < F) {
        float v = 0.0f;
#pragma unroll
        for (int w = 0; w < 8; ++w) v += s_part[w][tid];
        const int np = n >> 1;
        out[(size_t)m * F + tid] = v / (float)(np * np);
    }
}

}  // namespace fvp

using namespace fvp;

template <int F>
static int launch_weightnet(const float *maps, int M, int n, const WeightNetParams &prm, float *out, cudaStream_t s) {
    const size_t smem = sizeof(float) * (size_t)(n + 2) * (n + 3);
    auto kern = weightnet_features_kernel<F>;
    if (smem > 48 * 1024) FVP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<M, 256, smem, s>>>(maps, n, prm, out);
    return check_launch("weightnet_features_kernel");
}

template <int F>
static int launch_weightnet_tf32(const float *maps, int M, int n, const WeightNetParams &prm, float *out, cudaStream_t s) {
    const int pitch = ((n + 2 + 22) / 32) * 32 + 9;   // >= n + 2 and = 9 (mod 32)
    const size_t smem = sizeof(float) * (size_t)(n + 2) * pitch;
    auto kern = weightnet_features_tf32_kernel<F>;
    if (smem > 48 * 1024) FVP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<M, 256, smem, s>>>(maps, n, pitch, prm, out);
    return check_launch("weightnet_features_tf32_kernel");
}

extern "C" int fvp_weightnet_features_tf32(const float *maps, int M, int n, const float *w /*HOST [F][9]*/,
                                           const float *b /*HOST [F]*/, int F, float *out, void *stream) {
    FVP_REQUIRE(maps && w && b && out, "fvp_weightnet_features_tf32: null pointer");
    FVP_REQUIRE(n >= 8 && n <= 160 && n % 8 == 0, "fvp_weightnet_features_tf32: map side %d must be a multiple of 8 in [8,160]", n);
    FVP_REQUIRE(F == 16 || F == 32 || F == 64, "fvp_weightnet_features_tf32: F must be 16, 32 or 64 (got %d)", F);
    FVP_REQUIRE((reinterpret_cast<uintptr_t>(maps) & 15) == 0, "fvp_weightnet_features_tf32: maps must be 16-byte aligned");
    if (M == 0) return FVP_OK;
    FVP_REQUIRE(M > 0, "fvp_weightnet_features_tf32: M < 0");
    WeightNetParams prm;
    for (int i = 0; i < F * 9; ++i) prm.w[i] = w[i];
    for (int i = 0; i < F; ++i) prm.b[i] = b[i];
    cudaStream_t s = (cudaStream_t)stream;
    switch (F) {
        case 16: return launch_weightnet_tf32<16>(maps, M, n, prm, out, s);
        case 32: return launch_weightnet_tf32<32>(maps, M, n, prm, out, s);
        default: return launch_weightnet_tf32<64>(maps, M, n, prm, out, s);
    }
}

extern "C" int fvp_weightnet_features(const float *maps, int M, int n, const float *w /*HOST [F][9]*/,
                                      const float *b /*HOST [F]*/, int F, float *out, void *stream) {
    FVP_REQUIRE(maps && w && b && out, "fvp_weightnet_features: null pointer");
    FVP_REQUIRE(n >= 2 && n <= 200, "fvp_weightnet_features: map side %d outside [2,200]", n);
    FVP_REQUIRE(F == 16 || F == 32 || F == 64, "fvp_weightnet_features: F must be 16, 32 or 64 (got %d)", F);
    if (M == 0) return FVP_OK;
    FVP_REQUIRE(M > 0, "fvp_weightnet_features: M < 0");
    WeightNetParams prm;
    for (int i = 0; i < F * 9; ++i) prm.w[i] = w[i];
    for (int i = 0; i < F; ++i) prm.b[i] = b[i];
    cudaStream_t s = (cudaStream_t)stream;
    switch (F) {
        case 16: return launch_weightnet<16>(maps, M, n, prm, out, s);
        case 32: return launch_weightnet<32>(maps, M, n, prm, out, s);
        default: return launch_weightnet<64>(maps, M, n, prm, out, s);
    }
}
