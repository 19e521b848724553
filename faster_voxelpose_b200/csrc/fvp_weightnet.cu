// WeightNet front end as one kernel (reference lib/models/weight_net.py:36-88):
//   conv3x3(1 -> F, pad 1) + eval-mode BatchNorm (folded on the host) + MaxPool2d(2) + ReLU +
//   AdaptiveAvgPool2d(1)            per joint feature map [n][n] -> F numbers.
// The PyTorch path writes the [M][F][n][n] activation (7.5 GB for the Panoptic batch) and reads it back
// twice; here every map is staged once in shared memory and only [M][F] leaves the SM.
// ReLU and max-pool commute (both monotone), so relu(max(conv) + b) is evaluated per pooled cell.
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
