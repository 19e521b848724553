// Memory-bound glue of the 2-D conv nets' inference path (reference
// lib/models/cnns_2d.py:100-185 hourglass, :238-240 P2PNet output layer).  cuDNN keeps the dense
// convolutions; these three kernels replace the ATen elementwise passes around them, which the launch
// list of the Panoptic step showed at 2.6 ms of 8.2 ms for P2PNet (profiles/r01_launches_v4.csv):
//   * 2x2/stride-2 max-pool on channels-last activations (MaxPool2d(2), cnns_2d.py:113,116)
//   * ConvTranspose2d(k=2, s=2) re-expressed by the caller as a 1x1 conv to 4*C channels (a,b,c order);
//     here the pixel shuffle back to [2H,2W,C] rides with the skip addition  (cnns_2d.py:122-126)
//   * P2PNet's 1x1 output layer fused with that last shuffle + skip add, writing the planar
//     [N][Cout][H][W] maps the soft-argmax and WeightNet kernels consume (no NHWC->NCHW copy)
// All tensors fp32; activations channels-last ([N][H][W][C], C % 4 == 0).
#include "fvp_common.cuh"

namespace fvp {

__device__ __forceinline__ float4 ldg4(const float4 *p) { return __ldg(p); }

__device__ __forceinline__ float max_prop(float a, float b) {  // NaN-propagating, as ATen's max-pool
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float4 max4(float4 a, float4 b) {
    return make_float4(max_prop(a.x, b.x), max_prop(a.y, b.y), max_prop(a.z, b.z), max_prop(a.w, b.w));
}
__device__ __forceinline__ float4 add4(float4 a, float4 b) {
    return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}

// out[n][oh][ow][c] = max over the 2x2 window; floor semantics for odd H/W
__global__ void __launch_bounds__(256)
maxpool2_nhwc_kernel(const float4 *__restrict__ x, int H, int W, int C4, int OH, int OW, long total,
                     float4 *__restrict__ out) {
    for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += gridDim.x * 256L) {
        const int c = (int)(i % C4);
        long r = i / C4;
        const int ow = (int)(r % OW);
        r /= OW;
        const int oh = (int)(r % OH);
        const long n = r / OH;
        const float4 *p = x + ((n * H + 2 * oh) * W + 2 * ow) * C4 + c;
        const float4 a = ldg4(p), b = ldg4(p + C4), d = ldg4(p + (long)W * C4), e = ldg4(p + (long)W * C4 + C4);
        out[i] = max4(max4(a, b), max4(d, e));
    }
}

// out[n][2h+a][2w+b][c] = y[n][h][w][(a*2+b)*C + c] + s[n][2h+a][2w+b][c]
__global__ void __launch_bounds__(256)
upshuffle_add_kernel(const float4 *__restrict__ y, const float4 *__restrict__ s, int H, int W, int C4, long total,
                     float4 *__restrict__ out) {
    const int W2 = 2 * W, H2 = 2 * H;
    for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += gridDim.x * 256L) {
        const int c = (int)(i % C4);
        long r = i / C4;
        const int x2 = (int)(r % W2);
        r /= W2;
        const int y2 = (int)(r % H2);
        const long n = r / H2;
        const float4 *p = y + (((n * H + (y2 >> 1)) * W + (x2 >> 1)) * 4 + ((y2 & 1) * 2 + (x2 & 1))) * C4 + c;
        out[i] = add4(ldg4(p), ldg4(s + i));
    }
}

constexpr int PW_CIN = 32;
constexpr int PW_MAX_COUT = 20;

struct PointwiseParams {
    float w[PW_MAX_COUT * PW_CIN];  // [cout][cin], rows >= cout are zero
    float b[PW_MAX_COUT];
};

// out[n][co][y][x] = b[co] + sum_c w[co][c] * (a[n][y][x][c] + up[n][y/2][x/2][(y&1)*2+(x&1)][c])
// One thread per pixel: its 128-byte channel vector(s) come in as eight 16-byte loads, the planar stores
// are coalesced over x.  COUT is a compile-time bound of the accumulator array.
template <int COUT, bool UP>
__global__ void __launch_bounds__(128)
pointwise_planar_kernel(const float4 *__restrict__ a, const float4 *__restrict__ up, int H, int W, long pixels,
                        const __grid_constant__ PointwiseParams prm, int cout, float *__restrict__ out) {
    const long i = blockIdx.x * 128L + threadIdx.x;
    if (i >= pixels) return;
    const int x = (int)(i % W);
    long r = i / W;
    const int y = (int)(r % H);
    const long n = r / H;
    const float4 *pa = a + i * (PW_CIN / 4);
    const float4 *pu = nullptr;
    if (UP) pu = up + (((n * (H >> 1) + (y >> 1)) * (W >> 1) + (x >> 1)) * 4 + ((y & 1) * 2 + (x & 1))) * (PW_CIN / 4);
    float acc[COUT];
#pragma unroll
    for (int co = 0; co < COUT; ++co) acc[co] = prm.b[co];
#pragma unroll
    for (int q = 0; q < PW_CIN / 4; ++q) {
        float4 v = ldg4(pa + q);
        if (UP) v = add4(v, ldg4(pu + q));
#pragma unroll
        for (int co = 0; co < COUT; ++co) {
            const float *w = prm.w + co * PW_CIN + q * 4;
            acc[co] = fmaf(w[3], v.w, fmaf(w[2], v.z, fmaf(w[1], v.y, fmaf(w[0], v.x, acc[co]))));
        }
    }
    const long plane = (long)H * W;
    float *po = out + n * cout * plane + (long)y * W + x;
#pragma unroll
    for (int co = 0; co < COUT; ++co)
        if (co < cout) po[co * plane] = acc[co];
}

}  // namespace fvp

using namespace fvp;

static int elementwise_grid(long total) {
    int sms = 148;
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long want = (total + 255) / 256, cap = (long)sms * 16;
    return (int)(want < cap ? want : cap);
}

extern "C" int fvp_maxpool2_nhwc(const float *x, int N, int H, int W, int C, float *out, void *stream) {
    FVP_REQUIRE(x && out, "fvp_maxpool2_nhwc: null pointer");
    FVP_REQUIRE(N >= 0 && H >= 2 && W >= 2 && C > 0 && C % 4 == 0, "fvp_maxpool2_nhwc: bad shape N=%d H=%d W=%d C=%d", N,
                H, W, C);
    const int OH = H / 2, OW = W / 2, C4 = C / 4;
    const long total = (long)N * OH * OW * C4;
    if (total == 0) return FVP_OK;
    maxpool2_nhwc_kernel<<<elementwise_grid(total), 256, 0, (cudaStream_t)stream>>>((const float4 *)x, H, W, C4, OH, OW,
                                                                                   total, (float4 *)out);
    return check_launch("maxpool2_nhwc_kernel");
}

extern "C" int fvp_upshuffle_add(const float *y, const float *s, int N, int H, int W, int C, float *out, void *stream) {
    FVP_REQUIRE(y && s && out, "fvp_upshuffle_add: null pointer");
    FVP_REQUIRE(N >= 0 && H >= 1 && W >= 1 && C > 0 && C % 4 == 0, "fvp_upshuffle_add: bad shape N=%d H=%d W=%d C=%d", N,
                H, W, C);
    const int C4 = C / 4;
    const long total = (long)N * (2 * H) * (2 * W) * C4;
    if (total == 0) return FVP_OK;
    upshuffle_add_kernel<<<elementwise_grid(total), 256, 0, (cudaStream_t)stream>>>((const float4 *)y, (const float4 *)s,
                                                                                   H, W, C4, total, (float4 *)out);
    return check_launch("upshuffle_add_kernel");
}

template <bool UP>
static int launch_pointwise(const float *a, const float *up, int H, int W, long pixels, const PointwiseParams &prm,
                            int cout, float *out, cudaStream_t s) {
    const unsigned blocks = (unsigned)((pixels + 127) / 128);
    if (cout <= 16)
        pointwise_planar_kernel<16, UP><<<blocks, 128, 0, s>>>((const float4 *)a, (const float4 *)up, H, W, pixels, prm,
                                                               cout, out);
    else
        pointwise_planar_kernel<PW_MAX_COUT, UP><<<blocks, 128, 0, s>>>((const float4 *)a, (const float4 *)up, H, W,
                                                                        pixels, prm, cout, out);
    return check_launch("pointwise_planar_kernel");
}

extern "C" int fvp_pointwise_planar(const float *a, const float *up, int N, int H, int W, int C, const float *w,
                                    const float *b, int Cout, float *out, void *stream) {
    FVP_REQUIRE(a && w && b && out, "fvp_pointwise_planar: null pointer");
    FVP_REQUIRE(C == PW_CIN, "fvp_pointwise_planar: C must be %d (got %d)", PW_CIN, C);
    FVP_REQUIRE(Cout >= 1 && Cout <= PW_MAX_COUT, "fvp_pointwise_planar: Cout %d outside [1,%d]", Cout, PW_MAX_COUT);
    FVP_REQUIRE(N >= 0 && H >= 1 && W >= 1, "fvp_pointwise_planar: bad shape N=%d H=%d W=%d", N, H, W);
    FVP_REQUIRE(!up || (H % 2 == 0 && W % 2 == 0), "fvp_pointwise_planar: the shuffled addend needs even H, W");
    const long pixels = (long)N * H * W;
    if (pixels == 0) return FVP_OK;
    PointwiseParams prm;
    for (int i = 0; i < PW_MAX_COUT * PW_CIN; ++i) prm.w[i] = i < Cout * C ? w[i] : 0.0f;
    for (int i = 0; i < PW_MAX_COUT; ++i) prm.b[i] = i < Cout ? b[i] : 0.0f;
    cudaStream_t s = (cudaStream_t)stream;
    return up ? launch_pointwise<true>(a, up, H, W, pixels, prm, Cout, out, s)
              : launch_pointwise<false>(a, nullptr, H, W, pixels, prm, Cout, out, s);
}
