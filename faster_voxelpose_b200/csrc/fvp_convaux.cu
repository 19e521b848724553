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
    float rb[PW_CIN];               // RES: bias of the residual block whose add + ReLU is finished here
};

// out[n][co][y][x] = b[co] + sum_c w[co][c] * (a[n][y][x][c] + up[n][y/2][x/2][(y&1)*2+(x&1)][c])
// A warp owns 32 consecutive pixels (a flat index over n, y, x).  Its 4 KB of channel vectors come in as coalesced 16-byte
// loads -- lane l of pass k loads chunk l % 8 of pixel 4k + l / 8, the elementwise part (residual epilogue, shuffled
// addend) is applied chunk-wise -- and are staged in a padded shared-memory tile; then a lane is a pixel: it reads its 32
// channels (stride 33: conflict-free), forms the dot products and stores planar maps coalesced over x.  COUT is a
// compile-time bound of the accumulator array.
// RES: `a` is the second convolution of a residual block WITHOUT its epilogue; the block's output
// relu(a + rb + res) (cnns_2d.py:75-76) is formed here, on the way in -- the library's fused conv + add + ReLU costs
// 0.44 ms on the 960 planes of 64 x 64 where the plain convolution costs 0.20 ms.
template <int COUT, bool UP, bool RES>
__global__ void __launch_bounds__(128)
pointwise_planar_kernel(const float4 *__restrict__ a, const float4 *__restrict__ up, const float4 *__restrict__ res, int H, int W,
                        long pixels, const __grid_constant__ PointwiseParams prm, int cout, float *__restrict__ out) {
    __shared__ float tile[4][32 * 33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long p0 = (blockIdx.x * 4L + warp) * 32;      // first pixel of this warp
    if (p0 >= pixels) return;
    float *t = tile[warp];
    const int chunk = lane & 7;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int px = 4 * k + (lane >> 3);
        const long i = p0 + px;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < pixels) {
            v = ldg4(a + i * (PW_CIN / 4) + chunk);
            if (RES) {
                const float4 z = ldg4(res + i * (PW_CIN / 4) + chunk);
                v.x = fmaxf(v.x + prm.rb[4 * chunk + 0] + z.x, 0.0f); v.y = fmaxf(v.y + prm.rb[4 * chunk + 1] + z.y, 0.0f);
                v.z = fmaxf(v.z + prm.rb[4 * chunk + 2] + z.z, 0.0f); v.w = fmaxf(v.w + prm.rb[4 * chunk + 3] + z.w, 0.0f);
            }
            if (UP) {
                const int x = (int)(i % W);
                const long r = i / W;
                const int y = (int)(r % H);
                const long n = r / H;
                v = add4(v, ldg4(up + (((n * (H >> 1) + (y >> 1)) * (W >> 1) + (x >> 1)) * 4 + ((y & 1) * 2 + (x & 1))) * (PW_CIN / 4) + chunk));
            }
        }
        float *d = t + px * 33 + 4 * chunk;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    }
    __syncwarp();
    const long i = p0 + lane;
    if (i >= pixels) return;
    float acc[COUT];
#pragma unroll
    for (int co = 0; co < COUT; ++co) acc[co] = prm.b[co];
    const float *mine = t + lane * 33;
#pragma unroll
    for (int c = 0; c < PW_CIN; ++c) {
        const float v = mine[c];
#pragma unroll
        for (int co = 0; co < COUT; ++co) acc[co] = fmaf(prm.w[co * PW_CIN + c], v, acc[co]);
    }
    const int x = (int)(i % W);
    const long r = i / W;
    const int y = (int)(r % H);
    const long n = r / H;
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

template <bool UP, bool RES>
static int launch_pointwise(const float *a, const float *up, const float *res, int H, int W, long pixels, const PointwiseParams &prm,
                            int cout, float *out, cudaStream_t s) {
    const unsigned blocks = (unsigned)((pixels + 127) / 128);
    if (cout <= 16)
        pointwise_planar_kernel<16, UP, RES><<<blocks, 128, 0, s>>>((const float4 *)a, (const float4 *)up, (const float4 *)res, H, W,
                                                                    pixels, prm, cout, out);
    else
        pointwise_planar_kernel<PW_MAX_COUT, UP, RES><<<blocks, 128, 0, s>>>((const float4 *)a, (const float4 *)up, (const float4 *)res,
                                                                             H, W, pixels, prm, cout, out);
    return check_launch("pointwise_planar_kernel");
}

static int pointwise_impl(const char *who, const float *a, const float *res, const float *res_bias, const float *up, int N, int H,
                          int W, int C, const float *w, const float *b, int Cout, float *out, void *stream) {
    FVP_REQUIRE(a && w && b && out, "%s: null pointer", who);
    FVP_REQUIRE((res == nullptr) == (res_bias == nullptr), "%s: residual input and its bias go together", who);
    FVP_REQUIRE(C == PW_CIN, "%s: C must be %d (got %d)", who, PW_CIN, C);
    FVP_REQUIRE(Cout >= 1 && Cout <= PW_MAX_COUT, "%s: Cout %d outside [1,%d]", who, Cout, PW_MAX_COUT);
    FVP_REQUIRE(N >= 0 && H >= 1 && W >= 1, "%s: bad shape N=%d H=%d W=%d", who, N, H, W);
    FVP_REQUIRE(!up || (H % 2 == 0 && W % 2 == 0), "%s: the shuffled addend needs even H, W", who);
    const long pixels = (long)N * H * W;
    if (pixels == 0) return FVP_OK;
    PointwiseParams prm;
    for (int i = 0; i < PW_MAX_COUT * PW_CIN; ++i) prm.w[i] = i < Cout * C ? w[i] : 0.0f;
    for (int i = 0; i < PW_MAX_COUT; ++i) prm.b[i] = i < Cout ? b[i] : 0.0f;
    for (int i = 0; i < PW_CIN; ++i) prm.rb[i] = res_bias ? res_bias[i] : 0.0f;
    cudaStream_t s = (cudaStream_t)stream;
    if (res) return up ? launch_pointwise<true, true>(a, up, res, H, W, pixels, prm, Cout, out, s)
                       : launch_pointwise<false, true>(a, nullptr, res, H, W, pixels, prm, Cout, out, s);
    return up ? launch_pointwise<true, false>(a, up, nullptr, H, W, pixels, prm, Cout, out, s)
              : launch_pointwise<false, false>(a, nullptr, nullptr, H, W, pixels, prm, Cout, out, s);
}

extern "C" int fvp_pointwise_planar(const float *a, const float *up, int N, int H, int W, int C, const float *w,
                                    const float *b, int Cout, float *out, void *stream) {
    return pointwise_impl("fvp_pointwise_planar", a, nullptr, nullptr, up, N, H, W, C, w, b, Cout, out, stream);
}

extern "C" int fvp_pointwise_planar_res(const float *a, const float *res, const float *res_bias, const float *up, int N, int H,
                                        int W, int C, const float *w, const float *b, int Cout, float *out, void *stream) {
    FVP_REQUIRE(res && res_bias, "fvp_pointwise_planar_res: null residual input / bias");
    return pointwise_impl("fvp_pointwise_planar_res", a, res, res_bias, up, N, H, W, C, w, b, Cout, out, stream);
}
