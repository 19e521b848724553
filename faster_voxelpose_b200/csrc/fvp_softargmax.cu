// Three-plane soft-argmax and weighted fusion.
#include <cstdint>

#include "fvp_common.cuh"

namespace fvp {

__device__ __forceinline__ float warp_max(float v) {
    for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// One CTA per (plane, person, joint) map of M cells: softmax(beta*x) expectation of the cell
// coordinates and the maximum probability (= 1/sum exp(beta*x - max)).
__global__ void __launch_bounds__(256)
softargmax_kernel(const float *__restrict__ feat, const float *__restrict__ grid, int NJ, int M, float beta,
                  float *__restrict__ pose, float *__restrict__ maxprob) {
    __shared__ float s_red[3][8];
    const int map = blockIdx.x;         // (a*N + n)*J + j
    const int a = map / NJ;
    const float *x = feat + (size_t)map * M;
    const float2 *g = reinterpret_cast<const float2 *>(grid) + (size_t)a * M;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    float mx = -INFINITY;
    for (int i = threadIdx.x; i < M; i += blockDim.x) mx = fmaxf(mx, beta * __ldg(x + i));
    mx = warp_max(mx);
    if (lane == 0) s_red[0][warp] = mx;
    __syncthreads();
    mx = s_red[0][0];
    for (int w = 1; w < 8; ++w) mx = fmaxf(mx, s_red[0][w]);
    __syncthreads();

    float s = 0.f, sx = 0.f, sy = 0.f;
    for (int i = threadIdx.x; i < M; i += blockDim.x) {
        const float e = expf(beta * __ldg(x + i) - mx);
        const float2 c = __ldg(g + i);
        s += e;
        sx = fmaf(e, c.x, sx);
        sy = fmaf(e, c.y, sy);
    }
    s = warp_sum(s); sx = warp_sum(sx); sy = warp_sum(sy);
    if (lane == 0) { s_red[0][warp] = s; s_red[1][warp] = sx; s_red[2][warp] = sy; }
    __syncthreads();
    if (threadIdx.x == 0) {
        s = sx = sy = 0.f;
        for (int w = 0; w < 8; ++w) { s += s_red[0][w]; sx += s_red[1][w]; sy += s_red[2][w]; }
        pose[2 * (size_t)map] = sx / s;
        pose[2 * (size_t)map + 1] = sy / s;
        maxprob[map] = 1.0f / s;
    }
}

// The same for maps of exactly 16 cells per thread (64 x 64 planes at 256 threads): the map is read once with
// 16-byte loads and stays in registers between the maximum and the exponential pass.
__global__ void __launch_bounds__(256)
softargmax16_kernel(const float *__restrict__ feat, const float *__restrict__ grid, int NJ, float beta,
                    float *__restrict__ pose, float *__restrict__ maxprob) {
    constexpr int M = 4096;
    __shared__ float s_red[3][8];
    const int map = blockIdx.x, a = map / NJ, tid = threadIdx.x;
    const float4 *x = reinterpret_cast<const float4 *>(feat + (size_t)map * M);
    const float4 *g = reinterpret_cast<const float4 *>(grid + (size_t)a * M * 2);
    const int lane = tid & 31, warp = tid >> 5;
    float4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        v[k] = __ldg(x + tid + 256 * k);
        v[k].x *= beta; v[k].y *= beta; v[k].z *= beta; v[k].w *= beta;
    }
    float mx = -INFINITY;
#pragma unroll
    for (int k = 0; k < 4; ++k) mx = fmaxf(mx, fmaxf(fmaxf(v[k].x, v[k].y), fmaxf(v[k].z, v[k].w)));
    mx = warp_max(mx);
    if (lane == 0) s_red[0][warp] = mx;
    __syncthreads();
    mx = s_red[0][0];
#pragma unroll
    for (int w = 1; w < 8; ++w) mx = fmaxf(mx, s_red[0][w]);
    __syncthreads();
    float s = 0.f, sx = 0.f, sy = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int cell = 4 * (tid + 256 * k);                 // float4 k holds cells cell .. cell+3
        const float4 c01 = __ldg(g + cell / 2), c23 = __ldg(g + cell / 2 + 1);
        const float e0 = expf(v[k].x - mx), e1 = expf(v[k].y - mx), e2 = expf(v[k].z - mx), e3 = expf(v[k].w - mx);
        s += (e0 + e1) + (e2 + e3);
        sx = fmaf(e0, c01.x, fmaf(e1, c01.z, fmaf(e2, c23.x, fmaf(e3, c23.z, sx))));
        sy = fmaf(e0, c01.y, fmaf(e1, c01.w, fmaf(e2, c23.y, fmaf(e3, c23.w, sy))));
    }
    s = warp_sum(s); sx = warp_sum(sx); sy = warp_sum(sy);
    if (lane == 0) { s_red[0][warp] = s; s_red[1][warp] = sx; s_red[2][warp] = sy; }
    __syncthreads();
    if (tid == 0) {
        s = sx = sy = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) { s += s_red[0][w]; sx += s_red[1][w]; sy += s_red[2][w]; }
        pose[2 * (size_t)map] = sx / s;
        pose[2 * (size_t)map + 1] = sy / s;
        maxprob[map] = 1.0f / s;
    }
}

// One thread per (person, joint): add plane offsets (joint_localization_net.py:168-171), fuse the
// three planes with normalised WeightNet weights (:83-120); thread j == 0 also averages the
// per-map maximum probabilities into the person confidence (:50-51).
__global__ void fuse_kernel(float *__restrict__ pose, const float *__restrict__ maxprob,
                            const float *__restrict__ offset, const float *__restrict__ weights, int N, int J,
                            float *__restrict__ fused, float *__restrict__ conf) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * J) return;
    const int n = i / J, j = i % J;
    const size_t S = (size_t)N * J;
    const float ox = offset[3 * n], oy = offset[3 * n + 1], oz = offset[3 * n + 2];
    float2 *p = reinterpret_cast<float2 *>(pose);
    float2 xy = p[i], xz = p[S + i], yz = p[2 * S + i];
    xy.x = __fadd_rn(xy.x, ox); xy.y = __fadd_rn(xy.y, oy);
    xz.x = __fadd_rn(xz.x, ox); xz.y = __fadd_rn(xz.y, oz);
    yz.x = __fadd_rn(yz.x, oy); yz.y = __fadd_rn(yz.y, oz);
    p[i] = xy; p[S + i] = xz; p[2 * S + i] = yz;
    const float wxy = weights[i], wxz = weights[S + i], wyz = weights[2 * S + i];
    const float sx = __fadd_rn(wxy, wxz), sy = __fadd_rn(wxy, wyz), sz = __fadd_rn(wxz, wyz);
    fused[3 * (size_t)i + 0] =
        __fadd_rn(__fmul_rn(__fdiv_rn(wxy, sx), xy.x), __fmul_rn(__fdiv_rn(wxz, sx), xz.x));
    fused[3 * (size_t)i + 1] =
        __fadd_rn(__fmul_rn(__fdiv_rn(wxy, sy), xy.y), __fmul_rn(__fdiv_rn(wyz, sy), yz.x));
    fused[3 * (size_t)i + 2] =
        __fadd_rn(__fmul_rn(__fdiv_rn(wxz, sz), xz.y), __fmul_rn(__fdiv_rn(wyz, sz), yz.y));
    if (j == 0 && conf != nullptr) {
        float acc = 0.f;
        for (int a = 0; a < 3; ++a)
            for (int q = 0; q < J; ++q) acc += maxprob[(size_t)a * S + (size_t)n * J + q];
        conf[n] = acc / (float)(3 * J);
    }
}

}  // namespace fvp

using namespace fvp;

extern "C" int fvp_softargmax(const float *feat, const float *grid, int N, int J, int M, float beta, float *pose,
                              float *maxprob, void *stream) {
    FVP_REQUIRE(feat && grid && pose && maxprob, "fvp_softargmax: null pointer");
    FVP_REQUIRE(J > 0 && M > 0, "fvp_softargmax: bad dims");
    if (N == 0) return FVP_OK;
    FVP_REQUIRE(N > 0, "fvp_softargmax: N < 0");
    if (M == 4096 && ((uintptr_t)feat & 15) == 0 && ((uintptr_t)grid & 15) == 0)
        softargmax16_kernel<<<3 * N * J, 256, 0, (cudaStream_t)stream>>>(feat, grid, N * J, beta, pose, maxprob);
    else
        softargmax_kernel<<<3 * N * J, 256, 0, (cudaStream_t)stream>>>(feat, grid, N * J, M, beta, pose, maxprob);
    return check_launch("softargmax_kernel");
}

extern "C" int fvp_fuse(float *pose, const float *maxprob, const float *offset, const float *weights, int N, int J,
                        float *fused, float *conf, void *stream) {
    FVP_REQUIRE(pose && maxprob && offset && weights && fused, "fvp_fuse: null pointer");
    FVP_REQUIRE(J > 0, "fvp_fuse: bad dims");
    if (N == 0) return FVP_OK;
    FVP_REQUIRE(N > 0, "fvp_fuse: N < 0");
    fuse_kernel<<<ceil_div(N * J, 128), 128, 0, (cudaStream_t)stream>>>(pose, maxprob, offset, weights, N, J, fused, conf);
    return check_launch("fuse_kernel");
}
