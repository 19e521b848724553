// C2CNet inference in ONE kernel (reference lib/models/cnns_1d.py:165-241 C2CNet,
// :14-163 its blocks; called from lib/models/human_detection_net.py:203-206 on the P z-columns of every
// frame).  The net is tiny -- [B*P][J][20] in, [B*P][20] out, 32/64/128 channels at lengths 20/10/5,
// ~5 MFLOP per column -- and as ~150 separate cuDNN/ATen launches it cost 1 ms of a 14.5 ms step.  Here a
// CTA keeps S columns' activations in shared memory for the whole network and streams the (BatchNorm-
// folded) weights from L2 once per CTA.
//
// The caller (lib/models/cnns.py) flattens the module tree into a short op list over twelve activation
// buffers per column: four of each resolution class
//      class 0: 32 rows x (20 + 2*3 halo)   class 1: 64 rows x (10 + 2)   class 2: 128 rows x (5 + 2)
// Halo cells are zeroed once and never written: they are the zero padding of the convolutions.
// Work split of a conv: thread = (output channel, run of 5 consecutive positions); with 128*S threads every
// layer of the hourglass (32x20, 64x10, 128x5 outputs per column) occupies exactly all threads.
#include <algorithm>

#include "fvp_common.cuh"

namespace fvp {

constexpr int C2C_L = 20;          // column length the buffer classes are laid out for
constexpr int C2C_RUN = 5;         // positions per thread
constexpr int C2C_MAX_OPS = 32;
constexpr int C2C_SAMPLE = 4 * 32 * 26 + 4 * 64 * 12 + 4 * 128 * 7;  // floats per column: 9984

enum { C2C_CONV = 0, C2C_POOL = 1, C2C_UP = 2, C2C_OUT = 3 };

struct C2COp {
    int type, in, out, add;   // buffer ids (class*4 + index), add < 0: none
    int cin, cout_log2, k, relu;
    int w_off, b_off;         // offsets into the flat weight buffer
};

struct C2CParams {
    const float *x;        // [n][J][20]
    const float *weights;  // flat, device
    float *out;            // [n][20]
    int n, J, nops;
    C2COp ops[C2C_MAX_OPS];
};

struct Buf {
    int off, LP, halo, L;
};
__device__ __forceinline__ Buf buf_of(int id) {
    const int cls = id >> 2, idx = id & 3;
    Buf b;
    if (cls == 0) { b.LP = 26; b.halo = 3; b.L = 20; b.off = idx * 32 * 26; }
    else if (cls == 1) { b.LP = 12; b.halo = 1; b.L = 10; b.off = 4 * 32 * 26 + idx * 64 * 12; }
    else { b.LP = 7; b.halo = 1; b.L = 5; b.off = 4 * 32 * 26 + 4 * 64 * 12 + idx * 128 * 7; }
    return b;
}

// out[co][l] = act(b[co] + sum_{ci,kk} W[ci][kk][co] * in[ci][l + kk - K/2] (+ add[co][l]))
template <int K, int S>
__device__ __forceinline__ void conv_op(float *sm, const C2COp &op, const float *__restrict__ wts) {
    const Buf bi = buf_of(op.in), bo = buf_of(op.out);
    const int cout = 1 << op.cout_log2, t = threadIdx.x;
    const int co = t & (cout - 1), run = t >> op.cout_log2, runs_per = bi.L / C2C_RUN;
    if (run >= S * runs_per) return;
    const int s = run / runs_per, l0 = (run - s * runs_per) * C2C_RUN;
    const float *xin = sm + s * C2C_SAMPLE + bi.off + bi.halo - K / 2 + l0;
    const float *w = wts + op.w_off + co;
    float acc[C2C_RUN];
    const float bias = __ldg(wts + op.b_off + co);
#pragma unroll
    for (int j = 0; j < C2C_RUN; ++j) acc[j] = bias;
    // The filters come from L2 (one coalesced 128-byte line per warp and tap); with a dozen warps per SM that
    // latency is only hidden if the next group's filters are in flight while this group is multiplied.
    constexpr int G = K == 7 ? 2 : 8;  // input channels per group
    float wn[G][K];
    auto fetch = [&](int c0) {
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int kk = 0; kk < K; ++kk)
                wn[g][kk] = (c0 + g < op.cin) ? __ldg(w + (((c0 + g) * K + kk) << op.cout_log2)) : 0.0f;
    };
    fetch(0);
    for (int c0 = 0; c0 < op.cin; c0 += G) {
        float wc[G][K];
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int kk = 0; kk < K; ++kk) wc[g][kk] = wn[g][kk];
        fetch(c0 + G);
#pragma unroll
        for (int g = 0; g < G; ++g) {
            if (c0 + g >= op.cin) break;
            float xv[C2C_RUN + K - 1];
#pragma unroll
            for (int j = 0; j < C2C_RUN + K - 1; ++j) xv[j] = xin[(c0 + g) * bi.LP + j];
#pragma unroll
            for (int kk = 0; kk < K; ++kk)
#pragma unroll
                for (int j = 0; j < C2C_RUN; ++j) acc[j] = fmaf(wc[g][kk], xv[j + kk], acc[j]);
        }
    }
    float *dst = sm + s * C2C_SAMPLE + bo.off + co * bo.LP + bo.halo + l0;
    if (op.add >= 0) {
        const Buf ba = buf_of(op.add);
        const float *ad = sm + s * C2C_SAMPLE + ba.off + co * ba.LP + ba.halo + l0;
#pragma unroll
        for (int j = 0; j < C2C_RUN; ++j) acc[j] += ad[j];
    }
#pragma unroll
    for (int j = 0; j < C2C_RUN; ++j) dst[j] = op.relu ? fmaxf(acc[j], 0.0f) : acc[j];
}

// ConvTranspose1d(k=2, s=2) + BN + ReLU, then + skip:  out[co][2l+a] = relu(b[co] + sum_ci W[ci][a][co] in[ci][l]) + add
template <int S>
__device__ __forceinline__ void up_op(float *sm, const C2COp &op, const float *__restrict__ wts) {
    const Buf bi = buf_of(op.in), bo = buf_of(op.out);
    const int cout = 1 << op.cout_log2, t = threadIdx.x;
    const int cv = t & (2 * cout - 1), a = cv >> op.cout_log2, co = cv & (cout - 1);
    const int run = t >> (op.cout_log2 + 1), runs_per = bi.L / C2C_RUN;
    if (run >= S * runs_per) return;
    const int s = run / runs_per, l0 = (run - s * runs_per) * C2C_RUN;
    const float *xin = sm + s * C2C_SAMPLE + bi.off + bi.halo + l0;
    const float *w = wts + op.w_off + cv;
    float acc[C2C_RUN];
    const float bias = __ldg(wts + op.b_off + co);
#pragma unroll
    for (int j = 0; j < C2C_RUN; ++j) acc[j] = bias;
    constexpr int G = 16;  // same filter pipelining as conv_op
    float wn[G];
    auto fetch = [&](int c0) {
#pragma unroll
        for (int g = 0; g < G; ++g) wn[g] = (c0 + g < op.cin) ? __ldg(w + (c0 + g) * 2 * cout) : 0.0f;
    };
    fetch(0);
    for (int c0 = 0; c0 < op.cin; c0 += G) {
        float wc[G];
#pragma unroll
        for (int g = 0; g < G; ++g) wc[g] = wn[g];
        fetch(c0 + G);
#pragma unroll
        for (int g = 0; g < G; ++g) {
            if (c0 + g >= op.cin) break;
#pragma unroll
            for (int j = 0; j < C2C_RUN; ++j) acc[j] = fmaf(wc[g], xin[(c0 + g) * bi.LP + j], acc[j]);
        }
    }
    const Buf ba = buf_of(op.add);
    const float *ad = sm + s * C2C_SAMPLE + ba.off + co * ba.LP + ba.halo + 2 * l0 + a;
    float *dst = sm + s * C2C_SAMPLE + bo.off + co * bo.LP + bo.halo + 2 * l0 + a;
#pragma unroll
    for (int j = 0; j < C2C_RUN; ++j) dst[2 * j] = fmaxf(acc[j], 0.0f) + ad[2 * j];
}

template <int S>
__global__ void __launch_bounds__(128 * S)
c2c_forward_kernel(const __grid_constant__ C2CParams p) {
    extern __shared__ __align__(16) float sm[];
    const int t = threadIdx.x, nt = 128 * S;
    const int n0 = blockIdx.x * S;
    for (int e = t; e < S * C2C_SAMPLE; e += nt) sm[e] = 0.0f;
    __syncthreads();
    {   // input columns -> buffer 0 (class 0, rows 0..J-1)
        const Buf b = buf_of(0);
        for (int e = t; e < S * p.J * C2C_L; e += nt) {
            const int s = e / (p.J * C2C_L), r = e - s * (p.J * C2C_L), j = r / C2C_L, l = r - j * C2C_L;
            if (n0 + s < p.n) sm[s * C2C_SAMPLE + b.off + j * b.LP + b.halo + l] = __ldg(p.x + (size_t)(n0 + s) * p.J * C2C_L + r);
        }
    }
    __syncthreads();
    for (int i = 0; i < p.nops; ++i) {
        const C2COp &op = p.ops[i];
        if (op.type == C2C_CONV) {
            if (op.k == 3) conv_op<3, S>(sm, op, p.weights);
            else if (op.k == 1) conv_op<1, S>(sm, op, p.weights);
            else conv_op<7, S>(sm, op, p.weights);
        } else if (op.type == C2C_UP) {
            up_op<S>(sm, op, p.weights);
        } else if (op.type == C2C_POOL) {
            const Buf bi = buf_of(op.in), bo = buf_of(op.out);
            for (int e = t; e < S * op.cin * bo.L; e += nt) {
                const int s = e / (op.cin * bo.L), r = e - s * (op.cin * bo.L), c = r / bo.L, l = r - c * bo.L;
                const float *src = sm + s * C2C_SAMPLE + bi.off + c * bi.LP + bi.halo + 2 * l;
                sm[s * C2C_SAMPLE + bo.off + c * bo.LP + bo.halo + l] = fmaxf(src[0], src[1]);
            }
        } else {  // C2C_OUT: 1x1 conv to one channel, straight to global memory
            const Buf bi = buf_of(op.in);
            if (t < S * C2C_L) {
                const int s = t / C2C_L, l = t - s * C2C_L;
                float acc = __ldg(p.weights + op.b_off);
                const float *src = sm + s * C2C_SAMPLE + bi.off + bi.halo + l;
                for (int ci = 0; ci < op.cin; ++ci) acc = fmaf(__ldg(p.weights + op.w_off + ci), src[ci * bi.LP], acc);
                if (n0 + s < p.n) p.out[(size_t)(n0 + s) * C2C_L + l] = acc;
            }
        }
        __syncthreads();
    }
}

}  // namespace fvp

using namespace fvp;

template <int S>
static int launch_c2c(const C2CParams &p, cudaStream_t s) {
    const size_t smem = sizeof(float) * (size_t)S * C2C_SAMPLE;
    auto kern = c2c_forward_kernel<S>;
    FVP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)ceil_div(p.n, S), 128 * S, smem, s>>>(p);
    return check_launch("c2c_forward_kernel");
}

extern "C" int fvp_c2c_forward(const float *x, int n, int J, int L, const float *weights, const int32_t *ops /*HOST*/,
                               int nops, float *out, void *stream) {
    FVP_REQUIRE(x && weights && ops && out, "fvp_c2c_forward: null pointer");
    FVP_REQUIRE(L == C2C_L, "fvp_c2c_forward: column length must be %d (got %d)", C2C_L, L);
    FVP_REQUIRE(J >= 1 && J <= 32, "fvp_c2c_forward: J %d outside [1,32]", J);
    FVP_REQUIRE(nops >= 1 && nops <= C2C_MAX_OPS, "fvp_c2c_forward: %d ops outside [1,%d]", nops, C2C_MAX_OPS);
    if (n == 0) return FVP_OK;
    FVP_REQUIRE(n > 0, "fvp_c2c_forward: n < 0");
    C2CParams p;
    p.x = x; p.weights = weights; p.out = out; p.n = n; p.J = J; p.nops = nops;
    const int rows[3] = {32, 64, 128};
    for (int i = 0; i < nops; ++i) {
        C2COp &o = p.ops[i];
        const int32_t *r = ops + 10 * i;
        o.type = r[0]; o.in = r[1]; o.out = r[2]; o.add = r[3]; o.cin = r[4]; o.cout_log2 = r[5]; o.k = r[6]; o.relu = r[7];
        o.w_off = r[8]; o.b_off = r[9];
        FVP_REQUIRE(o.type >= C2C_CONV && o.type <= C2C_OUT, "fvp_c2c_forward: op %d has unknown type %d", i, o.type);
        FVP_REQUIRE(o.in >= 0 && o.in < 12 && (o.type == C2C_OUT || (o.out >= 0 && o.out < 12)) && o.add < 12,
                    "fvp_c2c_forward: op %d buffer id out of range", i);
        FVP_REQUIRE(o.cin >= 1 && o.cin <= rows[o.in >> 2], "fvp_c2c_forward: op %d reads %d rows of a %d-row buffer", i,
                    o.cin, rows[o.in >> 2]);
        const int ci = o.in >> 2, co = o.out >> 2;
        if (o.type == C2C_CONV) {
            FVP_REQUIRE(ci == co && o.out != o.in && o.out != o.add, "fvp_c2c_forward: conv op %d buffer classes/aliasing", i);
            FVP_REQUIRE(o.k == 1 || o.k == 3 || (o.k == 7 && ci == 0), "fvp_c2c_forward: op %d kernel size %d", i, o.k);
            FVP_REQUIRE((1 << o.cout_log2) <= rows[co] && (1 << o.cout_log2) >= 16, "fvp_c2c_forward: op %d Cout", i);
            FVP_REQUIRE(o.add < 0 || (o.add >> 2) == co, "fvp_c2c_forward: op %d add buffer class", i);
        } else if (o.type == C2C_UP) {
            FVP_REQUIRE(co == ci - 1 && o.add >= 0 && (o.add >> 2) == co && o.add != o.out,
                        "fvp_c2c_forward: up op %d buffer classes", i);
            FVP_REQUIRE((1 << o.cout_log2) <= rows[co] && (2 << o.cout_log2) * (C2C_L >> ci) / C2C_RUN <= 128,
                        "fvp_c2c_forward: up op %d Cout", i);
        } else if (o.type == C2C_POOL) {
            FVP_REQUIRE(co == ci + 1 && o.cin <= rows[co], "fvp_c2c_forward: pool op %d buffer classes", i);
        } else {
            FVP_REQUIRE(ci == 0, "fvp_c2c_forward: the output op reads a full-resolution buffer");
        }
        if (o.type == C2C_CONV)  // every thread = (channel, run): all runs must fit the block
            FVP_REQUIRE((1 << o.cout_log2) * ((C2C_L >> ci) / C2C_RUN) <= 128, "fvp_c2c_forward: op %d does not fit", i);
    }
    // columns per CTA: one while every CTA can be resident at once (40 KB of shared memory each -> 5 per SM; measured
    // 0.21 ms for 320 columns against 0.22-0.29 ms for 2-4 columns per CTA), more only to bound the weight re-reads
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int S = n <= 5 * sms ? 1 : n <= 10 * sms ? 2 : 4;
    cudaStream_t s = (cudaStream_t)stream;
    switch (S) {
        case 1: return launch_c2c<1>(p, s);
        case 2: return launch_c2c<2>(p, s);
        case 3: return launch_c2c<3>(p, s);
        default: return launch_c2c<4>(p, s);
    }
}
