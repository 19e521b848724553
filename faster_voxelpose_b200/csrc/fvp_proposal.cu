// Proposal-layer kernels: 3x3 max-equality NMS + top-k, the
// bbox / z-column gathers at the top-k cells, and the 7-float proposal records.
#include <cfloat>
#include <mutex>

#include "fvp_common.cuh"

namespace fvp {

// Total order used for the top-k: larger value first, then smaller flat index.
// NaN sorts above +inf (torch.topk treats NaN as the largest value); -0 == +0.
__device__ __forceinline__ unsigned long long topk_key(float v, int idx) {
    unsigned u;
    if (v != v) {
        u = 0xffffffffu;
    } else {
        u = __float_as_uint(v + 0.0f);
        u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    }
    return ((unsigned long long)u << 32) | (unsigned long long)(0xffffffffu - (unsigned)idx);
}

// One CTA per frame.  Pass 1 writes keep*x (proposal.py:49-52) into the workspace -- shared memory when the map fits
// (`in_smem`), else the caller's scratch; pass 2 extracts the P largest entries (P is MAX_PEOPLE, 5..32) by repeated block-wide
// arg-max over one cached key per thread: only the thread that owned the extracted cell rescans its cells, so a round is a
// shuffle tree, eight shared-memory words and one barrier (it was a rescan of the whole map from global memory: 61 us for the
// bench's 32 maps of 80 x 80, now 12).
__global__ void __launch_bounds__(256)
nms_topk_kernel(const float *__restrict__ hm, int X, int Y, int P, float *__restrict__ scratch, int in_smem,
                float *__restrict__ vals, long long *__restrict__ idx2, long long *__restrict__ flat) {
    extern __shared__ float s_kept[];
    __shared__ unsigned long long s_best[2][8];
    const int b = blockIdx.x, n = X * Y;
    const float *m = hm + (size_t)b * n;
    float *kept = in_smem ? s_kept : scratch + (size_t)b * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int x = i / Y, y = i % Y;
        float mx = -INFINITY;
        for (int dx = -1; dx <= 1; ++dx)
            for (int dy = -1; dy <= 1; ++dy) {
                const int xx = x + dx, yy = y + dy;
                if (xx < 0 || yy < 0 || xx >= X || yy >= Y) continue;  // -inf padding
                const float v = __ldg(m + xx * Y + yy);
                if (v > mx || v != v) mx = v;                          // max_pool2d keeps NaN
            }
        const float c = __ldg(m + i);
        kept[i] = __fmul_rn(c == mx ? 1.0f : 0.0f, c);                 // keep.float() * inputs
    }
    // every thread reads back only what it wrote itself (cells threadIdx.x + k * blockDim.x): no barrier needed
    auto local_best = [&]() {
        unsigned long long best = 0ull;
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const float v = kept[i];
            if (v == -INFINITY) continue;  // consumed (or a genuine -inf, which can never rank)
            const unsigned long long key = topk_key(v, i);
            best = key > best ? key : best;
        }
        return best;
    };
    unsigned long long mine = local_best();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (int)(blockDim.x >> 5);
    for (int p = 0; p < P; ++p) {
        unsigned long long best = mine;
        for (int o = 16; o; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other > best ? other : best;
        }
        if (lane == 0) s_best[p & 1][warp] = best;
        __syncthreads();   // the other buffer was last read before the previous round's barrier
        best = s_best[p & 1][0];
        for (int w = 1; w < nwarp; ++w) best = s_best[p & 1][w] > best ? s_best[p & 1][w] : best;
        const int idx = best ? (int)(0xffffffffu - (unsigned)(best & 0xffffffffull)) : 0;
        if ((int)threadIdx.x == idx % (int)blockDim.x) {   // the owner of the cell reports it and retires it
            const size_t o = (size_t)b * P + p;
            vals[o] = kept[idx];
            flat[o] = idx;
            idx2[2 * o] = idx / X;      // get_index2D divides by shape[1] == X
            idx2[2 * o + 1] = idx % X;
            kept[idx] = -INFINITY;
            mine = local_best();
        }
    }
}

// bbox[B][2][X][Y] -> match_bbox[B][P][2];  cubes[B][J][X][Y][Z] -> cols[B][P][J][Z]
__global__ void gather_topk_kernel(const float *__restrict__ bbox, const float *__restrict__ cubes,
                                   const long long *__restrict__ flat, int B, int P, int J, int XY, int Z,
                                   float *__restrict__ match_bbox, float *__restrict__ cols) {
    const int per = J * Z + 2;
    const long total = (long)B * P * per;
    for (long t = (long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long)gridDim.x * blockDim.x) {
        const int e = (int)(t % per);
        const long bp = t / per;
        const int b = (int)(bp / P);
        const long cell = flat[bp];
        if (e < J * Z) {
            if (cols != nullptr && cubes != nullptr) {
                const int j = e / Z, z = e % Z;
                cols[bp * J * Z + e] = __ldg(cubes + (((size_t)b * J + j) * XY + cell) * Z + z);
            }
        } else if (match_bbox != nullptr && bbox != nullptr) {
            const int a = e - J * Z;
            match_bbox[bp * 2 + a] = __ldg(bbox + ((size_t)b * 2 + a) * XY + cell);
        }
    }
}

struct RecordConsts {
    float scale[3], bias[3], min_score;
};

__global__ void proposal_records_kernel(const long long *__restrict__ idx2, const float *__restrict__ conf2d,
                                        const float *__restrict__ hm1d, const float *__restrict__ bbox, int BP, int Z,
                                        const __grid_constant__ RecordConsts c, float *__restrict__ rec) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= BP) return;
    const float *h = hm1d + (size_t)i * Z;
    int best = 0;
    float bv = h[0];
    for (int z = 1; z < Z; ++z) {
        const float v = h[z];
        if (v > bv || (v != v && bv == bv)) { bv = v; best = z; }  // topk(1): first maximum, NaN largest
    }
    const float conf = __fmul_rn(conf2d[i], bv);
    float *r = rec + (size_t)i * 7;
    r[0] = __fadd_rn(__fmul_rn((float)idx2[2 * i], c.scale[0]), c.bias[0]);
    r[1] = __fadd_rn(__fmul_rn((float)idx2[2 * i + 1], c.scale[1]), c.bias[1]);
    r[2] = __fadd_rn(__fmul_rn((float)best, c.scale[2]), c.bias[2]);
    r[3] = __fsub_rn(conf > c.min_score ? 1.0f : 0.0f, 1.0f);
    r[4] = conf;
    r[5] = bbox[2 * i];
    r[6] = bbox[2 * i + 1];
}

}  // namespace fvp

using namespace fvp;

extern "C" int fvp_nms_topk(const float *hm, int B, int X, int Y, int P, float *scratch, float *vals, int64_t *idx2,
                            int64_t *flat, void *stream) {
    FVP_REQUIRE(hm && scratch && vals && idx2 && flat, "fvp_nms_topk: null pointer");
    FVP_REQUIRE(B > 0 && X > 0 && Y > 0, "fvp_nms_topk: bad dims");
    FVP_REQUIRE(P > 0 && P <= X * Y, "fvp_nms_topk: need 0 < P <= X*Y");
    const size_t smem = sizeof(float) * (size_t)X * Y;
    const bool in_smem = smem <= 40 * 1024;   // else the caller's scratch is the workspace
    nms_topk_kernel<<<B, 256, in_smem ? smem : 0, (cudaStream_t)stream>>>(hm, X, Y, P, scratch, in_smem ? 1 : 0, vals,
                                                                          reinterpret_cast<long long *>(idx2),
                                                                          reinterpret_cast<long long *>(flat));
    return check_launch("nms_topk_kernel");
}

extern "C" int fvp_gather_topk(const float *bbox, const float *cubes, const int64_t *flat, int B, int P, int J, int X,
                               int Y, int Z, float *match_bbox, float *cols, void *stream) {
    FVP_REQUIRE(flat != nullptr, "fvp_gather_topk: null flat index");
    FVP_REQUIRE(B > 0 && P > 0 && J > 0 && X > 0 && Y > 0 && Z > 0, "fvp_gather_topk: bad dims");
    FVP_REQUIRE((bbox != nullptr) == (match_bbox != nullptr), "fvp_gather_topk: bbox in/out must both be set or null");
    FVP_REQUIRE((cubes != nullptr) == (cols != nullptr), "fvp_gather_topk: cubes/cols must both be set or null");
    const long total = (long)B * P * (J * Z + 2);
    const int blocks = (int)std::min<long>((total + 255) / 256, 148L * 8);
    gather_topk_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(bbox, cubes, reinterpret_cast<const long long *>(flat), B,
                                                               P, J, X * Y, Z, match_bbox, cols);
    return check_launch("gather_topk_kernel");
}

extern "C" int fvp_proposal_records(const int64_t *idx2, const float *conf2d, const float *hm1d, const float *match_bbox,
                                    int BP, int Z, const float scale[3], const float bias[3], float min_score, float *rec,
                                    void *stream) {
    FVP_REQUIRE(idx2 && conf2d && hm1d && match_bbox && rec && scale && bias, "fvp_proposal_records: null pointer");
    FVP_REQUIRE(BP > 0 && Z > 0, "fvp_proposal_records: bad dims");
    RecordConsts c;
    for (int a = 0; a < 3; ++a) { c.scale[a] = scale[a]; c.bias[a] = bias[a]; }
    c.min_score = min_score;
    proposal_records_kernel<<<ceil_div(BP, 128), 128, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const long long *>(idx2), conf2d, hm1d, match_bbox, BP, Z, c, rec);
    return check_launch("proposal_records_kernel");
}

// ------------------------------------------------------------------------------------------------------------
// fvp_nms_topk_static: the same kernel for callers that cannot allocate (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  The workspace of fvp_nms_topk lives in a file-scope device array of
// FVP_NMS_STATIC_MAX_CELLS floats (4 MiB, held by every GPU that loads the library until exit).  On the host the
// array's NAME is only a stand-in for the symbol, not a device address: the address is looked up once per device
// with cudaGetSymbolAddress while that device is current, and that pointer is what the kernel receives.
// ------------------------------------------------------------------------------------------------------------

static __device__ __align__(256) float g_nms_static_scratch[FVP_NMS_STATIC_MAX_CELLS];

namespace {

constexpr int MAX_DEVICES = 64;
std::mutex g_nms_static_mutex;
float *g_nms_static_addr[MAX_DEVICES] = {};  // per device: device address of g_nms_static_scratch (null = not looked up)

// No allocation: cudaGetDevice + (first call per device) cudaGetSymbolAddress.
int nms_static_scratch(float **out) {
    int dev = -1;
    FVP_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return fvp::fail(FVP_E_ARG, "fvp_nms_topk_static: device ordinal %d not in [0,%d)", dev, MAX_DEVICES);
    std::lock_guard<std::mutex> lock(g_nms_static_mutex);
    if (g_nms_static_addr[dev] == nullptr) {
        void *p = nullptr;
        FVP_CUDA(cudaGetSymbolAddress(&p, g_nms_static_scratch));
        if (p == nullptr) return fvp::fail(FVP_E_CUDA, "fvp_nms_topk_static: cudaGetSymbolAddress returned a null address on device %d", dev);
        g_nms_static_addr[dev] = static_cast<float *>(p);
    }
    *out = g_nms_static_addr[dev];
    return FVP_OK;
}

}  // namespace

extern "C" int fvp_nms_topk_static(const float *hm, int B, int X, int Y, int P, float *vals, int64_t *idx2, int64_t *flat,
                                   void *stream) {
    FVP_REQUIRE(hm && vals && idx2 && flat, "fvp_nms_topk_static: null pointer");
    FVP_REQUIRE(B > 0 && X > 0 && Y > 0, "fvp_nms_topk_static: bad dims");
    FVP_REQUIRE((long long)B * X * Y <= (long long)FVP_NMS_STATIC_MAX_CELLS,
                "fvp_nms_topk_static: B*X*Y = %lld cells exceeds the built-in workspace of %d cells "
                "(FVP_NMS_STATIC_MAX_CELLS); use fvp_nms_topk with caller-provided scratch",
                (long long)B * X * Y, (int)FVP_NMS_STATIC_MAX_CELLS);
    FVP_REQUIRE(P > 0 && P <= X * Y, "fvp_nms_topk_static: need 0 < P <= X*Y");
    float *scratch = nullptr;
    const int rc = nms_static_scratch(&scratch);
    if (rc != FVP_OK) return rc;
    // Nothing to clear: pass 1 of the kernel overwrites every cell [b][i] it later reads, whatever the last call left.
    nms_topk_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(hm, X, Y, P, scratch, 0, vals, reinterpret_cast<long long *>(idx2),
                                                       reinterpret_cast<long long *>(flat));
    return check_launch("nms_topk_kernel (static workspace)");
}
