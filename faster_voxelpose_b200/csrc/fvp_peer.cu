// z-slab plane combine over peer memory: ONE kernel that reads every rank's partial planes
// straight out of the peers' HBM (NVLink / NVSwitch P2P loads) and writes the full planes locally -- the
// reduce-max of the xy plane and the all-gather of the xz / yz slabs in one pass, instead of one NCCL
// all-reduce plus two all-gathers and their pad / trim copies.
#include <algorithm>

#include "fvp_common.cuh"

namespace fvp {

constexpr int MAX_PEERS = 16;

struct PeerPlanes {
    const unsigned *xy[MAX_PEERS];  // partial xy plane of every rank, [n_xy]
    const unsigned *xz[MAX_PEERS];  // xz slab of every rank, [rows_xz][zs_r]   (null when not wanted)
    const unsigned *yz[MAX_PEERS];  // yz slab of every rank, [rows_yz][zs_r]
    int zstart[MAX_PEERS + 1];      // slab of rank r is z in [zstart[r], zstart[r+1])
    int R;
};

// Peer lines may sit in this SM's L1 from the previous step: every peer load is cache-volatile.
__device__ __forceinline__ uint4 ld_peer4(const unsigned *p) { return __ldcv(reinterpret_cast<const uint4 *>(p)); }

__global__ void __launch_bounds__(256)
planes_combine_peers_kernel(const __grid_constant__ PeerPlanes pp, long n_xy, long rows_xz, long rows_yz, int nz,
                            unsigned *__restrict__ out_xy, unsigned *__restrict__ out_xz,
                            unsigned *__restrict__ out_yz) {
    const long tid = (long)blockIdx.x * blockDim.x + threadIdx.x, nthr = (long)gridDim.x * blockDim.x;
    // xy: max over the ranks.  Values are >= 0 or NaN, so the unsigned order is the float order with NaN on
    // top -- the same rule the projection kernel's atomicMax applies inside one GPU.
    const long n4 = n_xy >> 2;
    for (long i = tid; i < n4; i += nthr) {
        uint4 m = ld_peer4(pp.xy[0] + 4 * i);
        for (int r = 1; r < pp.R; ++r) {
            const uint4 v = ld_peer4(pp.xy[r] + 4 * i);
            m.x = max(m.x, v.x); m.y = max(m.y, v.y); m.z = max(m.z, v.z); m.w = max(m.w, v.w);
        }
        reinterpret_cast<uint4 *>(out_xy)[i] = m;
    }
    for (long i = 4 * n4 + tid; i < n_xy; i += nthr) {
        unsigned m = __ldcv(pp.xy[0] + i);
        for (int r = 1; r < pp.R; ++r) m = max(m, __ldcv(pp.xy[r] + i));
        out_xy[i] = m;
    }
    // xz / yz: row `row`, cell z comes from the rank that owns z
    for (int which = 0; which < 2; ++which) {
        unsigned *out = which ? out_yz : out_xz;
        if (out == nullptr) continue;
        const long rows = which ? rows_yz : rows_xz;
        for (long i = tid; i < rows * nz; i += nthr) {
            const long row = i / nz;
            const int z = (int)(i - row * nz);
            int r = 0;
            while (z >= pp.zstart[r + 1]) ++r;
            const int zs = pp.zstart[r + 1] - pp.zstart[r];
            const unsigned *src = which ? pp.yz[r] : pp.xz[r];
            out[i] = __ldcv(src + row * zs + (z - pp.zstart[r]));
        }
    }
}

// The same combine through the NVSwitch (NVLS): every pointer marked mc is a MULTICAST address of a symmetric
// buffer.  multimem.ld_reduce returns the maximum of the addressed word over all ranks' copies, reduced inside
// the switch; multimem.st writes a word into every rank's copy.  Rank r reduces its 1/R of the xy plane and
// broadcasts it, and broadcasts its own xz / yz slab into place: per rank n_xy/R words in and n_xy/R + slab
// words out, against R * n_xy words in for the pull kernel above.
__device__ __forceinline__ unsigned mc_ld_max(const unsigned *mc) {
    unsigned v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.max.u32 %0, [%1];" : "=r"(v) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st(unsigned *mc, unsigned v) {
    asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}

__global__ void __launch_bounds__(256)
planes_combine_multimem_kernel(const unsigned *mc_xy_partial, unsigned *mc_xy_full, long lo, long hi,
                               const unsigned *__restrict__ xz_slab, unsigned *mc_xz_full, long rows_xz,
                               const unsigned *__restrict__ yz_slab, unsigned *mc_yz_full, long rows_yz, int z0,
                               int zs, int nz) {
    const long tid = (long)blockIdx.x * blockDim.x + threadIdx.x, nthr = (long)gridDim.x * blockDim.x;
    for (long i = lo + tid; i < hi; i += nthr) mc_st(mc_xy_full + i, mc_ld_max(mc_xy_partial + i));
    for (int which = 0; which < 2; ++which) {
        const unsigned *slab = which ? yz_slab : xz_slab;
        unsigned *full = which ? mc_yz_full : mc_xz_full;
        if (slab == nullptr || zs == 0) continue;
        const long rows = which ? rows_yz : rows_xz;
        for (long e = tid; e < rows * zs; e += nthr) {
            const long row = e / zs;
            mc_st(full + row * nz + z0 + (int)(e - row * zs), slab[e]);
        }
    }
}

}  // namespace fvp

using namespace fvp;

extern "C" int fvp_planes_combine_multimem(const void *mc_xy_partial, void *mc_xy_full, const float *xz_slab,
                                           void *mc_xz_full, const float *yz_slab, void *mc_yz_full, int R, int rank,
                                           const int32_t *zstart, long n_xy, long rows_xz, long rows_yz, void *stream) {
    FVP_REQUIRE(mc_xy_partial && mc_xy_full && zstart, "fvp_planes_combine_multimem: null pointer");
    FVP_REQUIRE(R >= 1 && R <= MAX_PEERS && rank >= 0 && rank < R, "fvp_planes_combine_multimem: bad rank %d / R %d", rank, R);
    FVP_REQUIRE(n_xy > 0 && rows_xz >= 0 && rows_yz >= 0, "fvp_planes_combine_multimem: bad sizes");
    FVP_REQUIRE((xz_slab == nullptr) == (mc_xz_full == nullptr) && (yz_slab == nullptr) == (mc_yz_full == nullptr),
                "fvp_planes_combine_multimem: xz / yz need both the local slab and the multicast output");
    const int z0 = zstart[rank], zs = zstart[rank + 1] - zstart[rank], nz = zstart[R];
    FVP_REQUIRE(zstart[0] == 0 && zs >= 0 && nz > 0, "fvp_planes_combine_multimem: bad zstart");
    // this rank's share of the xy plane, in whole 128-byte lines
    const long per = ((n_xy + R - 1) / R + 31) / 32 * 32;
    const long lo = std::min(n_xy, (long)rank * per), hi = std::min(n_xy, lo + per);
    const long work = std::max(hi - lo, std::max(xz_slab ? rows_xz * zs : 0L, yz_slab ? rows_yz * zs : 0L));
    const int blocks = (int)std::min<long>((work + 255) / 256, 148L * 8);
    planes_combine_multimem_kernel<<<std::max(blocks, 1), 256, 0, (cudaStream_t)stream>>>(
        static_cast<const unsigned *>(mc_xy_partial), static_cast<unsigned *>(mc_xy_full), lo, hi,
        reinterpret_cast<const unsigned *>(xz_slab), static_cast<unsigned *>(mc_xz_full), rows_xz,
        reinterpret_cast<const unsigned *>(yz_slab), static_cast<unsigned *>(mc_yz_full), rows_yz, z0, zs, nz);
    return check_launch("planes_combine_multimem_kernel");
}

extern "C" int fvp_planes_combine_peers(const void *const *xy_peers, const void *const *xz_peers,
                                        const void *const *yz_peers, int R, const int32_t *zstart, long n_xy,
                                        long rows_xz, long rows_yz, float *out_xy, float *out_xz, float *out_yz,
                                        void *stream) {
    FVP_REQUIRE(xy_peers && zstart && out_xy, "fvp_planes_combine_peers: null pointer");
    FVP_REQUIRE(R >= 1 && R <= MAX_PEERS, "fvp_planes_combine_peers: R must be in 1..%d", MAX_PEERS);
    FVP_REQUIRE(n_xy > 0 && rows_xz >= 0 && rows_yz >= 0, "fvp_planes_combine_peers: bad sizes");
    FVP_REQUIRE((out_xz == nullptr) == (xz_peers == nullptr) && (out_yz == nullptr) == (yz_peers == nullptr),
                "fvp_planes_combine_peers: xz / yz need both the peer slabs and the output");
    FVP_REQUIRE((reinterpret_cast<uintptr_t>(out_xy) & 15) == 0, "fvp_planes_combine_peers: out_xy must be 16-byte aligned");
    PeerPlanes pp;
    pp.R = R;
    FVP_REQUIRE(zstart[0] == 0, "fvp_planes_combine_peers: zstart[0] must be 0");
    for (int r = 0; r <= R; ++r) {
        pp.zstart[r] = zstart[r];
        FVP_REQUIRE(r == 0 || zstart[r] >= zstart[r - 1], "fvp_planes_combine_peers: zstart must not decrease");
    }
    const int nz = zstart[R];
    FVP_REQUIRE(nz > 0, "fvp_planes_combine_peers: empty z range");
    for (int r = 0; r < R; ++r) {
        const bool owns = zstart[r + 1] > zstart[r];
        pp.xy[r] = static_cast<const unsigned *>(xy_peers[r]);
        pp.xz[r] = xz_peers ? static_cast<const unsigned *>(xz_peers[r]) : nullptr;
        pp.yz[r] = yz_peers ? static_cast<const unsigned *>(yz_peers[r]) : nullptr;
        FVP_REQUIRE(pp.xy[r] != nullptr && (reinterpret_cast<uintptr_t>(pp.xy[r]) & 15) == 0,
                    "fvp_planes_combine_peers: xy of rank %d is null or not 16-byte aligned", r);
        FVP_REQUIRE(!owns || ((!out_xz || pp.xz[r]) && (!out_yz || pp.yz[r])),
                    "fvp_planes_combine_peers: rank %d owns z cells but its slab pointer is null", r);
    }
    for (int r = R; r < MAX_PEERS; ++r) pp.xy[r] = pp.xz[r] = pp.yz[r] = nullptr;
    const long work = std::max(n_xy / 4, std::max(out_xz ? rows_xz * nz : 0L, out_yz ? rows_yz * nz : 0L));
    const int blocks = (int)std::min<long>((work + 255) / 256, 148L * 8);
    planes_combine_peers_kernel<<<std::max(blocks, 1), 256, 0, (cudaStream_t)stream>>>(
        pp, n_xy, rows_xz, rows_yz, nz, reinterpret_cast<unsigned *>(out_xy), reinterpret_cast<unsigned *>(out_xz),
        reinterpret_cast<unsigned *>(out_yz));
    return check_launch("planes_combine_peers_kernel");
}
