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

}  // namespace fvp

using namespace fvp;

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
