// Input-heatmap producer: the reference builds the network's input heatmaps
// from 2-D poses on the CPU, one numpy Gaussian per (person, joint)
// (lib/dataset/JointsDataset.py:368-447 generate_input_heatmap, :265-279 compute_human_scale; used for
// the 'pred' and 'gt' heatmap sources, :207-258), then ships 4*J*H*W bytes per view to the GPU.  Here
// the 2-D poses (8*N*J bytes per view) are what crosses PCIe and the maps are splatted in HBM.
//
// Arithmetic follows the reference as numpy 2.x evaluates it: everything up to the exponential is
// float64 (the float32 arange meets np.float64 scalars), the value is rounded to float32 when it is
// stored, persons are combined with max and the final clip(0,1) is the identity on (0,1].
#include "fvp_common.cuh"

namespace fvp {

struct SplatParams {
    const float *joints;       // [M][N][J][2] network-input image pixels (after the resize affine)
    const unsigned char *vis;  // [M][N][J] or null (all visible)
    const int *count;          // [M] people per map set, or null (N everywhere)
    unsigned *out;             // [M][J][H][W], zero-filled by the caller
    int N, J, H, W;
    double fsx, fsy, sigma;    // feat_stride = image_size / heatmap_size
};

__global__ void __launch_bounds__(128)
splat_heatmaps_kernel(const __grid_constant__ SplatParams p) {
    const int j = blockIdx.x, n = blockIdx.y, m = blockIdx.z;
    if (p.count != nullptr && n >= p.count[m]) return;
    const float *pose = p.joints + (((size_t)m * p.N + n) * p.J) * 2;
    if (p.vis != nullptr && p.vis[((size_t)m * p.N + n) * p.J + j] == 0) return;

    // compute_human_scale(pose / feat_stride, ones): every joint counts, visible or not
    double minx = 0.0, maxx = 0.0, miny = 0.0, maxy = 0.0;
    for (int k = 0; k < p.J; ++k) {
        const double x = (double)pose[2 * k] / p.fsx, y = (double)pose[2 * k + 1] / p.fsy;
        if (k == 0) { minx = maxx = x; miny = maxy = y; }
        minx = fmin(minx, x); maxx = fmax(maxx, x);
        miny = fmin(miny, y); maxy = fmax(maxy, y);
    }
    const double ext = fmax(maxy - miny, maxx - minx);
    const double human_scale = 2.0 * fmin(fmax(ext * ext, 0.25 * 96.0 * 96.0), 4.0 * 96.0 * 96.0);
    const double cur_sigma = p.sigma * sqrt(human_scale / (96.0 * 96.0));
    const double tmp_size = cur_sigma * 3.0;

    // int() truncates toward zero
    const int mu_x = (int)((double)pose[2 * j] / p.fsx), mu_y = (int)((double)pose[2 * j + 1] / p.fsy);
    const int ul0 = (int)(mu_x - tmp_size), ul1 = (int)(mu_y - tmp_size);
    const int br0 = (int)(mu_x + tmp_size + 1.0), br1 = (int)(mu_y + tmp_size + 1.0);
    if (ul0 >= p.W || ul1 >= p.H || br0 < 0 || br1 < 0) return;

    const double size = 2.0 * tmp_size + 1.0;
    const int len = (int)ceil(size);           // np.arange(0, size, 1) has ceil(size) entries
    const double c0 = floor(size / 2.0);       // x0 = y0 = size // 2
    const double denom = 2.0 * (cur_sigma * cur_sigma);
    const int x_lo = max(0, ul0), x_hi = min(min(br0, p.W), ul0 + len);
    const int y_lo = max(0, ul1), y_hi = min(min(br1, p.H), ul1 + len);
    const int nx = x_hi - x_lo, ny = y_hi - y_lo;
    if (nx <= 0 || ny <= 0) return;
    unsigned *plane = p.out + ((size_t)m * p.J + j) * p.H * p.W;
    for (int e = threadIdx.x; e < nx * ny; e += 128) {
        const int py = y_lo + e / nx, px = x_lo + e % nx;
        const double dx = (double)(px - ul0) - c0, dy = (double)(py - ul1) - c0;
        const float g = (float)exp(-(dx * dx + dy * dy) / denom);
        atomicMax(plane + (size_t)py * p.W + px, __float_as_uint(g));  // g >= 0: unsigned order == float order
    }
}

}  // namespace fvp

using namespace fvp;

extern "C" int fvp_splat_heatmaps(const float *joints, const unsigned char *vis, const int32_t *count, int M, int N,
                                  int J, float sigma, int img_w, int img_h, int hm_w, int hm_h, float *out,
                                  void *stream) {
    FVP_REQUIRE(joints && out, "fvp_splat_heatmaps: null pointer");
    FVP_REQUIRE(M >= 0 && N >= 0 && J > 0 && J <= 65535, "fvp_splat_heatmaps: bad dims M=%d N=%d J=%d", M, N, J);
    FVP_REQUIRE(img_w > 0 && img_h > 0 && hm_w > 0 && hm_h > 0 && sigma > 0.0f, "fvp_splat_heatmaps: bad sizes");
    FVP_REQUIRE(M <= 65535 && N <= 65535, "fvp_splat_heatmaps: M or N too large for the launch grid");
    cudaStream_t s = (cudaStream_t)stream;
    const size_t bytes = sizeof(float) * (size_t)M * J * hm_h * hm_w;
    if (bytes == 0) return FVP_OK;
    FVP_CUDA(cudaMemsetAsync(out, 0, bytes, s));
    if (N == 0) return FVP_OK;
    SplatParams p;
    p.joints = joints; p.vis = vis; p.count = count; p.out = reinterpret_cast<unsigned *>(out);
    p.N = N; p.J = J; p.H = hm_h; p.W = hm_w;
    p.fsx = (double)img_w / (double)hm_w; p.fsy = (double)img_h / (double)hm_h; p.sigma = (double)sigma;
    splat_heatmaps_kernel<<<dim3((unsigned)J, (unsigned)N, (unsigned)M), 128, 0, s>>>(p);
    return check_launch("splat_heatmaps_kernel");
}
