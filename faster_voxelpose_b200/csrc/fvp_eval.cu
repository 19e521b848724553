// Accuracy evaluation, the numeric part (reference lib/dataset/panoptic.py:224-246: every predicted
// pose is compared with every ground-truth person of its frame -- mean distance over the person's visible joints -- and
// keeps its closest person).  The reference does this per frame on the host while the poses sit on the device; here one
// kernel does it for the whole evaluation set so that only [F][P] errors and indices travel.  Double precision, as numpy
// computes it (float32 poses minus float64 ground truth).  The AP / recall bookkeeping over the sorted scores stays on the
// host (evaluation.py): it is sequential by definition (greedy claims in score order).
#include "fvp_common.cuh"

namespace fvp {

// One warp per (frame, predicted pose); lanes stride over the joints of one ground-truth person at a time.
__global__ void __launch_bounds__(128)
pose_match_kernel(const float *__restrict__ pred, const double *__restrict__ gt, const float *__restrict__ vis,
                  const int *__restrict__ ngt, int F, int P, int J, int G, double *__restrict__ err, int *__restrict__ best) {
    const int lane = threadIdx.x & 31;
    const long item = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (item >= (long)F * P) return;
    const int f = (int)(item / P);
    const float *pp = pred + item * J * 5;
    const int n = min(ngt[f], G);
    double best_err = __longlong_as_double(0x7ff0000000000000LL);   // +inf
    int best_g = -1;
    const bool valid = pp[3] >= 0.0f && n > 0;                      // match flag of the pose (joint 0, component 3)
    if (valid) {
        for (int g = 0; g < n; ++g) {
            const double *gg = gt + ((size_t)f * G + g) * J * 3;
            const float *vv = vis + ((size_t)f * G + g) * J;
            double sum = 0.0, cnt = 0.0;
            for (int j = lane; j < J; j += 32) {
                if (vv[j] > 0.1f) {
                    const double dx = (double)pp[j * 5 + 0] - gg[j * 3 + 0], dy = (double)pp[j * 5 + 1] - gg[j * 3 + 1],
                                 dz = (double)pp[j * 5 + 2] - gg[j * 3 + 2];
                    sum += sqrt(dx * dx + dy * dy + dz * dz);
                    cnt += 1.0;
                }
            }
            for (int o = 16; o; o >>= 1) {
                sum += __shfl_xor_sync(0xffffffffu, sum, o);
                cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
            }
            const double mean = cnt > 0.0 ? sum / cnt : __longlong_as_double(0x7ff0000000000000LL);   // nobody visible: never closest
            if (mean < best_err || best_g < 0) { best_err = mean; best_g = g; }   // first minimum, as numpy's argmin
        }
    }
    if (lane == 0) {
        err[item] = best_err;
        best[item] = best_g;
    }
}

}  // namespace fvp

using namespace fvp;

extern "C" int fvp_pose_match(const float *pred, const double *gt, const float *vis, const int32_t *ngt, int F, int P, int J,
                              int G, double *err, int32_t *best, void *stream) {
    FVP_REQUIRE(pred && gt && vis && ngt && err && best, "fvp_pose_match: null pointer");
    FVP_REQUIRE(F >= 0 && P >= 1 && J >= 1 && G >= 1, "fvp_pose_match: bad dims F=%d P=%d J=%d G=%d", F, P, J, G);
    if (F == 0) return FVP_OK;
    const long items = (long)F * P;
    pose_match_kernel<<<(unsigned)((items + 3) / 4), 128, 0, (cudaStream_t)stream>>>(pred, gt, vis, ngt, F, P, J, G, err, best);
    return check_launch("pose_match_kernel");
}
