// csrc/k_undistort.cu — Frame::UndistortKeyPoints (reference src/Frame.cc:403-433): cv::undistortPoints(mat, mat, mK, mDistCoef,
// cv::Mat(), mK) on the N x 2 keypoint coordinates, then mvKeysUn[i] = mvKeys[i] with the corrected pt.  SURVEY.md §8f rank 4.
//
// OpenCV's routine works in double: normalise with 1/fx, 1/fy, five fixed-point iterations of the inverse Brown-Conrady model
// (k1 k2 p1 p2 k3; the default TermCriteria is COUNT = 5), re-project with P = K, round once to float.  One thread per keypoint,
// individually rounded double operations in OpenCV's order (the library is built with --fmad=false), so the result equals
// cv2.undistortPoints bit for bit (tests/test_undistort.py).  Latency-sized work: 2000 points.
#include "orbx_internal.h"

namespace orbx {

__global__ void __launch_bounds__(128) k_undistort(const orbx_keypoint* __restrict__ in, int n, double fx, double fy, double cx, double cy,
                                                    double k1, double k2, double p1, double p2, double k3, orbx_keypoint* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    orbx_keypoint kp = in[i];
    const double ifx = 1. / fx, ify = 1. / fy;
    const double u = (double)kp.x, v = (double)kp.y;
    double x = (u - cx) * ifx, y = (v - cy) * ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; ++j) {
        const double r2 = x * x + y * y;
        // k[5..7] (rational model) are zero: the numerator is (1 + ((0*r2 + 0)*r2 + 0)*r2) = 1 exactly
        const double icdist = 1. / (1. + ((k3 * r2 + k2) * r2 + k1) * r2);
        if (icdist < 0) {  // OpenCV >= 4: a negative factor means the model left its domain; keep the normalised input
            x = (u - cx) * ifx;
            y = (v - cy) * ify;
            break;
        }
        const double deltaX = 2 * p1 * x * y + p2 * (r2 + 2 * x * x);  // thin-prism terms k[8..11] are zero
        const double deltaY = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y;
        x = (x0 - deltaX) * icdist;
        y = (y0 - deltaY) * icdist;
    }
    const double xx = fx * x + 0. * y + cx, yy = 0. * x + fy * y + cy, ww = 1. / (0. * x + 0. * y + 1.);
    kp.x = (float)(xx * ww);
    kp.y = (float)(yy * ww);
    out[i] = kp;
}

int launch_undistort(const orbx_keypoint* in, int n, const float K[4], const float* dist, int ndist, orbx_keypoint* out, cudaStream_t s) {
    if (n <= 0) return ORBX_OK;
    const double k[5] = {ndist > 0 ? dist[0] : 0., ndist > 1 ? dist[1] : 0., ndist > 2 ? dist[2] : 0., ndist > 3 ? dist[3] : 0.,
                         ndist > 4 ? dist[4] : 0.};
    k_undistort<<<(n + 127) / 128, 128, 0, s>>>(in, n, (double)K[0], (double)K[1], (double)K[2], (double)K[3], k[0], k[1], k[2], k[3], k[4],
                                                out);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
