// What every extractor stores per keypoint besides the descriptor: the cv::KeyPoint itself and the Frame fields
// derived from it (row N1: Core/Frame.cpp:75-117,231-241,251-281) -- undistorted position, 64 x 48 grid cell,
// back-projected 3D point.  Called by one lane per keypoint.
#pragma once
#include "common.cuh"

namespace rgbdfe {

__device__ __forceinline__ void store_frame_keypoint(const rgbdfe_frame& F, int gi, const rgbdfe_keypoint& kp, const FramePrep& prep,
    const void* depth_img, int depth_stride, int unproject)
{
    const float px = kp.x, py = kp.y;
    F.kps[gi] = kp;
    F.outlier[gi] = 0;
    // Frame::undistortKeyPoints: cv::undistortPoints(pts, K, distCoef, noArray(), K) -- 5 fixed-point iterations of the
    // inverse distortion in double, then back through K, narrowed to float (OpenCV undistort.dispatch.cpp)
    float ux = px, uy = py;
    if (prep.undistort) {
        double x = (px - prep.cx) * prep.ifx, y = (py - prep.cy) * prep.ify; // float coordinates widen exactly
        const double x0 = x, y0 = y;
        for (int it = 0; it < 5; it++) {
            const double r2 = x * x + y * y;
            const double icdist = 1.0 / (1.0 + ((prep.k[4] * r2 + prep.k[1]) * r2 + prep.k[0]) * r2); // k4..k6 are zero: numerator 1
            if (icdist < 0) {
                x = x0, y = y0;
                break;
            }
            const double dx = 2.0 * prep.k[2] * x * y + prep.k[3] * (r2 + 2.0 * x * x);
            const double dy = prep.k[2] * (r2 + 2.0 * y * y) + 2.0 * prep.k[3] * x * y;
            x = (x0 - dx) * icdist;
            y = (y0 - dy) * icdist;
        }
        ux = (float)(prep.fx * x + prep.cx);
        uy = (float)(prep.fy * y + prep.cy);
    }
    F.pt_un[gi] = make_float2(ux, uy);
    // Frame::posInGrid on the undistorted position
    {
        const int gx = (int)roundf(__fmul_rn(__fsub_rn(ux, prep.minX), prep.gridWInv));
        const int gy = (int)roundf(__fmul_rn(__fsub_rn(uy, prep.minY), prep.gridHInv));
        F.cell[gi] = (gx < 0 || gx >= 64 || gy < 0 || gy >= 48) ? -1 : gx * 48 + gy;
    }
    if (unproject) {
        // Frame::uprojectCamera: depth at the raw position, back-projection of the undistorted one.  The depth image is
        // float metres, or the raw 16-bit image: then cv::Mat::convertTo(CV_32F, factor) happens here
        const size_t di = (size_t)(int)py * depth_stride + (int)px;
        const float z = prep.depth16_factor > 0.0f ? __fmul_rn((float)static_cast<const unsigned short*>(depth_img)[di], prep.depth16_factor)
                                                   : static_cast<const float*>(depth_img)[di];
        rgbdfe_point3f P = { 0.f, 0.f, 0.f };
        if (z > 0) {
            P.x = __fmul_rn(__fmul_rn(__fsub_rn(ux, prep.cxf), z), prep.invfx);
            P.y = __fmul_rn(__fmul_rn(__fsub_rn(uy, prep.cyf), z), prep.invfy);
            P.z = z;
        }
        F.xyz[gi] = P;
    }
}

} // namespace rgbdfe
