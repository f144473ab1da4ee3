// imgproc_compat.hpp -- the three host-side OpenCV imgproc calls of the TRAINING helpers, for builds without
// OpenCV: cv::resize / cv::rotate as used by shapeInfo_producer::transform (/root/reference/line2Dup.h:379-405)
// and cv::erode as used on the object mask by extractTemplate (/root/reference/line2Dup.cpp:458), plus the
// nearest-neighbour mask halving of ColorGradientPyramid::pyrDown (:439).  8-bit images only.
// Nothing of the match() hot path is here: blur / Sobel / phase / pyrDown of the image run on the GPU.
//
// Semantics follow OpenCV 4.x: INTER_LINEAR for 8U uses 11-bit fixed-point coefficients computed in float from
// the source coordinate (dst + 0.5) * scale - 0.5, horizontal pass first, vertical pass with the
// (b * (row >> 4)) >> 16 form and a final (+2) >> 2 rounding.
#ifndef SBM_IMGPROC_COMPAT_HPP
#define SBM_IMGPROC_COMPAT_HPP

#include "cv_compat.hpp"

namespace cv
{
enum BorderTypes { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4, BORDER_DEFAULT = 4 };
enum InterpolationFlags { INTER_NEAREST = 0, INTER_LINEAR = 1 };
enum RotateFlags { ROTATE_90_CLOCKWISE = 0, ROTATE_180 = 1, ROTATE_90_COUNTERCLOCKWISE = 2 };

namespace compat_detail
{
// per-axis bilinear table: source index of the left/top tap and the two 11-bit weights
struct LinearAxis
{
    std::vector<int> idx;
    std::vector<short> w0, w1;
    LinearAxis(int n_dst, int n_src, double scale, bool clamp_in_table) : idx(n_dst), w0(n_dst), w1(n_dst)
    {
        for (int d = 0; d < n_dst; ++d)
        {
            float pos = (float)((d + 0.5) * scale - 0.5);
            int i = (int)std::floor(pos);
            float frac = pos - i;
            if (clamp_in_table) // x axis: taps are clamped and the fraction dropped at the borders
            {
                if (i < 0)
                    i = 0, frac = 0.f;
                if (i >= n_src - 1)
                    i = n_src - 1, frac = 0.f;
            }
            idx[d] = i;
            w0[d] = saturate_cast<short>((1.f - frac) * 2048);
            w1[d] = saturate_cast<short>(frac * 2048);
        }
    }
};
inline int clampIndex(int v, int n) { return v < 0 ? 0 : (v >= n ? n - 1 : v); }
} // namespace compat_detail

inline void resize(const Mat &src, Mat &dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR)
{
    CV_Assert(!src.empty() && src.depth() == CV_8U);
    const int sh = src.rows, sw = src.cols, cn = src.channels();
    int dw = dsize.width, dh = dsize.height;
    double sx, sy;
    if (dw > 0 && dh > 0)
        sx = (double)sw / dw, sy = (double)sh / dh;
    else
    {
        CV_Assert(fx > 0 && fy > 0);
        dw = (int)lrint(sw * fx), dh = (int)lrint(sh * fy);
        sx = 1.0 / fx, sy = 1.0 / fy; // the requested factor, not the size ratio
    }
    CV_Assert(dw > 0 && dh > 0);
    Mat out(dh, dw, src.type());
    if (interpolation == INTER_NEAREST)
    {
        for (int y = 0; y < dh; ++y)
        {
            const uchar *srow = src.ptr(std::min((int)std::floor(y * sy), sh - 1));
            uchar *drow = out.ptr(y);
            for (int x = 0; x < dw; ++x)
                memcpy(drow + (size_t)x * cn, srow + (size_t)std::min((int)std::floor(x * sx), sw - 1) * cn, cn);
        }
    }
    else
    {
        CV_Assert(interpolation == INTER_LINEAR);
        if (dw == sw && dh == sh)
            src.copyTo(out);
        else
        {
            const compat_detail::LinearAxis ax(dw, sw, sx, true), ay(dh, sh, sy, false);
            std::vector<int> top((size_t)dw * cn), bot((size_t)dw * cn);
            for (int y = 0; y < dh; ++y)
            {
                const uchar *r0 = src.ptr(compat_detail::clampIndex(ay.idx[y], sh));
                const uchar *r1 = src.ptr(compat_detail::clampIndex(ay.idx[y] + 1, sh));
                for (int x = 0; x < dw; ++x)
                {
                    const int i0 = ax.idx[x] * cn, i1 = std::min(ax.idx[x] + 1, sw - 1) * cn;
                    for (int c = 0; c < cn; ++c)
                    {
                        top[(size_t)x * cn + c] = r0[i0 + c] * ax.w0[x] + r0[i1 + c] * ax.w1[x];
                        bot[(size_t)x * cn + c] = r1[i0 + c] * ax.w0[x] + r1[i1 + c] * ax.w1[x];
                    }
                }
                uchar *drow = out.ptr(y);
                const int b0 = ay.w0[y], b1 = ay.w1[y];
                for (int k = 0; k < dw * cn; ++k)
                    drow[k] = (uchar)((((b0 * (top[k] >> 4)) >> 16) + ((b1 * (bot[k] >> 4)) >> 16) + 2) >> 2);
            }
        }
    }
    dst = out;
}

inline void rotate(const Mat &src, Mat &dst, int rotateCode)
{
    CV_Assert(!src.empty());
    const int h = src.rows, w = src.cols;
    const size_t es = src.elemSize();
    Mat out = rotateCode == ROTATE_180 ? Mat(h, w, src.type()) : Mat(w, h, src.type());
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x)
        {
            int oy = y, ox = x;
            switch (rotateCode)
            {
            case ROTATE_90_CLOCKWISE: oy = x, ox = h - 1 - y; break;
            case ROTATE_180: oy = h - 1 - y, ox = w - 1 - x; break;
            case ROTATE_90_COUNTERCLOCKWISE: oy = w - 1 - x, ox = y; break;
            default: CV_Error(Error::StsBadArg, "unknown rotate code");
            }
            memcpy(out.ptr(oy) + (size_t)ox * es, src.ptr(y) + (size_t)x * es, es);
        }
    dst = out;
}

// default 3x3 rectangular structuring element, one iteration
inline void erode(const Mat &src, Mat &dst, const Mat &kernel, Point anchor = Point(-1, -1), int iterations = 1,
                  int borderType = BORDER_CONSTANT)
{
    CV_Assert(src.type() == CV_8UC1 && kernel.empty() && iterations == 1);
    (void)anchor;
    const int h = src.rows, w = src.cols;
    Mat out(h, w, CV_8UC1);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x)
        {
            int m = 255;
            for (int j = y - 1; j <= y + 1; ++j)
                for (int i = x - 1; i <= x + 1; ++i)
                {
                    const bool inside = j >= 0 && j < h && i >= 0 && i < w;
                    if (!inside && borderType != BORDER_REPLICATE)
                        continue; // BORDER_CONSTANT for erosion = +inf: outside pixels never lower the minimum
                    m = std::min(m, (int)src.ptr(compat_detail::clampIndex(j, h))[compat_detail::clampIndex(i, w)]);
                }
            out.ptr(y)[x] = (uchar)m;
        }
    dst = out;
}
} // namespace cv

#endif // SBM_IMGPROC_COMPAT_HPP
