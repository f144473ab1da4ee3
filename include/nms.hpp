// nms.hpp -- the step every caller of Detector::match runs next: greedy non-maximum
// suppression over the matched templates' boxes.  Same entry point and argument meaning as the reference's
// header (/root/reference/nms.hpp:21-95, used at test.cpp:475-491 and test_jabil.cpp:133-149):
//
//     cv_dnn::NMSBoxes(boxes, scores, score_threshold, nms_threshold, indices, eta = 1, top_k = 0);
//
// Host code: match lists are a few hundred boxes, the loop is O(kept * n).
#ifndef SBM_B200_NMS_HPP
#define SBM_B200_NMS_HPP

#if defined(__has_include)
#if __has_include(<opencv2/core/core.hpp>) && !defined(SBM_FORCE_CV_COMPAT)
#include <opencv2/core/core.hpp>
#define SBM_NMS_HAVE_OPENCV 1
#endif
#endif
#ifndef SBM_NMS_HAVE_OPENCV
#include "sbm/cv_compat.hpp"
#endif

#include <algorithm>
#include <numeric>
#include <vector>

namespace cv_dnn
{
// intersection over union with the reference's arithmetic: Jaccard distance in double, the overlap as
// 1.f - (float)distance (nms.hpp:72-92); two empty boxes overlap completely
inline float boxOverlap(const cv::Rect &a, const cv::Rect &b)
{
    const int area_a = a.area(), area_b = b.area();
    if (area_a + area_b <= 0)
        return 1.f;
    const double inter = (a & b).area();
    const double distance = 1.0 - inter / (area_a + area_b - inter);
    return 1.f - static_cast<float>(distance);
}

// indices of the boxes that survive, best score first.  Boxes scoring <= score_threshold never take part; of
// the rest only the top_k best do (top_k > 0); a box is kept when it overlaps every already kept box by at
// most the current threshold, which decays by eta after each kept box while it is above 0.5 (eta < 1).
inline void NMSBoxes(const std::vector<cv::Rect> &bboxes, const std::vector<float> &scores, const float score_threshold,
                     const float nms_threshold, std::vector<int> &indices, const float eta = 1, const int top_k = 0)
{
    CV_Assert(bboxes.size() == scores.size());
    std::vector<int> order;
    order.reserve(scores.size());
    for (int i = 0; i < (int)scores.size(); ++i)
        if (scores[i] > score_threshold)
            order.push_back(i);
    // stable: equal scores keep their input order (nms.hpp:31-32)
    std::stable_sort(order.begin(), order.end(), [&](int i, int j) { return scores[i] > scores[j]; });
    if (top_k > 0 && top_k < (int)order.size())
        order.resize(top_k);

    float limit = nms_threshold;
    indices.clear();
    for (const int idx : order)
    {
        bool keep = true;
        for (size_t k = 0; k < indices.size() && keep; ++k)
            keep = boxOverlap(bboxes[idx], bboxes[indices[k]]) <= limit;
        if (!keep)
            continue;
        indices.push_back(idx);
        if (eta < 1 && limit > 0.5)
            limit *= eta;
    }
}
} // namespace cv_dnn

#endif // SBM_B200_NMS_HPP
