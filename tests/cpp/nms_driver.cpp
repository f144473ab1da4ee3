// nms_driver.cpp -- prints the indices cv_dnn::NMSBoxes keeps for the boxes read from stdin.
//   default build:        include/nms.hpp (this repo)
//   -DSBM_NMS_REFERENCE:  the reference's own nms.hpp (compiled against oracle/shim; test infrastructure only)
// stdin: n score_threshold nms_threshold eta top_k, then n lines "x y w h score".
#ifdef SBM_NMS_REFERENCE
#include "nms.hpp" // resolved to $(REFERENCE)/nms.hpp by the include path the test passes
#else
#include "../../include/nms.hpp"
#endif
#include <cstdio>

int main()
{
    int n, top_k;
    float st, nt, eta;
    if (scanf("%d %f %f %f %d", &n, &st, &nt, &eta, &top_k) != 5)
        return 2;
    std::vector<cv::Rect> boxes(n);
    std::vector<float> scores(n);
    for (int i = 0; i < n; ++i)
        if (scanf("%d %d %d %d %f", &boxes[i].x, &boxes[i].y, &boxes[i].width, &boxes[i].height, &scores[i]) != 5)
            return 2;
    std::vector<int> keep;
    cv_dnn::NMSBoxes(boxes, scores, st, nt, keep, eta, top_k);
    for (int k : keep)
        printf("%d\n", k);
    return 0;
}
