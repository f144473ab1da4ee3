// line2Dup_c.cpp -- a flat C view of the drop-in C++ classes (include/line2Dup.h) so that Python (ctypes) can drive
// the SAME host code a C++ caller links: Detector::addTemplate / addTemplate_rotate / match / readClasses /
// writeClasses / setGroup and shape_based_matching::shapeInfo_producer.  Built into libline2Dup_b200.so.
// Used by shape_based_matching_b200/line2dup.py (bench.py builds configuration M4's template set through it; the tests
// compare it with the reference's own code).  The array interchange mirrors oracle/ref_driver.cpp on purpose:
// tdesc[n * levels][5] = {width, height, tl_x, tl_y, nf}, feats[sum nf][3] = {x, y, label}.
#include "line2Dup.h"
#include "sbm_capi.h"

#include <cstring>
#include <sstream>

namespace
{
thread_local std::string g_err;

cv::Mat wrap(const uint8_t *img, int w, int h, int c)
{
    cv::Mat m(h, w, CV_MAKETYPE(CV_8U, c));
    memcpy(m.data, img, (size_t)w * h * c);
    return m;
}

std::vector<std::string> splitIds(const char *csv)
{
    std::vector<std::string> ids;
    if (!csv || !*csv)
        return ids;
    std::istringstream is(csv);
    std::string tok;
    while (std::getline(is, tok, ','))
        ids.push_back(tok);
    return ids;
}
} // namespace

#define L2D_TRY try {
#define L2D_CATCH(failval)          \
    }                               \
    catch (const std::exception &e) \
    {                               \
        g_err = e.what();           \
        return failval;             \
    }

extern "C"
{
const char *l2d_last_error() { return g_err.c_str(); }

void *l2d_create(int num_features, const int *T, int nT, float weak, float strong, int device)
{
    L2D_TRY
    auto *d = new line2Dup::Detector(num_features, std::vector<int>(T, T + nT), weak, strong);
    d->setDevice(device);
    return d;
    L2D_CATCH(nullptr)
}
void l2d_destroy(void *h) { delete (line2Dup::Detector *)h; }

int l2d_add_template(void *h, const uint8_t *img, int w, int hh, int c, const uint8_t *mask, const char *class_id, int num_features)
{
    L2D_TRY
    cv::Mat m = mask ? wrap(mask, w, hh, 1) : cv::Mat();
    return ((line2Dup::Detector *)h)->addTemplate(wrap(img, w, hh, c), class_id, m, -1.f, -1.f, 0, "none", num_features);
    L2D_CATCH(-2)
}
int l2d_add_template_rotate(void *h, const char *class_id, int zero_id, float theta, float cx, float cy)
{
    L2D_TRY
    return ((line2Dup::Detector *)h)->addTemplate_rotate(class_id, zero_id, theta, cv::Point2f(cx, cy));
    L2D_CATCH(-2)
}
int l2d_num_templates(void *h, const char *class_id) { return ((line2Dup::Detector *)h)->numTemplates(class_id); }

int l2d_class_sizes(void *h, const char *class_id, int *n, int *levels, int *total_feats)
{
    L2D_TRY
    auto *d = (line2Dup::Detector *)h;
    *n = d->numTemplates(class_id);
    *levels = d->pyramidLevels();
    int tot = 0;
    for (int t = 0; t < *n; ++t)
        for (const line2Dup::Template &tl : d->getTemplates(class_id, t))
            tot += (int)tl.features.size();
    *total_feats = tot;
    return 0;
    L2D_CATCH(-1)
}
int l2d_export_class(void *h, const char *class_id, int *tdesc, int *feats, float *thetas)
{
    L2D_TRY
    auto *d = (line2Dup::Detector *)h;
    size_t fo = 0, k = 0;
    const int n = d->numTemplates(class_id);
    for (int t = 0; t < n; ++t)
        for (const line2Dup::Template &tl : d->getTemplates(class_id, t))
        {
            int *o = tdesc + (k++) * 5;
            o[0] = tl.width, o[1] = tl.height, o[2] = tl.tl_x, o[3] = tl.tl_y, o[4] = (int)tl.features.size();
            for (const line2Dup::Feature &f : tl.features)
            {
                feats[fo * 3] = f.x, feats[fo * 3 + 1] = f.y, feats[fo * 3 + 2] = f.label;
                if (thetas)
                    thetas[fo] = f.theta;
                ++fo;
            }
        }
    return 0;
    L2D_CATCH(-1)
}
int l2d_read_class(void *h, const char *class_id, const char *fmt)
{
    L2D_TRY
    ((line2Dup::Detector *)h)->readClasses({class_id}, fmt);
    return ((line2Dup::Detector *)h)->numTemplates(class_id);
    L2D_CATCH(-1)
}
int l2d_write_classes(void *h, const char *fmt)
{
    L2D_TRY
    ((line2Dup::Detector *)h)->writeClasses(fmt);
    return 0;
    L2D_CATCH(-1)
}

// Detector::match -> sbm_match_rec-shaped records; class_idx indexes Detector::classIds() (std::map order)
int l2d_match(void *h, const uint8_t *img, int w, int hh, int c, const uint8_t *mask, float thr, const char *class_csv, sbm_match_rec *out,
              int cap, int *n_out)
{
    L2D_TRY
    auto *d = (line2Dup::Detector *)h;
    const std::vector<std::string> ids = d->classIds();
    std::vector<line2Dup::Match> m = d->match(wrap(img, w, hh, c), thr, splitIds(class_csv), mask ? wrap(mask, w, hh, 1) : cv::Mat());
    *n_out = (int)m.size();
    for (int i = 0; i < (int)m.size() && i < cap; ++i)
    {
        int k = 0;
        while (k < (int)ids.size() && ids[k] != m[i].class_id)
            ++k;
        out[i] = sbm_match_rec{m[i].x, m[i].y, m[i].similarity, k, m[i].template_id};
    }
    return 0;
    L2D_CATCH(-1)
}

int l2d_group_unique_id(void *id128)
{
    L2D_TRY
    const std::vector<unsigned char> id = line2Dup::Detector::groupUniqueId();
    memcpy(id128, id.data(), id.size());
    return 0;
    L2D_CATCH(-1)
}
int l2d_set_group(void *h, int n_ranks, int rank, const void *id128, int mode)
{
    L2D_TRY
    const unsigned char *p = (const unsigned char *)id128;
    ((line2Dup::Detector *)h)->setGroup(n_ranks, rank, std::vector<unsigned char>(p, p + SBM_GROUP_ID_BYTES), (line2Dup::Detector::GroupMode)mode);
    return 0;
    L2D_CATCH(-1)
}

// shape_based_matching::shapeInfo_producer (line2Dup.h:344-460 of the reference)
int l2d_shape_infos(const float *angle_range, int n_angle, const float *scale_range, int n_scale, float angle_step, float scale_step, float *out,
                    int cap, int *n_out)
{
    L2D_TRY
    shape_based_matching::shapeInfo_producer sp(cv::Mat(8, 8, CV_8UC1));
    sp.angle_range.assign(angle_range, angle_range + n_angle);
    sp.scale_range.assign(scale_range, scale_range + n_scale);
    sp.angle_step = angle_step;
    sp.scale_step = scale_step;
    sp.produce_infos();
    *n_out = (int)sp.infos.size();
    for (int i = 0; i < (int)sp.infos.size() && i < cap; ++i)
        out[2 * i] = sp.infos[i].angle, out[2 * i + 1] = sp.infos[i].scale;
    return 0;
    L2D_CATCH(-1)
}
int l2d_shape_transform(const uint8_t *img, int w, int hh, int c, float angle, float scale, int as_mask, uint8_t *out, int *ow, int *oh)
{
    L2D_TRY
    shape_based_matching::shapeInfo_producer sp(wrap(img, w, hh, c), as_mask ? wrap(img, w, hh, c) : cv::Mat());
    shape_based_matching::shapeInfo_producer::Info info(angle, scale);
    cv::Mat d = as_mask ? sp.mask_of(info) : sp.src_of(info);
    *ow = d.cols, *oh = d.rows;
    if (out)
        memcpy(out, d.data, (size_t)d.cols * d.rows * d.channels());
    return 0;
    L2D_CATCH(-1)
}
}
