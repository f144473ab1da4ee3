// ORACLE / TEST INFRASTRUCTURE -- never linked into, imported by or called from the product path.
//
// Builds the reference's OWN implementation into oracle/_ref/libsbm_ref.so: this translation unit
// textually includes /root/reference/line2Dup.cpp (found through -I/root/reference at build time,
// never copied into the repo) so that its file-static functions (spread, computeResponseMaps,
// linearize, similarity*, ...) and Detector's protected members are reachable, and exposes them
// through a small C interface for ctypes (tests/, bench.py --impl reference / cpu_baseline).
//
// OpenCV comes from oracle/shim (restated imgproc, pinned against cv2 4.13.0) + include/sbm/cv_compat.hpp.
// Build: see oracle/Makefile (-O3 -mavx2 -mfma -fopenmp; MIPP 256-bit; NOT -march=native, whose
// AVX-512 back end does not compile: /root/reference/MIPP/mipp_impl_AVX512.hxx:336).
#include "line2Dup.cpp"

#include <omp.h>

namespace
{
using namespace line2Dup;

struct RefMatch
{
    int x, y;
    float similarity;
    int class_idx;
    int template_id;
};

struct Probe : public Detector
{
    Probe(int nf, std::vector<int> T, float weak, float strong) : Detector(nf, T, weak, strong) {}

    typedef std::vector<cv::Mat> LM;
    struct Front
    {
        std::vector<cv::Mat> quantized, spreaded;
        std::vector<std::vector<cv::Mat>> response;
        LinearMemoryPyramid lm;
        std::vector<cv::Size> sizes;
    };

    // Same sequence of reference calls as Detector::match's first half (line2Dup.cpp:1084-1120).
    void front(const cv::Mat &src, const cv::Mat &mask, Front &f, bool keep) const
    {
        cv::Ptr<ColorGradientPyramid> q = modality->process(src, mask);
        f.lm = LinearMemoryPyramid(pyramid_levels, std::vector<LinearMemories>(1, LinearMemories(8)));
        f.sizes.clear();
        f.quantized.clear();
        f.spreaded.clear();
        f.response.clear();
        for (int l = 0; l < pyramid_levels; ++l)
        {
            if (l > 0)
                q->pyrDown();
            cv::Mat quant, spr;
            std::vector<cv::Mat> resp;
            q->quantize(quant);
            spread(quant, spr, T_at_level[l]);
            computeResponseMaps(spr, resp);
            for (int j = 0; j < 8; ++j)
                linearize(resp[j], f.lm[l][0][j], T_at_level[l]);
            f.sizes.push_back(quant.size());
            if (keep)
            {
                f.quantized.push_back(quant);
                f.spreaded.push_back(spr);
                f.response.push_back(resp);
            }
        }
    }

    std::vector<std::string> resolve(const std::vector<std::string> &ids) const
    {
        std::vector<std::string> out;
        if (ids.empty())
            for (auto &kv : class_templates)
                out.push_back(kv.first);
        else
            for (auto &id : ids)
                if (class_templates.count(id))
                    out.push_back(id);
        return out;
    }

    // Second half of Detector::match up to (not including) sort/unique (line2Dup.cpp:1124-1140).
    void candidates(const Front &f, float thr, const std::vector<std::string> &ids, std::vector<Match> &out) const
    {
        for (auto &id : resolve(ids))
            matchClass(f.lm, f.sizes, thr, out, id, class_templates.at(id));
    }

    TemplatesMap &templates() { return class_templates; }
    const std::vector<int> &Ts() const { return T_at_level; }
};

thread_local std::string g_err;

std::vector<std::string> splitIds(const char *csv)
{
    std::vector<std::string> ids;
    if (!csv || !*csv)
        return ids;
    std::string s(csv), tok;
    std::istringstream is(s);
    while (std::getline(is, tok, ','))
        ids.push_back(tok);
    return ids;
}

cv::Mat wrap(const uint8_t *img, int w, int h, int c)
{
    cv::Mat m(h, w, CV_MAKETYPE(CV_8U, c));
    memcpy(m.data, img, (size_t)w * h * c);
    return m;
}

int classIndex(Probe *p, const std::string &id)
{
    int i = 0;
    for (auto &kv : p->templates())
    {
        if (kv.first == id)
            return i;
        ++i;
    }
    return -1;
}

void exportMatches(Probe *p, const std::vector<Match> &m, RefMatch *out, int cap, int *n_out)
{
    *n_out = (int)m.size();
    for (int i = 0; i < (int)m.size() && i < cap; ++i)
        out[i] = RefMatch{m[i].x, m[i].y, m[i].similarity, classIndex(p, m[i].class_id), m[i].template_id};
}
} // namespace

#define REF_TRY try {
#define REF_CATCH                    \
    }                                \
    catch (const std::exception &e)  \
    {                                \
        g_err = e.what();            \
        return -1;                   \
    }

extern "C"
{
const char *ref_last_error() { return g_err.c_str(); }
int ref_simd_width() { return mipp::N<uint8_t>(); }

void *ref_create(int num_features, const int *T, int nT, float weak, float strong)
{
    return new Probe(num_features, std::vector<int>(T, T + nT), weak, strong);
}
void ref_destroy(void *h) { delete (Probe *)h; }

// Detector::readClasses (line2Dup.cpp:1577-1587) on one class id; fmt like ".../%s_templ.yaml".
int ref_read_class(void *h, const char *class_id, const char *fmt)
{
    REF_TRY
    ((Probe *)h)->readClasses({class_id}, fmt);
    return ((Probe *)h)->numTemplates(class_id);
    REF_CATCH
}
int ref_write_classes(void *h, const char *fmt)
{
    REF_TRY
    ((Probe *)h)->writeClasses(fmt);
    return 0;
    REF_CATCH
}

// Flat interchange: tdesc[n*levels][5] = {width,height,tl_x,tl_y,nf}; feats[sum nf][3] = {x,y,label}.
int ref_add_class(void *h, const char *class_id, int n, int levels, const int *tdesc, const int *feats)
{
    REF_TRY
    Probe *p = (Probe *)h;
    std::vector<std::vector<Template>> tps(n, std::vector<Template>(levels));
    size_t fo = 0;
    for (int t = 0; t < n; ++t)
        for (int l = 0; l < levels; ++l)
        {
            const int *d = tdesc + ((size_t)t * levels + l) * 5;
            Template &tp = tps[t][l];
            tp.width = d[0], tp.height = d[1], tp.tl_x = d[2], tp.tl_y = d[3], tp.pyramid_level = l;
            tp.sscale = 0, tp.orientation = 0, tp.tagFieldID = 0;
            tp.features.resize(d[4]);
            for (int i = 0; i < d[4]; ++i, ++fo)
            {
                tp.features[i] = Feature(feats[fo * 3], feats[fo * 3 + 1], feats[fo * 3 + 2]);
                tp.features[i].theta = 0;
            }
        }
    p->templates()[class_id] = tps;
    return n;
    REF_CATCH
}
int ref_class_sizes(void *h, const char *class_id, int *n, int *levels, int *total_feats)
{
    Probe *p = (Probe *)h;
    auto it = p->templates().find(class_id);
    if (it == p->templates().end())
        return -1;
    *n = (int)it->second.size();
    *levels = p->pyramidLevels();
    int tot = 0;
    for (auto &tp : it->second)
        for (auto &t : tp)
            tot += (int)t.features.size();
    *total_feats = tot;
    return 0;
}
int ref_export_class(void *h, const char *class_id, int *tdesc, int *feats, float *thetas)
{
    Probe *p = (Probe *)h;
    auto it = p->templates().find(class_id);
    if (it == p->templates().end())
        return -1;
    size_t fo = 0, k = 0;
    for (auto &tp : it->second)
        for (auto &t : tp)
        {
            int *d = tdesc + (k++) * 5;
            d[0] = t.width, d[1] = t.height, d[2] = t.tl_x, d[3] = t.tl_y, d[4] = (int)t.features.size();
            for (auto &f : t.features)
            {
                feats[fo * 3] = f.x, feats[fo * 3 + 1] = f.y, feats[fo * 3 + 2] = f.label;
                if (thetas)
                    thetas[fo] = f.theta;
                ++fo;
            }
        }
    return 0;
}

// Detector::addTemplate / addTemplate_rotate (line2Dup.cpp:1299-1353, :1409-1451)
int ref_add_template(void *h, const uint8_t *img, int w, int hh, int c, const uint8_t *mask, const char *class_id, int num_features)
{
    REF_TRY
    cv::Mat m = mask ? wrap(mask, w, hh, 1) : cv::Mat();
    return ((Probe *)h)->addTemplate(wrap(img, w, hh, c), class_id, m, -1.f, -1.f, 0, "none", num_features);
    REF_CATCH
}
int ref_add_template_rotate(void *h, const char *class_id, int zero_id, float theta, float cx, float cy)
{
    REF_TRY
    return ((Probe *)h)->addTemplate_rotate(class_id, zero_id, theta, cv::Point2f(cx, cy));
    REF_CATCH
}

// shape_based_matching::shapeInfo_producer (line2Dup.h:344-460): the `infos` sequence of produce_infos (float
// accumulation of `angle += angle_step` included) and src_of / mask_of of one Info.
int ref_shape_infos(const float *angle_range, int n_angle, const float *scale_range, int n_scale, float angle_step, float scale_step,
                    float *out, int cap, int *n_out)
{
    REF_TRY
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
    REF_CATCH
}
int ref_shape_transform(const uint8_t *img, int w, int hh, int c, float angle, float scale, int as_mask, uint8_t *out, int *ow, int *oh)
{
    REF_TRY
    shape_based_matching::shapeInfo_producer sp(wrap(img, w, hh, c), as_mask ? wrap(img, w, hh, c) : cv::Mat());
    shape_based_matching::shapeInfo_producer::Info info(angle, scale);
    cv::Mat d = as_mask ? sp.mask_of(info) : sp.src_of(info);
    *ow = d.cols, *oh = d.rows;
    if (out)
        memcpy(out, d.data, (size_t)d.cols * d.rows * d.channels());
    return 0;
    REF_CATCH
}

// Detector::match verbatim (line2Dup.cpp:1078-1150).  omp_threads<=0 keeps the OpenMP default.
int ref_match(void *h, const uint8_t *img, int w, int hh, int c, const uint8_t *mask, float thr, const char *class_csv, int omp_threads,
              RefMatch *out, int cap, int *n_out)
{
    REF_TRY
    if (omp_threads > 0)
        omp_set_num_threads(omp_threads);
    Probe *p = (Probe *)h;
    // mask: NULL or u8 [hh][w], handed to Detector::match as its `masks` argument (line2Dup.cpp:1078-1086)
    std::vector<Match> m = p->match(wrap(img, w, hh, c), thr, splitIds(class_csv), mask ? wrap(mask, w, hh, 1) : cv::Mat());
    exportMatches(p, m, out, cap, n_out);
    return 0;
    REF_CATCH
}

// Candidate list BEFORE sort/unique, in the reference's serial append order (1 OpenMP thread).
int ref_candidates(void *h, const uint8_t *img, int w, int hh, int c, const uint8_t *mask, float thr, const char *class_csv, RefMatch *out,
                   int cap, int *n_out)
{
    REF_TRY
    omp_set_num_threads(1);
    Probe *p = (Probe *)h;
    Probe::Front f;
    p->front(wrap(img, w, hh, c), mask ? wrap(mask, w, hh, 1) : cv::Mat(), f, false);
    std::vector<Match> m;
    p->candidates(f, thr, splitIds(class_csv), m);
    exportMatches(p, m, out, cap, n_out);
    return 0;
    REF_CATCH
}

// Intermediates of one pyramid level.  Any output pointer may be NULL.
//   quantized/spread: h_l*w_l bytes; response: 8*h_l*w_l; linmem: 8 * T*T * (w_l/T*h_l/T) bytes
//   ordered [label][grid row][cell] exactly as the reference's 8 Mats laid end to end.
int ref_front(void *h, const uint8_t *img, int w, int hh, int c, const uint8_t *mask, int level, uint8_t *quantized, uint8_t *spread_out,
              uint8_t *response, uint8_t *linmem)
{
    REF_TRY
    Probe *p = (Probe *)h;
    Probe::Front f;
    p->front(wrap(img, w, hh, c), mask ? wrap(mask, w, hh, 1) : cv::Mat(), f, true);
    const cv::Mat &q = f.quantized[level];
    size_t n = (size_t)q.rows * q.cols;
    if (quantized)
        memcpy(quantized, q.data, n);
    if (spread_out)
        memcpy(spread_out, f.spreaded[level].data, n);
    for (int j = 0; j < 8; ++j)
    {
        if (response)
            memcpy(response + j * n, f.response[level][j].data, n);
        if (linmem)
            memcpy(linmem + j * n, f.lm[level][0][j].data, n);
    }
    return 0;
    REF_CATCH
}

// Raw coarse similarity map (u16, H*W at the last level) of one template, through the same
// similarity_64 / similarity dispatch as matchClass (line2Dup.cpp:1189-1196).
int ref_similarity(void *h, const uint8_t *img, int w, int hh, int c, const uint8_t *mask, const char *class_id, int template_id, uint16_t *out)
{
    REF_TRY
    Probe *p = (Probe *)h;
    Probe::Front f;
    p->front(wrap(img, w, hh, c), mask ? wrap(mask, w, hh, 1) : cv::Mat(), f, false);
    const Template &t = p->getTemplates(class_id, template_id).back();
    int T = p->Ts().back();
    cv::Mat sim;
    if (t.features.size() < 64)
    {
        similarity_64(f.lm.back()[0], t, sim, f.sizes.back(), T);
        sim.convertTo(sim, CV_16U);
    }
    else
        similarity(f.lm.back()[0], t, sim, f.sizes.back(), T);
    memcpy(out, sim.data, (size_t)sim.rows * sim.cols * 2);
    return 0;
    REF_CATCH
}

// 16x16 local similarity patch at level `level` around centre (x,y) (line2Dup.cpp:1254-1261).
int ref_similarity_local(void *h, const uint8_t *img, int w, int hh, int c, const uint8_t *mask, const char *class_id, int template_id,
                         int level, int x, int y, uint16_t *out)
{
    REF_TRY
    Probe *p = (Probe *)h;
    Probe::Front f;
    p->front(wrap(img, w, hh, c), mask ? wrap(mask, w, hh, 1) : cv::Mat(), f, false);
    const Template &t = p->getTemplates(class_id, template_id)[level];
    int T = p->Ts()[level];
    cv::Mat sim;
    if (t.features.size() < 64)
    {
        similarityLocal_64(f.lm[level][0], t, sim, f.sizes[level], T, cv::Point(x, y));
        sim.convertTo(sim, CV_16U);
    }
    else
        similarityLocal(f.lm[level][0], t, sim, f.sizes[level], T, cv::Point(x, y));
    memcpy(out, sim.data, 512);
    return 0;
    REF_CATCH
}

// CPU baseline timing.  Mirrors the reference's two commented timer sites (line2Dup.cpp:1122, :1147):
// ms_front = linear-memory pyramid construction, ms_match = all matchClass calls + sort/unique,
// ms_total = Detector::match() as a whole, verbatim.  Each is the mean over `iters` runs.
int ref_time(void *h, const uint8_t *img, int w, int hh, int c, float thr, const char *class_csv, int omp_threads, int iters,
             double *ms_front, double *ms_match, double *ms_total, int *n_matches)
{
    REF_TRY
    if (omp_threads > 0)
        omp_set_num_threads(omp_threads);
    Probe *p = (Probe *)h;
    cv::Mat src = wrap(img, w, hh, c);
    std::vector<std::string> ids = splitIds(class_csv);
    typedef std::chrono::steady_clock clk;
    double tf = 0, tm = 0, tt = 0;
    size_t nm = 0;
    for (int it = 0; it < iters; ++it)
    {
        auto t0 = clk::now();
        Probe::Front f;
        p->front(src, cv::Mat(), f, false);
        auto t1 = clk::now();
        std::vector<Match> m;
        p->candidates(f, thr, ids, m);
        std::sort(m.begin(), m.end());
        m.erase(std::unique(m.begin(), m.end()), m.end());
        auto t2 = clk::now();
        std::vector<Match> m2 = p->match(src, thr, ids);
        auto t3 = clk::now();
        tf += std::chrono::duration<double, std::milli>(t1 - t0).count();
        tm += std::chrono::duration<double, std::milli>(t2 - t1).count();
        tt += std::chrono::duration<double, std::milli>(t3 - t2).count();
        nm = m2.size();
    }
    *ms_front = tf / iters, *ms_match = tm / iters, *ms_total = tt / iters;
    *n_matches = (int)nm;
    return 0;
    REF_CATCH
}

// ---- single OpenCV-restatement entry points, for pinning oracle/shim against cv2 (tests/test_oracle_pins.py) ----
int ref_cv_gaussian7(const uint8_t *img, int w, int hh, int c, uint8_t *out)
{
    REF_TRY
    cv::Mat d;
    cv::GaussianBlur(wrap(img, w, hh, c), d, cv::Size(7, 7), 0, 0, cv::BORDER_REPLICATE);
    memcpy(out, d.data, (size_t)w * hh * c);
    return 0;
    REF_CATCH
}
int ref_cv_sobel(const uint8_t *img, int w, int hh, int c, int dx, int dy, int16_t *out)
{
    REF_TRY
    cv::Mat d;
    cv::Sobel(wrap(img, w, hh, c), d, CV_16S, dx, dy, 3, 1.0, 0.0, cv::BORDER_REPLICATE);
    memcpy(out, d.data, (size_t)w * hh * c * 2);
    return 0;
    REF_CATCH
}
int ref_cv_phase(const float *x, const float *y, int n, float *out)
{
    REF_TRY
    cv::Mat mx(1, n, CV_32FC1), my(1, n, CV_32FC1), a;
    memcpy(mx.data, x, (size_t)n * 4);
    memcpy(my.data, y, (size_t)n * 4);
    cv::phase(mx, my, a, true);
    memcpy(out, a.data, (size_t)n * 4);
    return 0;
    REF_CATCH
}
int ref_cv_convert_u8(const float *x, int n, double alpha, uint8_t *out)
{
    REF_TRY
    cv::Mat m(1, n, CV_32FC1), d;
    memcpy(m.data, x, (size_t)n * 4);
    m.convertTo(d, CV_8U, alpha);
    memcpy(out, d.data, n);
    return 0;
    REF_CATCH
}
int ref_cv_pyrdown(const uint8_t *img, int w, int hh, int c, uint8_t *out)
{
    REF_TRY
    cv::Mat d;
    cv::pyrDown(wrap(img, w, hh, c), d, cv::Size(w / 2, hh / 2));
    memcpy(out, d.data, (size_t)(w / 2) * (hh / 2) * c);
    return 0;
    REF_CATCH
}
int ref_cv_resize_linear(const uint8_t *img, int w, int hh, int c, double fx, double fy, uint8_t *out, int *ow, int *oh)
{
    REF_TRY
    cv::Mat d;
    cv::resize(wrap(img, w, hh, c), d, cv::Size(), fx, fy);
    *ow = d.cols, *oh = d.rows;
    if (out)
        memcpy(out, d.data, (size_t)d.cols * d.rows * c);
    return 0;
    REF_CATCH
}
int ref_cv_erode3(const uint8_t *img, int w, int hh, uint8_t *out)
{
    REF_TRY
    cv::Mat d;
    cv::erode(wrap(img, w, hh, 1), d, cv::Mat(), cv::Point(-1, -1), 1, cv::BORDER_REPLICATE);
    memcpy(out, d.data, (size_t)w * hh);
    return 0;
    REF_CATCH
}

// Workload counters of one frame, derived from the reference's own functions:
//   coarse = sum_t (features inside the coarse image) * template_positions      (line2Dup.cpp:813-857)
//   coarse_cands = positions with score > threshold                              (line2Dup.cpp:1200-1216)
//   local  = 256 * sum over coarse candidates of features at level L-2           (line2Dup.cpp:1254-1261)
// (deeper pyramids: only the first refinement level is counted).
int ref_eval_counts(void *h, const uint8_t *img, int w, int hh, int c, float thr, const char *class_csv,
                    unsigned long long *coarse, unsigned long long *coarse_cands, unsigned long long *local)
{
    REF_TRY
    Probe *p = (Probe *)h;
    Probe::Front f;
    p->front(wrap(img, w, hh, c), cv::Mat(), f, false);
    const int L = p->pyramidLevels();
    const int T = p->Ts().back();
    const cv::Size sz = f.sizes.back();
    const int W = sz.width / T, H = sz.height / T;
    unsigned long long ce = 0, cc = 0, le = 0;
    for (auto &id : p->resolve(splitIds(class_csv)))
    {
        const auto &tps = p->templates().at(id);
        for (size_t t = 0; t < tps.size(); ++t)
        {
            const Template &tc = tps[t].back();
            int wf = (tc.width - 1) / T + 1, hf = (tc.height - 1) / T + 1;
            int tp = (H - hf) * W + (W - wf) + 1;
            int nv = 0;
            for (auto &ft : tc.features)
                if (!(ft.x < 0 || ft.x >= sz.width || ft.y < 0 || ft.y >= sz.height))
                    ++nv;
            ce += (unsigned long long)nv * (unsigned long long)std::max(tp, 0);
            cv::Mat sim;
            if (tc.features.size() < 64)
            {
                similarity_64(f.lm.back()[0], tc, sim, sz, T);
                sim.convertTo(sim, CV_16U);
            }
            else
                similarity(f.lm.back()[0], tc, sim, sz, T);
            int nf = (int)tc.features.size();
            unsigned long long n = 0;
            for (int r = 0; r < sim.rows; ++r)
                for (int q = 0; q < sim.cols; ++q)
                    if ((sim.at<ushort>(r, q) * 100.f) / (4 * nf) > thr)
                        ++n;
            cc += n;
            if (L >= 2)
                le += n * 256ull * (unsigned long long)tps[t][L - 2].features.size();
        }
    }
    *coarse = ce, *coarse_cands = cc, *local = le;
    return 0;
    REF_CATCH
}
}
