// line2Dup.cpp -- host side of the drop-in C++ API (include/line2Dup.h) over the C ABI (include/sbm_capi.h).
//
// What runs where:
//   Detector::match / matchBatch        -> sbm_match / sbm_match_batch: every stage on the GPU (csrc/sbm_kernels.cuh)
//   ColorGradientPyramid::update        -> sbm_quantized_orientations: blur, Sobel, phase, hysteresis, pyrDown on the GPU
//   extractTemplate / selectScatteredFeatures / cropTemplates / addTemplate_rotate, read*/write*
//                                       -> host code below: sequential selection over a few thousand candidates and
//                                          YAML persistence; each function cites the reference lines it follows.
// Built with the system g++ (no nvcc needed): shape_based_matching_b200/libline2Dup_b200.so.
#include "line2Dup.h"
#include "sbm_capi.h"

#include <algorithm>
#include <limits>
#include <mutex>

using namespace cv;

namespace line2Dup
{
namespace
{
[[noreturn]] void throwSbm(int rc, const char *what)
{
    // the reference signals bad input with CV_Assert / CV_Error -> cv::Exception; so do we
    const int code = rc == SBM_ERR_BAD_ARG ? Error::StsBadArg : Error::StsError;
    throw cv::Exception(code, std::string(what) + ": " + sbm_last_error(), __func__, __FILE__, __LINE__);
}

// one-hot quantised orientation -> label 0..7 (line2Dup.cpp:16-40)
int labelOf(int quantized)
{
    for (int b = 0; b < 8; ++b)
        if (quantized == (1 << b))
            return b;
    CV_Error(Error::StsBadArg, "Invalid value of quantized parameter");
}

// process-wide context used by the training front end (ColorGradientPyramid carries no Detector pointer in the
// reference's interface).  weak_threshold only gates the vote, it is passed per call through a context per value.
// Guarded by g_train_mu (Detector::setDevice may run on another thread); the cached contexts are destroyed at exit.
std::mutex g_train_mu;
int g_train_device = 0;
struct TrainContexts
{
    std::map<std::pair<int, float>, sbm_ctx *> ctx;
    ~TrainContexts()
    {
        for (auto &kv : ctx)
            sbm_destroy(kv.second);
    }
} g_train;

sbm_ctx *trainingContext(float weak)
{
    std::lock_guard<std::mutex> lock(g_train_mu);
    auto key = std::make_pair(g_train_device, weak);
    auto it = g_train.ctx.find(key);
    if (it != g_train.ctx.end())
        return it->second;
    const int T[2] = {4, 8};
    sbm_ctx *c = nullptr;
    int rc = sbm_create(T, 2, weak, 60.f, 63, g_train_device, &c);
    if (rc)
        throwSbm(rc, "sbm_create");
    g_train.ctx[key] = c;
    return c;
}

void setTrainingDevice(int device)
{
    std::lock_guard<std::mutex> lock(g_train_mu);
    g_train_device = device;
}

Mat continuous8u(const Mat &m)
{
    CV_Assert(m.depth() == CV_8U);
    return m.isContinuous() ? m : m.clone();
}
} // namespace

// ---------------------------------------------------------------- persistence (line2Dup.cpp:42-113)

void Feature::read(const FileNode &fn)
{
    FileNodeIterator fni = fn.begin();
    fni >> x >> y >> label;
}

void Feature::write(FileStorage &fs) const { fs << "[:" << x << y << label << "]"; }

void Template::read(const FileNode &fn)
{
    width = fn["width"];
    height = fn["height"];
    tl_x = fn["tl_x"];
    tl_y = fn["tl_y"];
    // written by this fork only; absent keys read as 0 / "" like an empty cv::FileNode (test/case*/ *_templ.yaml)
    sscale = fn["scale"];
    orientation = fn["orientation"];
    tagFieldID = fn["tagFieldID"];
    fn["fiducial_src"] >> fiducial_src;
    pyramid_level = fn["pyramid_level"];

    FileNode features_fn = fn["features"];
    features.resize(features_fn.size());
    FileNodeIterator it = features_fn.begin(), it_end = features_fn.end();
    for (int i = 0; it != it_end; ++it, ++i)
        features[i].read(*it);
}

void Template::write(FileStorage &fs) const
{
    fs << "width" << width;
    fs << "height" << height;
    fs << "tl_x" << tl_x;
    fs << "tl_y" << tl_y;
    fs << "scale" << sscale;
    fs << "orientation" << orientation;
    fs << "tagFieldID" << tagFieldID;
    fs << "fiducial_src" << fiducial_src;
    fs << "pyramid_level" << pyramid_level;
    fs << "features"
       << "[";
    for (int i = 0; i < (int)features.size(); ++i)
        features[i].write(fs);
    fs << "]";
}

// ---------------------------------------------------------------- ColorGradient (line2Dup.cpp:541-578)

ColorGradient::ColorGradient() : weak_threshold(30.0f), num_features(63), strong_threshold(60.0f) {}

ColorGradient::ColorGradient(float _weak_threshold, size_t _num_features, float _strong_threshold)
    : weak_threshold(_weak_threshold), num_features(_num_features), strong_threshold(_strong_threshold)
{
}

static const char CG_NAME[] = "ColorGradient";

std::string ColorGradient::name() const { return CG_NAME; }

void ColorGradient::read(const FileNode &fn)
{
    std::string type = fn["type"];
    CV_Assert(type == CG_NAME);
    weak_threshold = fn["weak_threshold"];
    num_features = int(fn["num_features"]);
    strong_threshold = fn["strong_threshold"];
}

void ColorGradient::write(FileStorage &fs) const
{
    fs << "type" << CG_NAME;
    fs << "weak_threshold" << weak_threshold;
    fs << "num_features" << int(num_features);
    fs << "strong_threshold" << strong_threshold;
}

// ---------------------------------------------------------------- ColorGradientPyramid (line2Dup.cpp:406-539)

ColorGradientPyramid::ColorGradientPyramid(const Mat &_src, const Mat &_mask, float _weak_threshold, size_t _num_features,
                                           float _strong_threshold)
    : src(_src), mask(_mask), pyramid_level(0), weak_threshold(_weak_threshold), num_features(_num_features),
      strong_threshold(_strong_threshold)
{
    CV_Assert(src.channels() == 1 || src.channels() == 3);
    update();
}

// quantizedOrientations + hysteresisGradient (:218-404) on the GPU.  steps = 0: of `src` as it is (constructor);
// steps = 1: of cv::pyrDown(src) (:431-433), which also becomes the new `src` -- one level per call, the chain is
// never recomputed from level 0.
void ColorGradientPyramid::update() { updateFrom(0); }

void ColorGradientPyramid::updateFrom(int steps)
{
    const Mat cur = continuous8u(src);
    int w = cur.cols, h = cur.rows;
    for (int l = 0; l < steps; ++l)
        w /= 2, h /= 2;
    magnitude.create(h, w, CV_32FC1);
    angle_ori.create(h, w, CV_32FC1);
    angle.create(h, w, CV_8UC1);
    Mat level_src(h, w, cur.type());
    int wo = 0, ho = 0;
    int rc = sbm_quantized_orientations(trainingContext(weak_threshold), cur.data, cur.cols, cur.rows, cur.channels(), cur.step, steps,
                                        magnitude.ptr<float>(), angle_ori.ptr<float>(), angle.data, steps ? level_src.data : nullptr, &wo,
                                        &ho);
    if (rc)
        throwSbm(rc, "sbm_quantized_orientations");
    CV_Assert(wo == w && ho == h);
    if (steps > 0)
        src = level_src;
}

// :424-444
void ColorGradientPyramid::pyrDown()
{
    num_features /= 2;
    ++pyramid_level;
    Size size(src.cols / 2, src.rows / 2);
    if (!mask.empty())
    {
        Mat next_mask;
        resize(mask, next_mask, size, 0.0, 0.0, INTER_NEAREST);
        mask = next_mask;
    }
    updateFrom(1); // cv::pyrDown of the image happens on the device
    CV_Assert(src.cols == size.width && src.rows == size.height);
}

// :446-450
void ColorGradientPyramid::quantize(Mat &dst) const
{
    dst = Mat::zeros(angle.size(), CV_8U);
    angle.copyTo(dst, mask);
}

// :452-539
bool ColorGradientPyramid::extractTemplate(Template &templ) const
{
    // features on the border distinguish the object from the background: shrink the mask by one pixel
    Mat local_mask;
    if (!mask.empty())
        erode(mask, local_mask, Mat(), Point(-1, -1), 1, BORDER_REPLICATE);

    std::vector<Candidate> candidates;
    const bool no_mask = local_mask.empty();
    const float threshold_sq = strong_threshold * strong_threshold;
    const int R = 5 / 2; // 5x5 non-maximum suppression
    const int rows = magnitude.rows, cols = magnitude.cols;
    Mat valid(rows, cols, CV_8UC1, Scalar(255));

    for (int r = R; r < rows - R; ++r)
    {
        const uchar *mask_r = no_mask ? nullptr : local_mask.ptr<uchar>(r);
        for (int c = R; c < cols - R; ++c)
        {
            if (!no_mask && !mask_r[c])
                continue;
            float score = 0;
            if (valid.at<uchar>(r, c) > 0)
            {
                score = magnitude.at<float>(r, c);
                bool is_max = true;
                for (int dr = -R; dr <= R && is_max; ++dr)
                    for (int dc = -R; dc <= R; ++dc)
                    {
                        if (dr == 0 && dc == 0)
                            continue;
                        if (score < magnitude.at<float>(r + dr, c + dc))
                        {
                            score = 0;
                            is_max = false;
                            break;
                        }
                    }
                if (is_max) // a local maximum invalidates its whole neighbourhood
                    for (int dr = -R; dr <= R; ++dr)
                        for (int dc = -R; dc <= R; ++dc)
                            if (dr != 0 || dc != 0)
                                valid.at<uchar>(r + dr, c + dc) = 0;
            }
            if (score > threshold_sq && angle.at<uchar>(r, c) > 0)
            {
                candidates.push_back(Candidate(c, r, labelOf(angle.at<uchar>(r, c)), score));
                candidates.back().f.theta = angle_ori.at<float>(r, c);
            }
        }
    }
    // a template needs a minimum of features (:514-520)
    if (candidates.size() < num_features)
    {
        if (candidates.size() <= 4)
            return false;
    }
    // stable: equal scores keep scan order
    std::stable_sort(candidates.begin(), candidates.end());
    // initial distance from the surplus of candidates
    float distance = static_cast<float>(candidates.size() / num_features + 1);
    if (!selectScatteredFeatures(candidates, templ.features, num_features, distance))
        return false;
    templ.width = -1; // set by cropTemplates
    templ.height = -1;
    templ.pyramid_level = pyramid_level;
    return true;
}

// :163-212 -- greedy selection with an adaptive minimum distance: grow the distance while a full pass still yields
// enough features, then relax it step by step until enough are found (or the distance drops under 3)
bool ColorGradientPyramid::selectScatteredFeatures(const std::vector<Candidate> &candidates, std::vector<Feature> &features,
                                                   size_t num_features, float distance)
{
    features.clear();
    float distance_sq = distance * distance;
    bool growing = true;
    const int n = (int)candidates.size();
    for (int i = 0;;)
    {
        const Feature &cf = candidates[i].f;
        bool keep = true;
        for (size_t j = 0; j < features.size() && keep; ++j)
        {
            const int dx = cf.x - features[j].x, dy = cf.y - features[j].y;
            keep = dx * dx + dy * dy >= distance_sq;
        }
        if (keep)
            features.push_back(cf);
        if (++i < n)
            continue;
        // one pass over the candidates is complete
        const bool num_ok = features.size() >= num_features;
        i = 0;
        if (growing)
        {
            if (num_ok)
            {
                features.clear(); // too many the first time: be pickier
                distance += 1.0f;
                distance_sq = distance * distance;
                continue;
            }
            growing = false;
        }
        distance -= 1.0f;
        distance_sq = distance * distance;
        if (num_ok || distance < 3)
            break;
    }
    return true;
}

// ---------------------------------------------------------------- cropTemplates (line2Dup.cpp:115-161)

static Rect cropTemplates(std::vector<Template> &templates)
{
    int min_x = std::numeric_limits<int>::max(), min_y = min_x;
    int max_x = std::numeric_limits<int>::min(), max_y = max_x;
    // bounding box of all features of all levels, in level-0 coordinates
    for (const Template &t : templates)
        for (const Feature &f : t.features)
        {
            const int x = f.x << t.pyramid_level, y = f.y << t.pyramid_level;
            min_x = std::min(min_x, x), max_x = std::max(max_x, x);
            min_y = std::min(min_y, y), max_y = std::max(max_y, y);
        }
    // even origin, so that it halves exactly at the next level
    if (min_x % 2 == 1)
        --min_x;
    if (min_y % 2 == 1)
        --min_y;
    for (Template &t : templates)
    {
        t.width = (max_x - min_x) >> t.pyramid_level;
        t.height = (max_y - min_y) >> t.pyramid_level;
        t.tl_x = min_x >> t.pyramid_level;
        t.tl_y = min_y >> t.pyramid_level;
        for (Feature &f : t.features)
            f.x -= t.tl_x, f.y -= t.tl_y;
    }
    return Rect(min_x, min_y, max_x - min_x, max_y - min_y);
}

// ---------------------------------------------------------------- Detector (line2Dup.cpp:1056-1599)

Detector::Detector() : ctx_(nullptr), templates_dirty_(true), device_(0)
{
    modality = makePtr<ColorGradient>();
    pyramid_levels = 2;
    T_at_level.push_back(4);
    T_at_level.push_back(8);
}

Detector::Detector(std::vector<int> T) : ctx_(nullptr), templates_dirty_(true), device_(0)
{
    modality = makePtr<ColorGradient>();
    pyramid_levels = (int)T.size();
    T_at_level = T;
}

Detector::Detector(int num_features, std::vector<int> T, float weak_thresh, float strong_thresh)
    : ctx_(nullptr), templates_dirty_(true), device_(0)
{
    modality = makePtr<ColorGradient>(weak_thresh, num_features, strong_thresh);
    pyramid_levels = (int)T.size();
    T_at_level = T;
}

Detector::~Detector()
{
    if (ctx_)
        sbm_destroy(ctx_);
}

void Detector::setDevice(int device)
{
    CV_Assert(!ctx_);
    device_ = device;
    setTrainingDevice(device);
}

std::vector<unsigned char> Detector::groupUniqueId()
{
    std::vector<unsigned char> id(SBM_GROUP_ID_BYTES);
    int rc = sbm_group_unique_id(id.data());
    if (rc)
        throwSbm(rc, "sbm_group_unique_id");
    return id;
}

void Detector::setGroup(int n_ranks, int rank, const std::vector<unsigned char> &unique_id, GroupMode mode)
{
    CV_Assert(unique_id.size() == SBM_GROUP_ID_BYTES && group_n_ == 0);
    sbm_ctx *c = context();
    int rc = sbm_clear_templates(c); // template sharding keeps a slice of every class: they are re-installed below
    if (!rc)
        rc = sbm_group_create(c, n_ranks, rank, unique_id.data(), (int)mode);
    if (rc)
        throwSbm(rc, "sbm_group_create");
    group_n_ = n_ranks, group_mode_ = (int)mode;
    templates_dirty_ = true;
}

sbm_ctx *Detector::context() const
{
    if (!ctx_)
    {
        int rc = sbm_create(T_at_level.data(), pyramid_levels, modality->weak_threshold, modality->strong_threshold,
                            (int)modality->num_features, device_, &ctx_);
        if (rc)
            throwSbm(rc, "sbm_create");
        templates_dirty_ = true;
    }
    return ctx_;
}

// class_templates -> device SoA, one class slot per class in std::map order (= the order match() visits them, :1124-1130)
void Detector::syncTemplates() const
{
    sbm_ctx *c = context();
    if (!templates_dirty_)
        return;
    int rc = sbm_clear_templates(c);
    if (rc)
        throwSbm(rc, "sbm_clear_templates");
    slots_.clear();
    int slot = 0;
    for (const auto &kv : class_templates)
    {
        std::vector<sbm_template_desc> desc;
        std::vector<sbm_feature> feats;
        for (const TemplatePyramid &tp : kv.second)
        {
            CV_Assert((int)tp.size() == pyramid_levels);
            for (const Template &t : tp)
            {
                CV_Assert(t.features.size() < 8192); // :811, :1195
                desc.push_back(sbm_template_desc{t.width, t.height, t.tl_x, t.tl_y, (int32_t)t.features.size()});
                for (const Feature &f : t.features)
                    feats.push_back(sbm_feature{f.x, f.y, f.label});
            }
        }
        rc = sbm_set_class_templates(c, slot, (int)kv.second.size(), desc.data(), feats.data());
        if (rc)
            throwSbm(rc, "sbm_set_class_templates");
        slots_.push_back(kv.first);
        ++slot;
    }
    templates_dirty_ = false;
}

namespace
{
// requested class ids -> class slots; unknown ids are skipped like in the reference (:1136-1138)
std::vector<int> slotsOf(const std::vector<std::string> &slots, const std::vector<std::string> &class_ids)
{
    std::vector<int> idx;
    for (const std::string &id : class_ids)
    {
        auto it = std::find(slots.begin(), slots.end(), id);
        idx.push_back(it == slots.end() ? -1 : (int)(it - slots.begin()));
    }
    return idx;
}
} // namespace

// :1078-1150
std::vector<Match> Detector::match(Mat source, float threshold, const std::vector<std::string> &class_ids, const Mat mask) const
{
    CV_Assert(mask.empty() || mask.size() == source.size()); // :1086
    CV_Assert(source.depth() == CV_8U && (source.channels() == 1 || source.channels() == 3));
    syncTemplates();
    const Mat img = continuous8u(source);
    Mat m8;
    if (!mask.empty())
        m8 = continuous8u(mask);
    const std::vector<int> idx = slotsOf(slots_, class_ids);
    if (!class_ids.empty() && std::all_of(idx.begin(), idx.end(), [](int v) { return v < 0; }))
        return std::vector<Match>(); // only unknown classes requested
    std::vector<sbm_match_rec> recs(1024);
    int n = 0;
    for (;;)
    {
        // in a template-sharding group the same call runs on every rank and returns the merged list (sbm_group_match)
        const bool sharded = group_n_ > 0 && group_mode_ != GROUP_FRAMES;
        int rc = (sharded ? sbm_group_match : sbm_match)(ctx_, img.data, img.cols, img.rows, img.channels(), img.step,
                                                         m8.empty() ? nullptr : m8.data, threshold, idx.empty() ? nullptr : idx.data(),
                                                         (int)idx.size(), recs.data(), (int)recs.size(), &n);
        if (rc == SBM_ERR_CAPACITY && n > (int)recs.size())
        {
            recs.resize(n);
            continue;
        }
        if (rc)
            throwSbm(rc, "sbm_match");
        break;
    }
    std::vector<Match> matches;
    matches.reserve(n);
    for (int i = 0; i < n; ++i)
        matches.emplace_back(recs[i].x, recs[i].y, recs[i].similarity, slots_[recs[i].class_idx], recs[i].template_id);
    return matches;
}

std::vector<std::vector<Match>> Detector::matchBatch(const std::vector<Mat> &sources, float threshold,
                                                     const std::vector<std::string> &class_ids) const
{
    const size_t lists = sources.size() * (group_n_ > 0 && group_mode_ == GROUP_FRAMES ? group_n_ : 1);
    std::vector<std::vector<Match>> out(lists);
    if (sources.empty())
        return out;
    syncTemplates();
    std::vector<Mat> imgs;
    std::vector<const uint8_t *> ptrs;
    for (const Mat &s : sources)
    {
        CV_Assert(s.depth() == CV_8U && s.size() == sources[0].size() && s.type() == sources[0].type());
        imgs.push_back(continuous8u(s));
        ptrs.push_back(imgs.back().data);
    }
    const std::vector<int> idx = slotsOf(slots_, class_ids);
    if (!class_ids.empty() && std::all_of(idx.begin(), idx.end(), [](int v) { return v < 0; }))
        return out;
    const int n = (int)sources.size(), nl = (int)lists;
    std::vector<sbm_match_rec> recs((size_t)256 * nl);
    std::vector<int> begin(nl + 1);
    for (;;)
    {
        int rc = (group_n_ > 0 ? sbm_group_match_batch : sbm_match_batch)(ctx_, ptrs.data(), nullptr, n, imgs[0].cols, imgs[0].rows,
                                                                          imgs[0].channels(), imgs[0].step, threshold,
                                                                          idx.empty() ? nullptr : idx.data(), (int)idx.size(), recs.data(),
                                                                          (int)recs.size(), begin.data());
        if (rc == SBM_ERR_CAPACITY && begin[nl] > (int)recs.size())
        {
            recs.resize(begin[nl]);
            continue;
        }
        if (rc)
            throwSbm(rc, "sbm_match_batch");
        break;
    }
    for (int f = 0; f < nl; ++f)
        for (int i = begin[f]; i < begin[f + 1]; ++i)
            out[f].emplace_back(recs[i].x, recs[i].y, recs[i].similarity, slots_[recs[i].class_idx], recs[i].template_id);
    return out;
}

// :1299-1353
int Detector::addTemplate(const Mat source, const std::string &class_id, const Mat &object_mask, float sscale, float orientation,
                          int tagFieldID, std::string fiducial_src, int num_features)
{
    std::vector<TemplatePyramid> &template_pyramids = class_templates[class_id];
    const int template_id = static_cast<int>(template_pyramids.size());
    TemplatePyramid tp(pyramid_levels);
    {
        Ptr<ColorGradientPyramid> qp = modality->process(source, object_mask);
        if (num_features > 0)
            qp->num_features = num_features;
        for (int l = 0; l < pyramid_levels; ++l)
        {
            if (l > 0)
                qp->pyrDown();
            const bool success = qp->extractTemplate(tp[l]);
            tp[l].sscale = sscale;
            tp[l].orientation = orientation;
            tp[l].tagFieldID = tagFieldID;
            tp[l].fiducial_src = fiducial_src;
            if (!success)
                return -1;
        }
    }
    cropTemplates(tp);
    template_pyramids.push_back(tp);
    templates_dirty_ = true;
    return template_id;
}

// :1395-1451 -- rotate the features of template zero_id about `center` instead of re-training on a rotated image
int Detector::addTemplate_rotate(const std::string &class_id, int zero_id, float theta, Point2f center)
{
    std::vector<TemplatePyramid> &template_pyramids = class_templates[class_id];
    const int template_id = static_cast<int>(template_pyramids.size());
    CV_Assert(zero_id >= 0 && zero_id < template_id);
    const TemplatePyramid &base = template_pyramids[zero_id];
    TemplatePyramid tp(pyramid_levels);
    const double ang = -theta / 180 * CV_PI;
    for (int l = 0; l < pyramid_levels; ++l)
    {
        if (l > 0)
            center /= 2;
        for (const Feature &f : base[l].features)
        {
            // rotation in float, like cv::Point2f arithmetic in the reference (:1395-1409)
            const Point2f d(f.x + base[l].tl_x - center.x, f.y + base[l].tl_y - center.y);
            Point2f r;
            r.x = (float)(std::cos(ang) * d.x - std::sin(ang) * d.y);
            r.y = (float)(std::sin(ang) * d.x + std::cos(ang) * d.y);
            r = r + center;
            Feature g;
            g.x = int(r.x + 0.5f);
            g.y = int(r.y + 0.5f);
            g.theta = f.theta - theta;
            while (g.theta > 360)
                g.theta -= 360;
            while (g.theta < 0)
                g.theta += 360;
            g.label = int(g.theta * 16 / 360 + 0.5f);
            g.label &= 7;
            tp[l].features.push_back(g);
        }
        tp[l].pyramid_level = l;
    }
    cropTemplates(tp);
    template_pyramids.push_back(tp);
    templates_dirty_ = true;
    return template_id;
}

const std::vector<Template> &Detector::getTemplates(const std::string &class_id, int template_id) const
{
    TemplatesMap::const_iterator i = class_templates.find(class_id);
    CV_Assert(i != class_templates.end());
    CV_Assert(i->second.size() > size_t(template_id));
    return i->second[template_id];
}

int Detector::numTemplates() const
{
    int ret = 0;
    for (const auto &kv : class_templates)
        ret += static_cast<int>(kv.second.size());
    return ret;
}

int Detector::numTemplates(const std::string &class_id) const
{
    TemplatesMap::const_iterator i = class_templates.find(class_id);
    return i == class_templates.end() ? 0 : static_cast<int>(i->second.size());
}

std::vector<std::string> Detector::classIds() const
{
    std::vector<std::string> ids;
    for (const auto &kv : class_templates)
        ids.push_back(kv.first);
    return ids;
}

void Detector::read(const FileNode &fn)
{
    class_templates.clear();
    pyramid_levels = fn["pyramid_levels"];
    fn["T"] >> T_at_level;
    modality = makePtr<ColorGradient>();
    modality->read(fn);
    if (ctx_) // parameters baked into the device context changed
    {
        sbm_destroy(ctx_);
        ctx_ = nullptr;
    }
    templates_dirty_ = true;
}

void Detector::write(FileStorage &fs) const
{
    fs << "pyramid_levels" << pyramid_levels;
    fs << "T" << T_at_level;
    modality->write(fs);
}

std::string Detector::readClass(const FileNode &fn, const std::string &class_id_override)
{
    std::string class_id;
    if (class_id_override.empty())
    {
        std::string class_id_tmp = fn["class_id"];
        CV_Assert(class_templates.find(class_id_tmp) == class_templates.end()); // must be a new class
        class_id = class_id_tmp;
    }
    else
        class_id = class_id_override;

    std::vector<TemplatePyramid> tps;
    FileNode tps_fn = fn["template_pyramids"];
    tps.resize(tps_fn.size());
    int expected_id = 0;
    for (FileNodeIterator it = tps_fn.begin(), end = tps_fn.end(); it != end; ++it, ++expected_id)
    {
        int template_id = (*it)["template_id"];
        CV_Assert(template_id == expected_id);
        FileNode templates_fn = (*it)["templates"];
        tps[template_id].resize(templates_fn.size());
        int idx = 0;
        for (FileNodeIterator t = templates_fn.begin(), tend = templates_fn.end(); t != tend; ++t)
            tps[template_id][idx++].read(*t);
    }
    class_templates.insert(TemplatesMap::value_type(class_id, tps)); // like std::map::insert: an existing key wins
    templates_dirty_ = true;
    return class_id;
}

void Detector::writeClass(const std::string &class_id, FileStorage &fs) const
{
    TemplatesMap::const_iterator it = class_templates.find(class_id);
    CV_Assert(it != class_templates.end());
    const std::vector<TemplatePyramid> &tps = it->second;
    fs << "class_id" << it->first;
    fs << "pyramid_levels" << pyramid_levels;
    fs << "template_pyramids"
       << "[";
    for (size_t i = 0; i < tps.size(); ++i)
    {
        fs << "{";
        fs << "template_id" << int(i);
        fs << "templates"
           << "[";
        for (const Template &t : tps[i])
        {
            fs << "{";
            t.write(fs);
            fs << "}";
        }
        fs << "]";
        fs << "}";
    }
    fs << "]";
}

void Detector::readClasses(const std::vector<std::string> &class_ids, const std::string &format)
{
    for (const std::string &class_id : class_ids)
    {
        FileStorage fs(cv::format(format.c_str(), class_id.c_str()), FileStorage::READ);
        readClass(fs.root());
    }
}

void Detector::writeClasses(const std::string &format) const
{
    for (const auto &kv : class_templates)
    {
        FileStorage fs(cv::format(format.c_str(), kv.first.c_str()), FileStorage::WRITE);
        writeClass(kv.first, fs);
    }
}

} // namespace line2Dup
