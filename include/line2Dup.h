// line2Dup.h -- drop-in public header of the B200 build of line2Dup.
//
// Same namespaces, class names, member names and signatures as /root/reference/line2Dup.h:113-460, so a
// translation unit written against the reference compiles against this header unchanged; link
// libline2Dup_b200.so (host glue, shape_based_matching_b200/csrc/line2Dup.cpp) + libsbm_b200.so (CUDA, C ABI in
// include/sbm_capi.h).  Detector::match runs entirely on the GPU; template bookkeeping, training's feature
// selection and (de)serialisation are host code.  There is no CPU matching path: constructing a Detector
// without a CUDA device throws cv::Exception.
//
// Not carried over: the Timer/ScopeTimer helpers and the csv-parser include (line2Dup.h:13-111, SURVEY §2 rows
// 12/14, out of scope) and the unused MIPP include.
#ifndef SBM_B200_LINE2DUP_H
#define SBM_B200_LINE2DUP_H

#if defined(__has_include)
#if __has_include(<opencv2/core/core.hpp>) && !defined(SBM_FORCE_CV_COMPAT)
#include <opencv2/core/core.hpp>
#include <opencv2/imgproc.hpp>
#define SBM_HAVE_OPENCV 1
#endif
#endif
#ifndef SBM_HAVE_OPENCV
#include "sbm/cv_compat.hpp"
#include "sbm/imgproc_compat.hpp"
#endif

#include <cassert>
#include <cfloat>
#include <cmath>
#include <map>
#include <memory>
#include <string>
#include <vector>

#define ANGLE_TOLERANCE FLT_EPSILON // line2Dup.h:8

struct sbm_ctx; // include/sbm_capi.h

namespace line2Dup
{

// line2Dup.h:116-129
struct Feature
{
    int x;
    int y;
    int label;
    float theta;

    void read(const cv::FileNode &fn);
    void write(cv::FileStorage &fs) const;

    Feature() : x(0), y(0), label(0), theta(0) {}
    Feature(int x, int y, int label);
};
inline Feature::Feature(int _x, int _y, int _label) : x(_x), y(_y), label(_label), theta(0) {}

// line2Dup.h:131-153
struct Template
{
    int width;
    int height;
    int tl_x;
    int tl_y;
    int pyramid_level;
    std::vector<Feature> features;

    float sscale;
    float orientation;

    int tagFieldID;
    std::string fiducial_src;

    Template() : width(0), height(0), tl_x(0), tl_y(0), pyramid_level(0), sscale(0), orientation(0), tagFieldID(0) {}
    void read(const cv::FileNode &fn);
    void write(cv::FileStorage &fs) const;
};

// line2Dup.h:155-200.  Training-side object: update() computes magnitude / angle / angle_ori of the current pyramid
// level on the GPU (sbm_quantized_orientations); extractTemplate() is host code.
class ColorGradientPyramid
{
public:
    ColorGradientPyramid(const cv::Mat &src, const cv::Mat &mask, float weak_threshold, size_t num_features,
                         float strong_threshold);

    void quantize(cv::Mat &dst) const;

    bool extractTemplate(Template &templ) const;

    void pyrDown();

public:
    void update();
    /// Candidate feature with a score
    struct Candidate
    {
        Candidate(int x, int y, int label, float score);

        /// Sort candidates with high score to the front
        bool operator<(const Candidate &rhs) const { return score > rhs.score; }

        Feature f;
        float score;
    };

    cv::Mat src;
    cv::Mat mask;

    int pyramid_level;
    cv::Mat angle;
    cv::Mat magnitude;
    cv::Mat angle_ori;

    float weak_threshold;
    size_t num_features;
    float strong_threshold;
    static bool selectScatteredFeatures(const std::vector<Candidate> &candidates, std::vector<Feature> &features,
                                        size_t num_features, float distance);

private:
    void updateFrom(int pyrdown_steps); // update() of `src` taken through that many cv::pyrDown steps on the device
};
inline ColorGradientPyramid::Candidate::Candidate(int x, int y, int label, float _score) : f(x, y, label), score(_score) {}

// line2Dup.h:202-220
class ColorGradient
{
public:
    ColorGradient();
    ColorGradient(float weak_threshold, size_t num_features, float strong_threshold);

    std::string name() const;

    float weak_threshold;
    size_t num_features;
    float strong_threshold;
    void read(const cv::FileNode &fn);
    void write(cv::FileStorage &fs) const;

    cv::Ptr<ColorGradientPyramid> process(const cv::Mat src, const cv::Mat &mask = cv::Mat()) const
    {
        return cv::makePtr<ColorGradientPyramid>(src, mask, weak_threshold, num_features, strong_threshold);
    }
};

// line2Dup.h:222-255
struct Match
{
    Match() {}

    Match(int x, int y, float similarity, const std::string &class_id, int template_id);

    /// Sort matches with high similarity to the front
    bool operator<(const Match &rhs) const
    {
        // Secondarily sort on template_id for the sake of duplicate removal
        if (similarity != rhs.similarity)
            return similarity > rhs.similarity;
        else
            return template_id < rhs.template_id;
    }

    bool operator==(const Match &rhs) const
    {
        return x == rhs.x && y == rhs.y && similarity == rhs.similarity && class_id == rhs.class_id;
    }

    int x;
    int y;
    float similarity;
    std::string class_id;
    int template_id;
};

inline Match::Match(int _x, int _y, float _similarity, const std::string &_class_id, int _template_id)
    : x(_x), y(_y), similarity(_similarity), class_id(_class_id), template_id(_template_id)
{
}

// line2Dup.h:257-333
class Detector
{
public:
    /// Empty constructor, initialize with read().
    Detector();

    Detector(std::vector<int> T);
    Detector(int num_features, std::vector<int> T, float weak_thresh = 30.0f, float strong_thresh = 60.0f);
    ~Detector();
    Detector(const Detector &) = delete;
    Detector &operator=(const Detector &) = delete;

    std::vector<Match> match(cv::Mat sources, float threshold,
                             const std::vector<std::string> &class_ids = std::vector<std::string>(),
                             const cv::Mat masks = cv::Mat()) const;

    /// Streaming front door: match() over equally sized frames, several frames in flight
    /// (sbm_match_batch).  result[i] is exactly match(sources[i], ...).
    std::vector<std::vector<Match>> matchBatch(const std::vector<cv::Mat> &sources, float threshold,
                                               const std::vector<std::string> &class_ids = std::vector<std::string>()) const;

    int addTemplate(const cv::Mat sources, const std::string &class_id, const cv::Mat &object_mask, float sscale = -1.0,
                    float orientation = -1.0, int tagFieldID = 0, std::string fiducial_src = "none", int num_features = 0);

    int addTemplate_rotate(const std::string &class_id, int zero_id, float theta, cv::Point2f center);

    const cv::Ptr<ColorGradient> &getModalities() const { return modality; }

    int getT(int pyramid_level) const { return T_at_level[pyramid_level]; }

    int pyramidLevels() const { return pyramid_levels; }

    const std::vector<Template> &getTemplates(const std::string &class_id, int template_id) const;

    int numTemplates() const;
    int numTemplates(const std::string &class_id) const;
    int numClasses() const { return static_cast<int>(class_templates.size()); }

    std::vector<std::string> classIds() const;

    void read(const cv::FileNode &fn);
    void write(cv::FileStorage &fs) const;

    std::string readClass(const cv::FileNode &fn, const std::string &class_id_override = "");
    void writeClass(const std::string &class_id, cv::FileStorage &fs) const;

    void readClasses(const std::vector<std::string> &class_ids, const std::string &format = "templates_%s.yml.gz");
    void writeClasses(const std::string &format = "templates_%s.yml.gz") const;

    /// CUDA device ordinal used by this detector (default 0; set before the first match/addTemplate).
    void setDevice(int device);

    /// Multi-GPU (sbm_group_* in include/sbm_capi.h): this detector becomes rank `rank` of `n_ranks`
    /// detectors, one per GPU of a box (one process or thread each), joined by one NCCL communicator.  Rank 0 calls
    /// groupUniqueId() and hands the 128 bytes to the others (file, pipe, MPI ...); every rank then calls setGroup
    /// (collective).  GROUP_TEMPLATES[_CONTIGUOUS]: the template loop of matchClass (line2Dup.cpp:1166-1170) is spread
    /// over the GPUs -- every rank keeps all its addTemplate / readClasses calls and passes the same frame, match()
    /// returns the unsharded list on every rank.  GROUP_FRAMES: the caller's frame loop (test_jabil.cpp:314-374) is
    /// spread -- matchBatch() takes this rank's frames and returns the lists of all ranks' frames, rank-major.
    enum GroupMode
    {
        GROUP_TEMPLATES = 0,
        GROUP_TEMPLATES_CONTIGUOUS = 1,
        GROUP_FRAMES = 2
    };
    static std::vector<unsigned char> groupUniqueId();
    void setGroup(int n_ranks, int rank, const std::vector<unsigned char> &unique_id, GroupMode mode = GROUP_TEMPLATES);
    int groupSize() const { return group_n_; }

protected:
    cv::Ptr<ColorGradient> modality;
    int pyramid_levels;
    std::vector<int> T_at_level;

    typedef std::vector<Template> TemplatePyramid;
    typedef std::map<std::string, std::vector<TemplatePyramid>> TemplatesMap;
    TemplatesMap class_templates;

private:
    // device side: created lazily, re-uploaded when the template set changed (match() stays logically const)
    sbm_ctx *context() const;
    void syncTemplates() const;
    mutable sbm_ctx *ctx_;
    mutable bool templates_dirty_;
    mutable std::vector<std::string> slots_; // class id of every uploaded class slot (std::map order)
    int device_;
    int group_n_ = 0, group_mode_ = -1; // 0: not in a group
};

} // namespace line2Dup

namespace shape_based_matching
{
// line2Dup.h:344-460
class shapeInfo_producer
{
public:
    cv::Mat src;
    cv::Mat mask;

    std::vector<float> angle_range;
    std::vector<float> scale_range;

    float angle_step = 15;
    float scale_step = 0.5;
    float eps = 0.00001f;

    class Info
    {
    public:
        float angle;
        float scale;

        Info(float angle_, float scale_)
        {
            angle = angle_;
            scale = scale_;
        }
    };
    std::vector<Info> infos;

    shapeInfo_producer(cv::Mat src, cv::Mat mask = cv::Mat())
    {
        this->src = src;
        if (mask.empty())
            this->mask = cv::Mat(src.size(), CV_8UC1, cv::Scalar(255)); // make sure we have masks
        else
            this->mask = mask;
    }

    // this fork: right-angle rotations + resize only (line2Dup.h:379-405): an angle within ANGLE_TOLERANCE of 90, 180
    // or 270 degrees turns the image by that quarter, any other angle leaves it as it is; then cv::resize by `scale`
    static cv::Mat transform(cv::Mat src, float angle, float scale)
    {
        static const struct
        {
            double degrees;
            int code;
        } quarter[3] = {{90.0, cv::ROTATE_90_CLOCKWISE}, {180.0, cv::ROTATE_180}, {270.0, cv::ROTATE_90_COUNTERCLOCKWISE}};
        cv::Mat turned = src;
        for (const auto &q : quarter)
            if (std::abs(angle - q.degrees) < ANGLE_TOLERANCE)
            {
                cv::rotate(src, turned, q.code);
                break;
            }
        cv::Mat dst;
        cv::resize(turned, dst, cv::Size(), scale, scale);
        return dst;
    }

    void produce_infos()
    {
        infos.clear();

        assert(angle_range.size() <= 2);
        assert(scale_range.size() <= 2);
        assert(angle_step > eps * 10);
        assert(scale_step > eps * 10);

        if (angle_range.size() == 0)
            angle_range.push_back(0);
        if (scale_range.size() == 0)
            scale_range.push_back(1);

        const bool one_angle = angle_range.size() == 1, one_scale = scale_range.size() == 1;
        if (!one_scale)
            assert(scale_range[1] > scale_range[0]);
        if (!one_angle)
            assert(angle_range[1] > angle_range[0]);
        // scale is the outer loop, angle the inner one (line2Dup.h:440-447); float accumulation as in the reference
        const float s_hi = one_scale ? scale_range[0] : scale_range[1] + eps;
        const float a_hi = one_angle ? angle_range[0] : angle_range[1] + eps;
        for (float scale = scale_range[0]; scale <= s_hi; scale += scale_step)
        {
            for (float angle = angle_range[0]; angle <= a_hi; angle += angle_step)
            {
                infos.emplace_back(angle, scale);
                if (one_angle)
                    break;
            }
            if (one_scale)
                break;
        }
    }

    cv::Mat src_of(const Info &info) { return transform(src, info.angle, info.scale); }

    cv::Mat mask_of(const Info &info) { return (transform(mask, info.angle, info.scale) > 0); }
};

} // namespace shape_based_matching

#endif // SBM_B200_LINE2DUP_H
