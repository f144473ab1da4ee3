// sbm_capi.cu -- host side of the C ABI declared in include/sbm_capi.h.
//
// Owns the device context of one Detector: template SoA per pyramid level, per-shape scratch (pyramid
// sources, quantised images, linear memories, candidate ping-pong buffers) and the stream the seven
// kernels of a frame are enqueued on.  The only CPU work per frame is ordering the (tiny) candidate list
// the way the reference's serial loop produces it and running the same std::sort / std::unique
// (line2Dup.cpp:1143-1145).  There is no CPU implementation of any pipeline stage in this library.
#include "sbm_capi.h"
#include "sbm_kernels.cuh"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

using namespace sbm;

static_assert(sizeof(Cand) == sizeof(sbm_device_rec), "Cand must mirror sbm_device_rec");

namespace
{
thread_local std::string g_err;

int fail(int code, const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CK(call)                                                                                      \
    do                                                                                                \
    {                                                                                                 \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess)                                                                        \
            return fail(SBM_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

struct ClassSet
{
    int n = 0;
    std::vector<sbm_template_desc> desc; // [n][levels]
    std::vector<sbm_feature> feats;
};

struct LevelBuf
{
    int T = 0, w = 0, h = 0, W = 0, H = 0;
    uint8_t *d_src = nullptr; // level >= 1 (level 0 is the frame itself)
    uint8_t *d_quant = nullptr;
    uint8_t *d_lm = nullptr;
    size_t lm_bytes = 0, lm_alloc = 0;
    uint8_t *d_spread = nullptr, *d_resp = nullptr; // debug taps
    // templates at this level
    TemplLevel *d_tl = nullptr;
    Feat *d_feats = nullptr;
};

template <typename T> void freeDev(T *&p)
{
    if (p)
        cudaFree(p);
    p = nullptr;
}
} // namespace

struct sbm_ctx
{
    int device = 0;
    std::vector<int> T;
    float weak = 30.f, strong = 60.f;
    int num_features = 63;
    int levels = 0;
    cudaStream_t own_stream = nullptr;
    std::mutex mu;

    std::vector<ClassSet> classes;
    std::vector<uint32_t> gid_begin; // per class slot (+ total)
    bool templates_dirty = true;
    uint32_t n_templates = 0;

    // frame shape
    int w = 0, h = 0, c = 0;
    bool have_frame = false;
    std::vector<LevelBuf> lv;
    uint8_t *d_frame = nullptr; // staging for host frames
    size_t frame_alloc = 0;
    uint8_t *d_mask = nullptr;
    size_t mask_alloc = 0;
    const uint8_t *last_src0 = nullptr;

    // coarse level, per (templates, shape)
    bool coarse_dirty = true;
    TemplCoarse *d_tcoarse = nullptr;
    uint32_t *d_coff = nullptr;
    std::vector<TemplCoarse> h_tcoarse;

    // active template list
    std::vector<int> active_classes; // as requested (may repeat)
    bool active_dirty = true;
    uint32_t *d_active = nullptr;
    uint32_t n_active = 0;
    uint64_t coarse_evals = 0;

    // candidates
    uint32_t cap = 1u << 18;
    Cand *d_buf[2] = {nullptr, nullptr};
    uint32_t *d_counts = nullptr; // [0..levels-1] stage counts, [8] overflow, [10..11] u64 local evals
    Cand *d_final = nullptr;      // where the last stage of the most recent frame wrote
    uint32_t *d_final_count = nullptr;
    uint32_t final_cap = 0;

    // host staging (pinned)
    uint32_t *h_counts = nullptr;
    Cand *h_recs = nullptr;
    uint32_t h_recs_cap = 0;

    bool debug = false, timing = false;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    float last_ms[4] = {0, 0, 0, 0};
    uint64_t stats[8] = {0};
    int launches = 0;
};

namespace
{
constexpr int COUNT_WORDS = 16;
constexpr uint32_t HEAD_RECS = 2048;

void freeShape(sbm_ctx *c)
{
    for (auto &l : c->lv)
    {
        freeDev(l.d_src);
        freeDev(l.d_quant);
        freeDev(l.d_lm);
        freeDev(l.d_spread);
        freeDev(l.d_resp);
    }
}

int checkShape(sbm_ctx *c, int w, int h, int ch)
{
    if (ch != 1 && ch != 3)
        return fail(SBM_ERR_BAD_ARG, "source must have 1 or 3 channels (got %d)", ch);
    if (w <= 0 || h <= 0)
        return fail(SBM_ERR_BAD_ARG, "empty source");
    int lw = w, lh = h;
    for (int l = 0; l < c->levels; ++l)
    {
        int T = c->T[l];
        if (l > 0)
            lw /= 2, lh /= 2;
        // CV_Assert((src.rows * src.cols) % 16 == 0)  line2Dup.cpp:639;  rows % T, cols % T  :751-752
        if (lw <= 0 || lh <= 0 || (lw % T) || (lh % T) || ((size_t)lw * lh) % 16)
            return fail(SBM_ERR_BAD_ARG,
                        "level %d is %dx%d: rows and cols must be multiples of T=%d and rows*cols of 16 "
                        "(line2Dup.cpp:639,751-752)", l, lw, lh, T);
        if (lw > 16383 || lh > 16383)
            return fail(SBM_ERR_BAD_ARG, "image too large");
    }
    return 0;
}

int ensureShape(sbm_ctx *c, int w, int h, int ch)
{
    if (c->w == w && c->h == h && c->c == ch && !c->lv.empty() && c->lv[0].d_quant &&
        (!c->debug || c->lv[0].d_spread))
        return 0;
    int rc = checkShape(c, w, h, ch);
    if (rc)
        return rc;
    freeShape(c);
    c->lv.resize(c->levels);
    int lw = w, lh = h;
    for (int l = 0; l < c->levels; ++l)
    {
        LevelBuf &L = c->lv[l];
        if (l > 0)
            lw /= 2, lh /= 2;
        TemplLevel *keep_tl = L.d_tl;
        Feat *keep_f = L.d_feats;
        L.T = c->T[l], L.w = lw, L.h = lh, L.W = lw / L.T, L.H = lh / L.T;
        L.d_tl = keep_tl, L.d_feats = keep_f;
        size_t px = (size_t)lw * lh;
        if (l > 0)
            CK(cudaMalloc(&L.d_src, px * ch));
        CK(cudaMalloc(&L.d_quant, px));
        L.lm_bytes = 8 * px;
        // tail slack: flat addressing lets the last memory be over-read: a whole memory for
        // the coarse scan past template_positions, 16 rows + 16 for the local patch, +8 for the realigning loads
        size_t WH = (size_t)L.W * L.H;
        L.lm_alloc = L.lm_bytes + std::max(WH, (size_t)16 * L.W + 16) + 128;
        CK(cudaMalloc(&L.d_lm, L.lm_alloc));
        CK(cudaMemset(L.d_lm, 0, L.lm_alloc));
        if (c->debug)
        {
            CK(cudaMalloc(&L.d_spread, px));
            CK(cudaMalloc(&L.d_resp, 8 * px));
        }
    }
    c->w = w, c->h = h, c->c = ch;
    c->coarse_dirty = true;
    c->have_frame = false;
    return 0;
}

int ensureTemplates(sbm_ctx *c)
{
    if (c->templates_dirty)
    {
        const int L = c->levels;
        c->gid_begin.assign(c->classes.size() + 1, 0);
        for (size_t k = 0; k < c->classes.size(); ++k)
            c->gid_begin[k + 1] = c->gid_begin[k] + (uint32_t)c->classes[k].n;
        c->n_templates = c->gid_begin.back();
        if (c->lv.size() != (size_t)L)
            c->lv.resize(L);
        for (int l = 0; l < L; ++l)
        {
            std::vector<TemplLevel> tl(c->n_templates);
            std::vector<Feat> ff;
            uint32_t gid = 0;
            for (auto &cs : c->classes)
            {
                size_t fo = 0;
                for (int t = 0; t < cs.n; ++t)
                    for (int ll = 0; ll < L; ++ll)
                    {
                        const sbm_template_desc &d = cs.desc[(size_t)t * L + ll];
                        if (ll == l)
                        {
                            TemplLevel &o = tl[gid++];
                            o.feat_begin = (uint32_t)ff.size();
                            o.nf = d.num_features;
                            o.width = d.width, o.height = d.height;
                            for (int i = 0; i < d.num_features; ++i)
                            {
                                const sbm_feature &f = cs.feats[fo + i];
                                Feat g;
                                g.x = (int16_t)std::max(-32768, std::min(32767, f.x));
                                g.y = (int16_t)std::max(-32768, std::min(32767, f.y));
                                g.label = f.label;
                                ff.push_back(g);
                            }
                        }
                        fo += d.num_features;
                    }
            }
            LevelBuf &LB = c->lv[l];
            freeDev(LB.d_tl);
            freeDev(LB.d_feats);
            CK(cudaMalloc(&LB.d_tl, std::max<size_t>(1, tl.size()) * sizeof(TemplLevel)));
            CK(cudaMalloc(&LB.d_feats, std::max<size_t>(1, ff.size()) * sizeof(Feat)));
            if (!tl.empty())
                CK(cudaMemcpy(LB.d_tl, tl.data(), tl.size() * sizeof(TemplLevel), cudaMemcpyHostToDevice));
            if (!ff.empty())
                CK(cudaMemcpy(LB.d_feats, ff.data(), ff.size() * sizeof(Feat), cudaMemcpyHostToDevice));
        }
        c->templates_dirty = false;
        c->coarse_dirty = true;
        c->active_dirty = true;
    }
    if (c->coarse_dirty && c->w > 0)
    {
        // similarity()'s per-template constants and per-feature start offsets for this frame shape
        // (line2Dup.cpp:813-838, accessLinearMemory :782-805)
        const int L = c->levels;
        const LevelBuf &LB = c->lv[L - 1];
        const int T = LB.T, W = LB.W, H = LB.H;
        const uint32_t WH = (uint32_t)W * H;
        c->h_tcoarse.assign(c->n_templates, TemplCoarse());
        std::vector<uint32_t> off;
        uint32_t gid = 0;
        for (auto &cs : c->classes)
        {
            size_t fo = 0;
            for (int t = 0; t < cs.n; ++t)
                for (int ll = 0; ll < L; ++ll)
                {
                    const sbm_template_desc &d = cs.desc[(size_t)t * L + ll];
                    if (ll == L - 1)
                    {
                        TemplCoarse &o = c->h_tcoarse[gid++];
                        o.off_begin = (uint32_t)off.size();
                        o.nf = d.num_features;
                        int wf = (d.width - 1) / T + 1, hf = (d.height - 1) / T + 1;
                        o.tp = (H - hf) * W + (W - wf) + 1;
                        // in-image features only (line2Dup.cpp:836), bucketed by the word shift of their start
                        // offset so that k_similarity runs four branch-free loops (integer adds commute)
                        std::vector<uint32_t> cls[4];
                        for (int i = 0; i < d.num_features; ++i)
                        {
                            const sbm_feature &f = cs.feats[fo + i];
                            if (f.x < 0 || f.x >= LB.w || f.y < 0 || f.y >= LB.h)
                                continue;
                            if (f.label < 0 || f.label > 7)
                                return fail(SBM_ERR_BAD_ARG, "feature label %d out of range", f.label);
                            uint32_t a = (uint32_t)f.label * T * T * WH + (uint32_t)((f.y % T) * T + f.x % T) * WH +
                                         (uint32_t)(f.y / T) * W + (uint32_t)(f.x / T);
                            cls[(a >> 2) & 3].push_back(a);
                        }
                        int nv = 0, stored = 0;
                        for (int k = 0; k < 4; ++k)
                        {
                            nv += (int)cls[k].size();
                            // pad with reads of the zeroed tail (same word shift): adds 0, keeps batches full
                            while (cls[k].size() % K5_U)
                                cls[k].push_back((uint32_t)LB.lm_bytes + 4u * k);
                            off.insert(off.end(), cls[k].begin(), cls[k].end());
                            stored += (int)cls[k].size();
                            o.cls_end[k] = (uint16_t)stored;
                        }
                        o.n_valid = nv;
                    }
                    fo += d.num_features;
                }
        }
        freeDev(c->d_tcoarse);
        freeDev(c->d_coff);
        CK(cudaMalloc(&c->d_tcoarse, std::max<size_t>(1, c->h_tcoarse.size()) * sizeof(TemplCoarse)));
        CK(cudaMalloc(&c->d_coff, std::max<size_t>(1, off.size()) * sizeof(uint32_t)));
        if (!c->h_tcoarse.empty())
            CK(cudaMemcpy(c->d_tcoarse, c->h_tcoarse.data(), c->h_tcoarse.size() * sizeof(TemplCoarse), cudaMemcpyHostToDevice));
        if (!off.empty())
            CK(cudaMemcpy(c->d_coff, off.data(), off.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        c->coarse_dirty = false;
        c->active_dirty = true;
    }
    return 0;
}

int ensureActive(sbm_ctx *c, const int *class_idx, int n_classes)
{
    std::vector<int> req;
    if (!class_idx || n_classes <= 0)
    {
        for (size_t k = 0; k < c->classes.size(); ++k)
            if (c->classes[k].n > 0)
                req.push_back((int)k);
    }
    else
    {
        for (int i = 0; i < n_classes; ++i) // unknown ids are silently skipped (line2Dup.cpp:1136-1138)
            if (class_idx[i] >= 0 && (size_t)class_idx[i] < c->classes.size() && c->classes[class_idx[i]].n > 0)
                req.push_back(class_idx[i]);
    }
    if (!c->active_dirty && req == c->active_classes)
        return 0;
    c->active_classes = req;
    std::vector<int> uniq = req;
    std::sort(uniq.begin(), uniq.end());
    uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
    std::vector<uint32_t> act;
    c->coarse_evals = 0;
    for (int k : uniq)
        for (uint32_t g = c->gid_begin[k]; g < c->gid_begin[k + 1]; ++g)
        {
            act.push_back(g);
            const TemplCoarse &t = c->h_tcoarse[g];
            c->coarse_evals += (uint64_t)t.n_valid * (uint64_t)std::max(t.tp, 0);
        }
    freeDev(c->d_active);
    CK(cudaMalloc(&c->d_active, std::max<size_t>(1, act.size()) * sizeof(uint32_t)));
    if (!act.empty())
        CK(cudaMemcpy(c->d_active, act.data(), act.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    c->n_active = (uint32_t)act.size();
    c->active_dirty = false;
    return 0;
}

int ensureCandBuffers(sbm_ctx *c)
{
    if (!c->d_buf[0])
    {
        CK(cudaMalloc(&c->d_buf[0], (size_t)c->cap * sizeof(Cand)));
        CK(cudaMalloc(&c->d_buf[1], (size_t)c->cap * sizeof(Cand)));
    }
    if (!c->d_counts)
    {
        CK(cudaMalloc(&c->d_counts, COUNT_WORDS * sizeof(uint32_t)));
        CK(cudaMallocHost(&c->h_counts, COUNT_WORDS * sizeof(uint32_t)));
    }
    return 0;
}

// Enqueue the whole pipeline for one frame resident at d_img.  user_recs/user_count: optional caller
// buffers for the last stage.
int enqueueFrame(sbm_ctx *c, const uint8_t *d_img, const uint8_t *d_mask, float threshold, cudaStream_t st, Cand *user_recs,
                 uint32_t user_cap, uint32_t *user_count)
{
    const int L = c->levels;
    int launches = 0;
    CK(cudaMemsetAsync(c->d_counts, 0, COUNT_WORDS * sizeof(uint32_t), st));
    if (user_count)
        CK(cudaMemsetAsync(user_count, 0, 2 * sizeof(uint32_t), st));
    if (c->timing)
        CK(cudaEventRecord(c->ev[0], st));
    const float weak_sq = c->weak * c->weak;
    const uint8_t *src = d_img;
    for (int l = 0; l < L; ++l)
    {
        LevelBuf &LB = c->lv[l];
        if (l > 0)
        {
            const LevelBuf &P = c->lv[l - 1];
            dim3 g((LB.w + 31) / 32, (LB.h + 7) / 8);
            if (c->c == 3)
                k_pyrdown<3><<<g, 256, 0, st>>>(src, P.w, P.h, LB.d_src, LB.w, LB.h);
            else
                k_pyrdown<1><<<g, 256, 0, st>>>(src, P.w, P.h, LB.d_src, LB.w, LB.h);
            src = LB.d_src;
            ++launches;
        }
        // 64x32 tiles unless that leaves the GPU under two waves of CTAs; then 64x16
        const bool tall = ((LB.w + K1_TW - 1) / K1_TW) * ((LB.h + 31) / 32) >= 2 * 148;
        const int th = tall ? 32 : 16;
        dim3 g1((LB.w + K1_TW - 1) / K1_TW, (LB.h + th - 1) / th);
        if (c->c == 3 && tall)
            k_orient<3, 32><<<g1, K1_THREADS, 0, st>>>(src, LB.w, LB.h, d_mask, l, c->w, weak_sq, LB.d_quant);
        else if (c->c == 3)
            k_orient<3, 16><<<g1, K1_THREADS, 0, st>>>(src, LB.w, LB.h, d_mask, l, c->w, weak_sq, LB.d_quant);
        else if (tall)
            k_orient<1, 32><<<g1, K1_THREADS, 0, st>>>(src, LB.w, LB.h, d_mask, l, c->w, weak_sq, LB.d_quant);
        else
            k_orient<1, 16><<<g1, K1_THREADS, 0, st>>>(src, LB.w, LB.h, d_mask, l, c->w, weak_sq, LB.d_quant);
        dim3 g2((LB.w + K2_PW - 1) / K2_PW, (LB.h + K2_PH - 1) / K2_PH);
        uint8_t *ds = c->debug ? LB.d_spread : nullptr, *dr = c->debug ? LB.d_resp : nullptr;
        if (LB.T == 4)
            k_spread_linearize<4><<<g2, K2_THREADS, 0, st>>>(LB.d_quant, LB.w, LB.h, LB.d_lm, ds, dr);
        else if (LB.T == 8)
            k_spread_linearize<8><<<g2, K2_THREADS, 0, st>>>(LB.d_quant, LB.w, LB.h, LB.d_lm, ds, dr);
        else if (LB.T == 2)
            k_spread_linearize<2><<<g2, K2_THREADS, 0, st>>>(LB.d_quant, LB.w, LB.h, LB.d_lm, ds, dr);
        else
            k_spread_linearize<16><<<g2, K2_THREADS, 0, st>>>(LB.d_quant, LB.w, LB.h, LB.d_lm, ds, dr);
        launches += 2;
    }
    if (c->timing)
        CK(cudaEventRecord(c->ev[1], st));

    // coarse similarity + threshold at the last level
    const LevelBuf &CL = c->lv[L - 1];
    const int HW = CL.W * CL.H;
    Cand *out = (L == 1 && user_recs) ? user_recs : c->d_buf[0];
    uint32_t out_cap = (L == 1 && user_recs) ? user_cap : c->cap;
    uint32_t *out_cnt = (L == 1 && user_count) ? user_count : c->d_counts + 0;
    if (c->n_active > 0)
    {
        dim3 g(c->n_active, (HW + K5_TILE - 1) / K5_TILE);
        k_similarity<<<g, K5_THREADS, 0, st>>>(CL.d_lm, c->d_coff, c->d_tcoarse, c->d_active, CL.W, HW, CL.T, threshold, out,
                                               out_cnt, out_cap, nullptr);
        ++launches;
    }
    if (c->timing)
        CK(cudaEventRecord(c->ev[3], st));
    Cand *in = out;
    uint32_t in_cap = out_cap;
    uint32_t *in_cnt = out_cnt;
    int stage = 1;
    for (int l = L - 2; l >= 0; --l, ++stage)
    {
        const LevelBuf &LB = c->lv[l];
        bool last = (l == 0);
        out = (last && user_recs) ? user_recs : c->d_buf[stage & 1];
        out_cap = (last && user_recs) ? user_cap : c->cap;
        out_cnt = (last && user_count) ? user_count : c->d_counts + stage;
        if (c->n_active > 0)
        {
            k_refine<<<148 * 4, K6_THREADS, 0, st>>>(LB.d_lm, LB.d_tl, LB.d_feats, in, in_cnt, in_cap, out, out_cnt, out_cap, LB.w,
                                                     LB.h, LB.T, threshold, nullptr, 0,
                                                     reinterpret_cast<unsigned long long *>(c->d_counts + 10));
            ++launches;
        }
        in = out, in_cap = out_cap, in_cnt = out_cnt;
    }
    if (c->timing)
        CK(cudaEventRecord(c->ev[2], st));
    c->d_final = in;
    c->d_final_count = in_cnt;
    c->final_cap = in_cap;
    c->launches = launches;
    CK(cudaGetLastError());
    c->have_frame = true;
    return 0;
}

int growCap(sbm_ctx *c, uint32_t need)
{
    uint32_t nc = c->cap;
    while (nc < need)
        nc *= 2;
    freeDev(c->d_buf[0]);
    freeDev(c->d_buf[1]);
    c->cap = nc;
    return ensureCandBuffers(c);
}

// (class occurrence order, template_id, coarse scan index) -> reference's serial candidate order
int orderRecords(sbm_ctx *c, const Cand *recs, uint32_t n, bool finalize, sbm_match_rec *out, int cap, int *n_out)
{
    std::vector<std::vector<const Cand *>> per_class(c->classes.size());
    for (uint32_t i = 0; i < n; ++i)
    {
        const Cand &r = recs[i];
        if (r.tgid >= c->n_templates)
            return fail(SBM_ERR_BAD_ARG, "record %u names template %u of %u", i, r.tgid, c->n_templates);
        size_t k = std::upper_bound(c->gid_begin.begin(), c->gid_begin.end(), r.tgid) - c->gid_begin.begin() - 1;
        per_class[k].push_back(&r);
    }
    for (auto &v : per_class)
        std::sort(v.begin(), v.end(), [](const Cand *a, const Cand *b) {
            return a->tgid != b->tgid ? a->tgid < b->tgid : a->key < b->key;
        });
    std::vector<sbm_match_rec> m;
    m.reserve(n);
    for (int k : c->active_classes)
        for (const Cand *r : per_class[k])
            m.push_back(sbm_match_rec{r->x, r->y, r->sim, k, (int32_t)(r->tgid - c->gid_begin[k])});
    if (finalize)
    {
        // Match::operator< and operator== (line2Dup.h:231-243), then the reference's own calls (:1143-1145)
        std::sort(m.begin(), m.end(), [](const sbm_match_rec &a, const sbm_match_rec &b) {
            if (a.similarity != b.similarity)
                return a.similarity > b.similarity;
            return a.template_id < b.template_id;
        });
        m.erase(std::unique(m.begin(), m.end(),
                            [](const sbm_match_rec &a, const sbm_match_rec &b) {
                                return a.x == b.x && a.y == b.y && a.similarity == b.similarity && a.class_idx == b.class_idx;
                            }),
                m.end());
    }
    *n_out = (int)m.size();
    if ((int)m.size() > cap)
        return fail(SBM_ERR_CAPACITY, "output holds %d matches, %zu needed", cap, m.size());
    if (!m.empty())
        memcpy(out, m.data(), m.size() * sizeof(sbm_match_rec));
    return 0;
}

int fetchRecords(sbm_ctx *c, cudaStream_t st, uint32_t *n_out)
{
    if (!c->h_recs)
    {
        c->h_recs_cap = std::max(c->cap, HEAD_RECS);
        CK(cudaMallocHost(&c->h_recs, (size_t)c->h_recs_cap * sizeof(Cand)));
    }
    // one round trip for the common case: counters + the first HEAD_RECS records together
    uint32_t head = std::min(HEAD_RECS, c->final_cap);
    CK(cudaMemcpyAsync(c->h_counts, c->d_counts, COUNT_WORDS * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(c->h_recs, c->d_final, (size_t)head * sizeof(Cand), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    uint32_t n = c->h_counts[c->d_final_count - c->d_counts];
    *n_out = n;
    for (int s = 0; s < c->levels; ++s)
        if (c->h_counts[s] > c->cap)
            return SBM_ERR_CAPACITY;
    if (n > head)
    {
        if (n > c->h_recs_cap)
        {
            cudaFreeHost(c->h_recs);
            c->h_recs_cap = n;
            CK(cudaMallocHost(&c->h_recs, (size_t)n * sizeof(Cand)));
            head = 0;
        }
        CK(cudaMemcpyAsync(c->h_recs + head, c->d_final + head, (size_t)(n - head) * sizeof(Cand), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    return 0;
}

int matchHost(sbm_ctx *c, const uint8_t *img, int w, int h, int ch, size_t stride, const uint8_t *mask, float threshold,
              const int *class_idx, int n_classes, bool finalize, sbm_match_rec *out, int cap, int *n_out)
{
    if (!c || !img || !n_out)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    int rc;
    if ((rc = ensureShape(c, w, h, ch)) || (rc = ensureTemplates(c)) || (rc = ensureActive(c, class_idx, n_classes)) ||
        (rc = ensureCandBuffers(c)))
        return rc;
    cudaStream_t st = c->own_stream;
    size_t row = (size_t)w * ch, bytes = row * h;
    if (c->frame_alloc < bytes)
    {
        freeDev(c->d_frame);
        CK(cudaMalloc(&c->d_frame, bytes));
        c->frame_alloc = bytes;
    }
    if (stride == 0 || stride == row)
        CK(cudaMemcpyAsync(c->d_frame, img, bytes, cudaMemcpyHostToDevice, st));
    else
        CK(cudaMemcpy2DAsync(c->d_frame, row, img, stride, row, h, cudaMemcpyHostToDevice, st));
    const uint8_t *dm = nullptr;
    if (mask)
    {
        size_t mb = (size_t)w * h;
        if (c->mask_alloc < mb)
        {
            freeDev(c->d_mask);
            CK(cudaMalloc(&c->d_mask, mb));
            c->mask_alloc = mb;
        }
        CK(cudaMemcpyAsync(c->d_mask, mask, mb, cudaMemcpyHostToDevice, st));
        dm = c->d_mask;
    }
    c->last_src0 = c->d_frame;
    for (int attempt = 0; attempt < 8; ++attempt)
    {
        if ((rc = enqueueFrame(c, c->d_frame, dm, threshold, st, nullptr, 0, nullptr)))
            return rc;
        uint32_t n = 0;
        rc = fetchRecords(c, st, &n);
        if (rc == SBM_ERR_CAPACITY)
        {
            uint32_t need = 0;
            for (int s = 0; s < c->levels; ++s)
                need = std::max(need, c->h_counts[s]);
            if ((rc = growCap(c, need)))
                return rc;
            continue;
        }
        if (rc)
            return rc;
        c->stats[0] = c->coarse_evals;
        c->stats[1] = c->h_counts[0];
        memcpy(&c->stats[2], &c->h_counts[10], sizeof(uint64_t));
        c->stats[3] = n;
        c->stats[4] = (uint64_t)c->launches;
        return orderRecords(c, c->h_recs, n, finalize, out, cap, n_out);
    }
    return fail(SBM_ERR_CAPACITY, "candidate buffers keep overflowing");
}
} // namespace

extern "C"
{
const char *sbm_last_error(void) { return g_err.c_str(); }

const char *sbm_build_info(void)
{
#define SBM_STR2(x) #x
#define SBM_STR(x) SBM_STR2(x)
    return "shape_based_matching_b200 csrc: nvcc " SBM_STR(__CUDACC_VER_MAJOR__) "." SBM_STR(__CUDACC_VER_MINOR__)
           " -gencode arch=compute_100a,code=sm_100a; kernels k_orient k_pyrdown k_spread_linearize k_similarity k_refine";
}

int sbm_create(const int *T, int levels, float weak, float strong, int num_features, int device, sbm_ctx **out)
{
    if (!out || !T || levels < 1 || levels > 6)
        return fail(SBM_ERR_BAD_ARG, "sbm_create: bad arguments (levels must be 1..6)");
    for (int l = 0; l < levels; ++l)
        if (T[l] != 2 && T[l] != 4 && T[l] != 8 && T[l] != 16)
            return fail(SBM_ERR_BAD_ARG, "T[%d]=%d: supported spread/stride values are 2, 4, 8, 16", l, T[l]);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(SBM_ERR_CUDA, "no CUDA device (%s): this library has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= ndev)
        return fail(SBM_ERR_BAD_ARG, "device %d of %d", device, ndev);
    CK(cudaSetDevice(device));
    sbm_ctx *c = new sbm_ctx();
    c->device = device;
    c->T.assign(T, T + levels);
    c->levels = levels;
    c->weak = weak, c->strong = strong, c->num_features = num_features;
    c->lv.resize(levels);
    CK(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 4; ++i)
        CK(cudaEventCreate(&c->ev[i]));
    *out = c;
    return 0;
}

int sbm_destroy(sbm_ctx *c)
{
    if (!c)
        return 0;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    freeShape(c);
    for (auto &l : c->lv)
    {
        freeDev(l.d_tl);
        freeDev(l.d_feats);
    }
    freeDev(c->d_frame);
    freeDev(c->d_mask);
    freeDev(c->d_tcoarse);
    freeDev(c->d_coff);
    freeDev(c->d_active);
    freeDev(c->d_buf[0]);
    freeDev(c->d_buf[1]);
    freeDev(c->d_counts);
    if (c->h_counts)
        cudaFreeHost(c->h_counts);
    if (c->h_recs)
        cudaFreeHost(c->h_recs);
    for (int i = 0; i < 4; ++i)
        if (c->ev[i])
            cudaEventDestroy(c->ev[i]);
    if (c->own_stream)
        cudaStreamDestroy(c->own_stream);
    delete c;
    return 0;
}

int sbm_set_class_templates(sbm_ctx *c, int class_idx, int n_templates, const sbm_template_desc *desc, const sbm_feature *feats)
{
    if (!c || class_idx < 0 || class_idx > 65535 || n_templates < 0 || (n_templates > 0 && (!desc || !feats)))
        return fail(SBM_ERR_BAD_ARG, "sbm_set_class_templates: bad arguments");
    std::lock_guard<std::mutex> lock(c->mu);
    size_t nd = (size_t)n_templates * c->levels, nf = 0;
    for (size_t i = 0; i < nd; ++i)
    {
        if (desc[i].num_features < 0)
            return fail(SBM_ERR_BAD_ARG, "negative feature count");
        if (desc[i].num_features >= 8192) // CV_Assert(templ.features.size() < 8192)  line2Dup.cpp:811, :1195
            return fail(SBM_ERR_BAD_ARG, "feature size too large (%d >= 8192)", desc[i].num_features);
        nf += desc[i].num_features;
    }
    if ((size_t)class_idx >= c->classes.size())
        c->classes.resize(class_idx + 1);
    ClassSet &cs = c->classes[class_idx];
    cs.n = n_templates;
    cs.desc.assign(desc, desc + nd);
    cs.feats.assign(feats, feats + nf);
    c->templates_dirty = true;
    return 0;
}

int sbm_clear_templates(sbm_ctx *c)
{
    if (!c)
        return fail(SBM_ERR_BAD_ARG, "null context");
    std::lock_guard<std::mutex> lock(c->mu);
    c->classes.clear();
    c->templates_dirty = true;
    return 0;
}

int sbm_num_templates(sbm_ctx *c, int class_idx)
{
    if (!c)
        return fail(SBM_ERR_BAD_ARG, "null context");
    if (class_idx < 0)
    {
        int n = 0;
        for (auto &cs : c->classes)
            n += cs.n;
        return n;
    }
    return (size_t)class_idx < c->classes.size() ? c->classes[class_idx].n : 0;
}

int sbm_match(sbm_ctx *c, const uint8_t *img, int w, int h, int ch, size_t stride, const uint8_t *mask, float threshold,
              const int *class_idx, int n_classes, sbm_match_rec *out, int cap, int *n_out)
{
    return matchHost(c, img, w, h, ch, stride, mask, threshold, class_idx, n_classes, true, out, cap, n_out);
}

int sbm_match_candidates(sbm_ctx *c, const uint8_t *img, int w, int h, int ch, size_t stride, const uint8_t *mask,
                         float threshold, const int *class_idx, int n_classes, sbm_match_rec *out, int cap, int *n_out)
{
    return matchHost(c, img, w, h, ch, stride, mask, threshold, class_idx, n_classes, false, out, cap, n_out);
}

int sbm_match_device(sbm_ctx *c, const void *d_img, int w, int h, int ch, const void *d_mask, float threshold,
                     const int *class_idx, int n_classes, void *d_records, int cap, void *d_count, void *stream)
{
    if (!c || !d_img)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    int rc;
    if ((rc = ensureShape(c, w, h, ch)) || (rc = ensureTemplates(c)) || (rc = ensureActive(c, class_idx, n_classes)) ||
        (rc = ensureCandBuffers(c)))
        return rc;
    cudaStream_t st = stream ? (cudaStream_t)stream : c->own_stream;
    c->last_src0 = (const uint8_t *)d_img;
    return enqueueFrame(c, (const uint8_t *)d_img, (const uint8_t *)d_mask, threshold, st, (Cand *)d_records,
                        d_records ? (uint32_t)cap : 0, (uint32_t *)d_count);
}

int sbm_device_results(sbm_ctx *c, void **d_records, void **d_count, int *cap)
{
    if (!c || !c->have_frame)
        return fail(SBM_ERR_STATE, "no frame has been processed");
    if (d_records)
        *d_records = c->d_final;
    if (d_count)
        *d_count = c->d_final_count;
    if (cap)
        *cap = (int)c->final_cap;
    return 0;
}

int sbm_finalize_records(sbm_ctx *c, const sbm_device_rec *recs, int n, const int *class_idx, int n_classes, int finalize,
                         sbm_match_rec *out, int cap, int *n_out)
{
    if (!c || (n > 0 && !recs) || !n_out)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    int rc;
    if ((rc = ensureTemplates(c)) || (rc = ensureActive(c, class_idx, n_classes)))
        return rc;
    return orderRecords(c, reinterpret_cast<const Cand *>(recs), (uint32_t)n, finalize != 0, out, cap, n_out);
}

int sbm_set_debug(sbm_ctx *c, int enable)
{
    if (!c)
        return fail(SBM_ERR_BAD_ARG, "null context");
    c->debug = enable != 0;
    return 0;
}

int sbm_set_timing(sbm_ctx *c, int enable)
{
    if (!c)
        return fail(SBM_ERR_BAD_ARG, "null context");
    c->timing = enable != 0;
    return 0;
}

int sbm_get_intermediate(sbm_ctx *c, int level, int kind, void *dst, size_t bytes, size_t *bytes_needed)
{
    if (!c || level < 0 || level >= c->levels)
        return fail(SBM_ERR_BAD_ARG, "bad level");
    std::lock_guard<std::mutex> lock(c->mu);
    if (!c->have_frame)
        return fail(SBM_ERR_STATE, "no frame has been processed");
    CK(cudaSetDevice(c->device));
    const LevelBuf &L = c->lv[level];
    size_t px = (size_t)L.w * L.h, need = 0;
    const void *src = nullptr;
    switch (kind)
    {
    case SBM_QUANTIZED: need = px, src = L.d_quant; break;
    case SBM_SPREAD: need = px, src = L.d_spread; break;
    case SBM_RESPONSE: need = 8 * px, src = L.d_resp; break;
    case SBM_LINEAR: need = 8 * px, src = L.d_lm; break;
    case SBM_PYRAMID_SRC: need = px * c->c, src = level == 0 ? c->last_src0 : L.d_src; break;
    default: return fail(SBM_ERR_BAD_ARG, "unknown intermediate kind %d", kind);
    }
    if (bytes_needed)
        *bytes_needed = need;
    if (!dst)
        return 0;
    if (!src)
        return fail(SBM_ERR_STATE, "intermediate not materialised (call sbm_set_debug(ctx,1) before the frame)");
    if (bytes != need)
        return fail(SBM_ERR_BAD_ARG, "buffer holds %zu bytes, %zu needed", bytes, need);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(dst, src, need, cudaMemcpyDeviceToHost));
    return 0;
}

int sbm_similarity_map(sbm_ctx *c, int class_idx, int template_id, uint16_t *dst, size_t bytes)
{
    if (!c || !dst)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lock(c->mu);
    if (!c->have_frame)
        return fail(SBM_ERR_STATE, "no frame has been processed");
    if (class_idx < 0 || (size_t)class_idx >= c->classes.size() || template_id < 0 || template_id >= c->classes[class_idx].n)
        return fail(SBM_ERR_BAD_ARG, "unknown class/template"); // CV_Assert in getTemplates (line2Dup.cpp:1454-1456)
    CK(cudaSetDevice(c->device));
    const LevelBuf &CL = c->lv[c->levels - 1];
    const int HW = CL.W * CL.H;
    if (bytes != (size_t)HW * 2)
        return fail(SBM_ERR_BAD_ARG, "buffer holds %zu bytes, %d needed", bytes, HW * 2);
    uint32_t gid = c->gid_begin[class_idx] + template_id;
    uint32_t *d_one = nullptr, *d_cnt = nullptr;
    uint16_t *d_dump = nullptr;
    Cand *d_sink = nullptr;
    CK(cudaMalloc(&d_one, 4));
    CK(cudaMalloc(&d_cnt, 4));
    CK(cudaMalloc(&d_sink, sizeof(Cand)));
    CK(cudaMalloc(&d_dump, (size_t)HW * 2));
    CK(cudaMemcpy(d_one, &gid, 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(d_cnt, 0, 4));
    dim3 g(1, (HW + K5_TILE - 1) / K5_TILE);
    k_similarity<<<g, K5_THREADS, 0, c->own_stream>>>(CL.d_lm, c->d_coff, c->d_tcoarse, d_one, CL.W, HW, CL.T, 1e30f, d_sink, d_cnt, 0,
                                                      d_dump);
    CK(cudaStreamSynchronize(c->own_stream));
    CK(cudaMemcpy(dst, d_dump, (size_t)HW * 2, cudaMemcpyDeviceToHost));
    cudaFree(d_one), cudaFree(d_cnt), cudaFree(d_sink), cudaFree(d_dump);
    return 0;
}

int sbm_similarity_local(sbm_ctx *c, int class_idx, int template_id, int level, int x, int y, uint16_t *dst)
{
    if (!c || !dst || level < 0 || level >= c->levels)
        return fail(SBM_ERR_BAD_ARG, "bad argument");
    std::lock_guard<std::mutex> lock(c->mu);
    if (!c->have_frame)
        return fail(SBM_ERR_STATE, "no frame has been processed");
    if (class_idx < 0 || (size_t)class_idx >= c->classes.size() || template_id < 0 || template_id >= c->classes[class_idx].n)
        return fail(SBM_ERR_BAD_ARG, "unknown class/template");
    CK(cudaSetDevice(c->device));
    const LevelBuf &LB = c->lv[level];
    Cand one;
    one.x = x, one.y = y, one.sim = 0, one.tgid = c->gid_begin[class_idx] + template_id, one.key = 0;
    Cand *d_in = nullptr, *d_out = nullptr;
    uint32_t *d_cnt = nullptr;
    unsigned long long *d_ev = nullptr;
    uint16_t *d_dump = nullptr;
    CK(cudaMalloc(&d_in, sizeof(Cand)));
    CK(cudaMalloc(&d_out, sizeof(Cand)));
    CK(cudaMalloc(&d_cnt, 8));
    CK(cudaMalloc(&d_ev, 8));
    CK(cudaMalloc(&d_dump, 512));
    uint32_t cnt[2] = {1, 0};
    CK(cudaMemcpy(d_in, &one, sizeof(Cand), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_cnt, cnt, 8, cudaMemcpyHostToDevice));
    CK(cudaMemset(d_ev, 0, 8));
    k_refine<<<1, K6_THREADS, 0, c->own_stream>>>(LB.d_lm, LB.d_tl, LB.d_feats, d_in, d_cnt, 1, d_out, d_cnt + 1, 1, LB.w, LB.h, LB.T,
                                                  1e30f, d_dump, 1, d_ev);
    CK(cudaStreamSynchronize(c->own_stream));
    CK(cudaMemcpy(dst, d_dump, 512, cudaMemcpyDeviceToHost));
    cudaFree(d_in), cudaFree(d_out), cudaFree(d_cnt), cudaFree(d_ev), cudaFree(d_dump);
    return 0;
}

int sbm_last_stats(sbm_ctx *c, uint64_t stats[8])
{
    if (!c || !stats)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    memcpy(stats, c->stats, sizeof(c->stats));
    stats[0] = c->coarse_evals;
    stats[4] = (uint64_t)c->launches;
    return 0;
}

int sbm_last_timing(sbm_ctx *c, float ms[4])
{
    if (!c || !ms)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    if (c->timing && c->have_frame)
    {
        CK(cudaSetDevice(c->device));
        CK(cudaEventSynchronize(c->ev[2]));
        cudaEventElapsedTime(&c->last_ms[0], c->ev[0], c->ev[2]);
        cudaEventElapsedTime(&c->last_ms[1], c->ev[0], c->ev[1]);
        cudaEventElapsedTime(&c->last_ms[2], c->ev[1], c->ev[3]);
        cudaEventElapsedTime(&c->last_ms[3], c->ev[3], c->ev[2]);
    }
    memcpy(ms, c->last_ms, sizeof(c->last_ms));
    return 0;
}
}
