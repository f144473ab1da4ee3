// ORACLE / TEST INFRASTRUCTURE -- plain scalar C++ restatement of the Detector::match hot path.
//
// This file is the checker, not the product: only tests/, __graft_entry__.smoke() and bench.py's CPU legs may load
// it (oracle/port.py).  It restates, loop by loop, what /root/reference/line2Dup.cpp computes on the match path,
// with the OpenCV calls replaced by their published integer / float definitions.  It is pinned by
// tests/test_oracle_port.py against tests/golden/*.npz, which are outputs of the reference's own code
// (oracle/_ref = line2Dup.cpp compiled verbatim) -- quantised images, linear memories, raw similarity maps,
// candidate lists and final lists must all be identical.
//
// Build: make -C oracle port   (g++ -O3, no dependencies)
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

namespace
{
inline int clampi(int v, int lo, int hi) { return v < lo ? lo : v > hi ? hi : v; }
inline int reflect101(int p, int n)
{
    if (n == 1)
        return 0;
    while (p < 0 || p >= n)
        p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

struct Image
{
    int w = 0, h = 0, c = 0;
    std::vector<uint8_t> px;
    uint8_t at(int y, int x, int ch) const { return px[((size_t)y * w + x) * c + ch]; }
};

// cv::GaussianBlur(src, dst, Size(7,7), 0, 0, BORDER_REPLICATE)  (line2Dup.cpp:320): 8-bit fixed-point kernel
// {8,28,56,72,56,28,8}/256 applied separably with ONE rounding: (sum + 2^15) >> 16.
Image gaussian7(const Image &s)
{
    static const int k[7] = {8, 28, 56, 72, 56, 28, 8};
    Image d = s;
    std::vector<int> hrow((size_t)s.w * s.c * s.h);
    for (int y = 0; y < s.h; ++y)
        for (int x = 0; x < s.w; ++x)
            for (int ch = 0; ch < s.c; ++ch)
            {
                int a = 0;
                for (int i = 0; i < 7; ++i)
                    a += k[i] * s.at(y, clampi(x + i - 3, 0, s.w - 1), ch);
                hrow[((size_t)y * s.w + x) * s.c + ch] = a;
            }
    for (int y = 0; y < s.h; ++y)
        for (int x = 0; x < s.w; ++x)
            for (int ch = 0; ch < s.c; ++ch)
            {
                int a = 0;
                for (int j = 0; j < 7; ++j)
                    a += k[j] * hrow[((size_t)clampi(y + j - 3, 0, s.h - 1) * s.w + x) * s.c + ch];
                d.px[((size_t)y * s.w + x) * s.c + ch] = (uint8_t)((a + 32768) >> 16);
            }
    return d;
}

// cv::phase(dx, dy, angle, true) = fastAtan32f, FMA form, degrees (line2Dup.cpp:327,398)
float fastAtan2Deg(float y, float x)
{
    const float p1 = 0.9997878412794807f * (float)(180 / M_PI), p3 = -0.3258083974640975f * (float)(180 / M_PI),
                p5 = 0.1555786518463281f * (float)(180 / M_PI), p7 = -0.04432655554792128f * (float)(180 / M_PI);
    float ax = std::fabs(x), ay = std::fabs(y), a, c, c2;
    if (ax >= ay)
    {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = std::fmaf(std::fmaf(std::fmaf(c2, p7, p5), c2, p3), c2, p1) * c;
    }
    else
    {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - std::fmaf(std::fmaf(std::fmaf(c2, p7, p5), c2, p3), c2, p1) * c;
    }
    if (x < 0)
        a = 180.f - a;
    if (y < 0)
        a = 360.f - a;
    return a;
}

// quantizedOrientations + hysteresisGradient (line2Dup.cpp:218-404): one-hot 8-bin label image
std::vector<uint8_t> quantizedOrientations(const Image &src, float weak)
{
    const Image b = gaussian7(src);
    const int w = src.w, h = src.h;
    std::vector<float> mag((size_t)w * h);
    std::vector<uint8_t> q((size_t)w * h);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x)
        {
            int bdx = 0, bdy = 0, bmag = -1;
            const int ym = clampi(y - 1, 0, h - 1), yp = clampi(y + 1, 0, h - 1), xm = clampi(x - 1, 0, w - 1), xp = clampi(x + 1, 0, w - 1);
            for (int ch = 0; ch < src.c; ++ch) // Sobel 3x3, BORDER_REPLICATE on the blurred image (:324-325, :343-344)
            {
                int dx = (b.at(ym, xp, ch) + 2 * b.at(y, xp, ch) + b.at(yp, xp, ch)) - (b.at(ym, xm, ch) + 2 * b.at(y, xm, ch) + b.at(yp, xm, ch));
                int dy = (b.at(yp, xm, ch) + 2 * b.at(yp, x, ch) + b.at(yp, xp, ch)) - (b.at(ym, xm, ch) + 2 * b.at(ym, x, ch) + b.at(ym, xp, ch));
                int m = dx * dx + dy * dy;
                if (m > bmag) // strongest channel, the first one on ties (:370-387)
                    bmag = m, bdx = dx, bdy = dy;
            }
            mag[(size_t)y * w + x] = (float)bmag;
            const float ang = fastAtan2Deg((float)bdy, (float)bdx);
            q[(size_t)y * w + x] = (uint8_t)std::nearbyintf(ang * (float)(16.0 / 360.0)); // convertTo(CV_8U, 16/360) (:225)
        }
    // zero the 1-px ring, fold 16 -> 8 directions (:229-257)
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x)
        {
            uint8_t &v = q[(size_t)y * w + x];
            v = (y == 0 || y == h - 1 || x == 0 || x == w - 1) ? 0 : (v & 7);
        }
    // 3x3 majority vote where the centre is strong enough (:262-310)
    std::vector<uint8_t> out((size_t)w * h, 0);
    const float thr = weak * weak;
    for (int y = 1; y < h - 1; ++y)
        for (int x = 1; x < w - 1; ++x)
        {
            if (!(mag[(size_t)y * w + x] > thr))
                continue;
            int hist[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int j = -1; j <= 1; ++j)
                for (int i = -1; i <= 1; ++i)
                    ++hist[q[(size_t)(y + j) * w + x + i]];
            int best = 0, idx = 0;
            for (int k = 0; k < 8; ++k)
                if (hist[k] > best)
                    best = hist[k], idx = k;
            if (best >= 5)
                out[(size_t)y * w + x] = (uint8_t)(1 << idx);
        }
    return out;
}

// cv::pyrDown(src, dst, Size(w/2, h/2)) (:431-433): {1,4,6,4,1}^2, BORDER_REFLECT_101, (sum + 128) >> 8
Image pyrDown(const Image &s)
{
    static const int k[5] = {1, 4, 6, 4, 1};
    Image d;
    d.w = s.w / 2, d.h = s.h / 2, d.c = s.c;
    d.px.resize((size_t)d.w * d.h * d.c);
    for (int y = 0; y < d.h; ++y)
        for (int x = 0; x < d.w; ++x)
            for (int ch = 0; ch < s.c; ++ch)
            {
                int a = 0;
                for (int j = 0; j < 5; ++j)
                    for (int i = 0; i < 5; ++i)
                        a += k[j] * k[i] * s.at(reflect101(2 * y + j - 2, s.h), reflect101(2 * x + i - 2, s.w), ch);
                d.px[((size_t)y * d.w + x) * d.c + ch] = (uint8_t)((a + 128) >> 8);
            }
    return d;
}

// spread (:616-630): OR over the forward TxT window, clipped at the right/bottom border
std::vector<uint8_t> spreadT(const std::vector<uint8_t> &q, int w, int h, int T)
{
    std::vector<uint8_t> s((size_t)w * h, 0);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x)
        {
            uint8_t v = 0;
            for (int j = 0; j < T && y + j < h; ++j)
                for (int i = 0; i < T && x + i < w; ++i)
                    v |= q[(size_t)(y + j) * w + x + i];
            s[(size_t)y * w + x] = v;
        }
    return s;
}

// computeResponseMaps (:632-747): max over the set bits of (4 - circular distance to ori), with the LUT's values
// {4 for distance 0, 3 (LUT3) for distance 1, 0 beyond}
uint8_t response(uint8_t spread, int ori)
{
    if (spread & (1 << ori))
        return 4;
    if (spread & ((1 << ((ori + 1) & 7)) | (1 << ((ori + 7) & 7))))
        return 3;
    return 0;
}

struct Feat
{
    int x, y, label;
};
struct Templ
{
    int width, height, tl_x, tl_y;
    std::vector<Feat> f;
};
struct Level
{
    int T = 0, w = 0, h = 0, W = 0, H = 0;
    std::vector<uint8_t> quant, spread, resp; // resp: [8][h][w]
    std::vector<uint8_t> lm;                  // [8][T*T][W*H] + zero tail (flat addressing may run past a memory)
};
struct Rec
{
    int x, y;
    float sim;
    int class_idx, template_id;
};

struct Oracle
{
    std::vector<int> T;
    float weak = 30.f;
    std::vector<std::vector<std::vector<Templ>>> classes; // [class][template][level]
    std::vector<Level> lv;                                // most recent frame
};

// quantize (:446-450) + spread + computeResponseMaps + linearize (:749-777) for every pyramid level (:1095-1120)
void buildPyramid(Oracle &o, Image img, const uint8_t *mask)
{
    const int L = (int)o.T.size();
    o.lv.assign(L, Level());
    std::vector<uint8_t> m;
    int mw = img.w;
    if (mask)
        m.assign(mask, mask + (size_t)img.w * img.h);
    for (int l = 0; l < L; ++l)
    {
        if (l > 0)
        {
            img = pyrDown(img);
            if (mask) // resize(mask, INTER_NEAREST) to half (:439)
            {
                std::vector<uint8_t> nm((size_t)img.w * img.h);
                for (int y = 0; y < img.h; ++y)
                    for (int x = 0; x < img.w; ++x)
                        nm[(size_t)y * img.w + x] = m[(size_t)(2 * y) * mw + 2 * x];
                m.swap(nm);
                mw = img.w;
            }
        }
        Level &V = o.lv[l];
        V.T = o.T[l], V.w = img.w, V.h = img.h, V.W = img.w / V.T, V.H = img.h / V.T;
        V.quant = quantizedOrientations(img, o.weak);
        if (mask)
            for (size_t i = 0; i < V.quant.size(); ++i)
                if (!m[i])
                    V.quant[i] = 0;
        V.spread = spreadT(V.quant, V.w, V.h, V.T);
        const size_t px = (size_t)V.w * V.h, WH = (size_t)V.W * V.H;
        V.resp.resize(8 * px);
        V.lm.assign(8 * px + WH + 16 * V.W + 64, 0);
        for (int ori = 0; ori < 8; ++ori)
            for (int y = 0; y < V.h; ++y)
                for (int x = 0; x < V.w; ++x)
                {
                    const uint8_t r = response(V.spread[(size_t)y * V.w + x], ori);
                    V.resp[ori * px + (size_t)y * V.w + x] = r;
                    // linear memory: row (y%T)*T + x%T, column (y/T)*W + x/T  (:759-775, accessLinearMemory :782-805)
                    V.lm[(size_t)ori * px + (size_t)((y % V.T) * V.T + x % V.T) * WH + (size_t)(y / V.T) * V.W + x / V.T] = r;
                }
    }
}

// similarity / similarity_64 (:807-858, :924-984): flat addressing over template_positions
void similarityMap(const Level &V, const Templ &t, std::vector<uint16_t> &dst)
{
    const int T = V.T, W = V.W, H = V.H;
    const size_t WH = (size_t)W * H, px = (size_t)V.w * V.h;
    dst.assign(WH, 0);
    const int wf = (t.width - 1) / T + 1, hf = (t.height - 1) / T + 1;
    const int span_x = W - wf, span_y = H - hf;
    const int tp = span_y * W + span_x + 1;
    for (const Feat &f : t.f)
    {
        if (f.x < 0 || f.x >= V.w || f.y < 0 || f.y >= V.h)
            continue;
        const uint8_t *p = &V.lm[(size_t)f.label * px + (size_t)((f.y % T) * T + f.x % T) * WH + (size_t)(f.y / T) * W + f.x / T];
        for (int j = 0; j < tp; ++j)
            dst[j] = (uint16_t)(dst[j] + p[j]);
    }
}

// similarityLocal / similarityLocal_64 (:860-922, :986-1048): 16x16 patch, rows of 16 consecutive bytes
void similarityLocal(const Level &V, const Templ &t, int cx, int cy, uint16_t dst[256])
{
    const int T = V.T, W = V.W, H = V.H;
    const size_t WH = (size_t)W * H, px = (size_t)V.w * V.h;
    memset(dst, 0, 256 * sizeof(uint16_t));
    const int ox = (cx / T - 8) * T, oy = (cy / T - 8) * T;
    for (const Feat &f0 : t.f)
    {
        const int x = f0.x + ox, y = f0.y + oy;
        if (x < 0 || y < 0 || x >= V.w || y >= V.h)
            continue;
        const uint8_t *p = &V.lm[(size_t)f0.label * px + (size_t)((y % T) * T + x % T) * WH + (size_t)(y / T) * W + x / T];
        for (int r = 0; r < 16; ++r)
            for (int c = 0; c < 16; ++c)
                dst[r * 16 + c] = (uint16_t)(dst[r * 16 + c] + p[(size_t)r * W + c]);
    }
}

// matchClass (:1160-1297) for one template: coarse scan, then refinement up the pyramid
void matchTemplate(const Oracle &o, const std::vector<Templ> &tp, float threshold, int class_idx, int template_id, std::vector<Rec> &out)
{
    const int L = (int)o.T.size();
    const Level &C = o.lv[L - 1];
    const Templ &tc = tp[L - 1];
    std::vector<uint16_t> sim;
    similarityMap(C, tc, sim);
    std::vector<Rec> cand;
    const int nf = (int)tc.f.size();
    const int off = C.T / 2 + (C.T % 2 - 1);
    for (int r = 0; r < C.H; ++r)
        for (int c = 0; c < C.W; ++c)
        {
            const int raw = sim[(size_t)r * C.W + c];
            const float score = (raw * 100.f) / (4 * nf); // :1206
            if (score > threshold)
                cand.push_back(Rec{c * C.T + off, r * C.T + off, score, class_idx, template_id});
        }
    for (int l = L - 2; l >= 0; --l) // :1221-1293
    {
        const Level &V = o.lv[l];
        const Templ &t = tp[l];
        const int T = V.T, border = 8 * T, offl = T / 2 + (T % 2 - 1);
        const int max_x = V.w - t.width - border, max_y = V.h - t.height - border;
        std::vector<Rec> kept;
        for (Rec m : cand)
        {
            int x = std::max(m.x * 2 + 1, border), y = std::max(m.y * 2 + 1, border);
            x = std::min(x, max_x), y = std::min(y, max_y);
            uint16_t patch[256];
            similarityLocal(V, t, x, y, patch);
            int best = 0, br = -1, bc = -1;
            for (int r = 0; r < 16; ++r)
                for (int c = 0; c < 16; ++c)
                    if (patch[r * 16 + c] > best)
                        best = patch[r * 16 + c], br = r, bc = c;
            m.x = (x / T - 8 + bc) * T + offl;
            m.y = (y / T - 8 + br) * T + offl;
            m.sim = (best * 100.f) / (4 * (int)t.f.size());
            if (!(m.sim < threshold)) // MatchPredicate (:1153-1158)
                kept.push_back(m);
        }
        cand.swap(kept);
    }
    out.insert(out.end(), cand.begin(), cand.end());
}
} // namespace

extern "C"
{
void *orc_create(const int *T, int levels, float weak)
{
    Oracle *o = new Oracle();
    o->T.assign(T, T + levels);
    o->weak = weak;
    return o;
}
void orc_destroy(void *h) { delete (Oracle *)h; }

// tdesc: [n][levels][5] = width, height, tl_x, tl_y, num_features; feats: [sum][3] = x, y, label
int orc_add_class(void *h, int n, const int *tdesc, const int *feats)
{
    Oracle *o = (Oracle *)h;
    const int L = (int)o->T.size();
    std::vector<std::vector<Templ>> cls(n, std::vector<Templ>(L));
    size_t fo = 0;
    for (int t = 0; t < n; ++t)
        for (int l = 0; l < L; ++l)
        {
            const int *d = tdesc + ((size_t)t * L + l) * 5;
            Templ &T = cls[t][l];
            T.width = d[0], T.height = d[1], T.tl_x = d[2], T.tl_y = d[3];
            for (int i = 0; i < d[4]; ++i, ++fo)
                T.f.push_back(Feat{feats[fo * 3], feats[fo * 3 + 1], feats[fo * 3 + 2]});
        }
    o->classes.push_back(cls);
    return (int)o->classes.size() - 1;
}

int orc_process(void *h, const uint8_t *img, int w, int hh, int c, const uint8_t *mask)
{
    Oracle *o = (Oracle *)h;
    Image im;
    im.w = w, im.h = hh, im.c = c;
    im.px.assign(img, img + (size_t)w * hh * c);
    buildPyramid(*o, im, mask);
    return 0;
}

// kind: 0 quantized, 1 spread, 2 response [8][h][w], 3 linear memories [8][T*T][W*H]
int orc_get(void *h, int level, int kind, uint8_t *dst)
{
    const Level &V = ((Oracle *)h)->lv[level];
    const size_t px = (size_t)V.w * V.h;
    const std::vector<uint8_t> &s = kind == 0 ? V.quant : kind == 1 ? V.spread : kind == 2 ? V.resp : V.lm;
    memcpy(dst, s.data(), kind <= 1 ? px : 8 * px);
    return 0;
}

int orc_similarity(void *h, int class_idx, int template_id, uint16_t *dst)
{
    Oracle *o = (Oracle *)h;
    const int L = (int)o->T.size();
    std::vector<uint16_t> sim;
    similarityMap(o->lv[L - 1], o->classes[class_idx][template_id][L - 1], sim);
    memcpy(dst, sim.data(), sim.size() * 2);
    return 0;
}

int orc_similarity_local(void *h, int class_idx, int template_id, int level, int x, int y, uint16_t *dst)
{
    Oracle *o = (Oracle *)h;
    similarityLocal(o->lv[level], o->classes[class_idx][template_id][level], x, y, dst);
    return 0;
}

// Detector::match (:1078-1150) on the most recent frame.  finalize = 0: the candidate list in the order one
// OpenMP thread appends it; 1: after std::sort + std::unique with Match::operator< / == (line2Dup.h:231-243).
int orc_match(void *h, float threshold, const int *class_idx, int n_classes, int finalize, Rec *out, int cap)
{
    Oracle *o = (Oracle *)h;
    std::vector<Rec> recs;
    for (int k = 0; k < n_classes; ++k)
    {
        const int ci = class_idx[k];
        if (ci < 0 || ci >= (int)o->classes.size())
            continue;
        for (size_t t = 0; t < o->classes[ci].size(); ++t)
            matchTemplate(*o, o->classes[ci][t], threshold, ci, (int)t, recs);
    }
    if (finalize)
    {
        std::sort(recs.begin(), recs.end(), [](const Rec &a, const Rec &b) {
            if (a.sim != b.sim)
                return a.sim > b.sim;
            return a.template_id < b.template_id;
        });
        recs.erase(std::unique(recs.begin(), recs.end(),
                               [](const Rec &a, const Rec &b) { return a.x == b.x && a.y == b.y && a.sim == b.sim && a.class_idx == b.class_idx; }),
                   recs.end());
    }
    const int n = (int)recs.size();
    if (n <= cap && n > 0)
        memcpy(out, recs.data(), (size_t)n * sizeof(Rec));
    return n;
}
}
