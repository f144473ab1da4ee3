// cv_compat.hpp -- the slice of the OpenCV C++ surface that the line2Dup API exposes.
//
// The reference's public header takes and returns cv::Mat / cv::Point2f / cv::Ptr / cv::FileNode
// (/root/reference/line2Dup.h:3-5, :216, :272-312, :344-460).  OpenCV's C++ headers are not part of
// this image, so the drop-in header (include/line2Dup.h) is written against this file when
// <opencv2/core.hpp> cannot be found.  Everything here is host-side container / persistence code:
// no image processing of the match() hot path lives in this header -- that is CUDA (csrc/).
//
// Contents: type macros, Exception / CV_Assert / CV_Error, Size, Point_, Rect, Scalar, a ref-counted
// Mat (+Mat_<T>), Ptr/makePtr, format(), and a FileStorage/FileNode/FileNodeIterator trio that reads
// and writes the OpenCV "YAML:1.0" dialect used by the reference's template files
// (/root/reference/test/case1/test_templ.yaml:1-20), plain or gzip-compressed.
#ifndef SBM_CV_COMPAT_HPP
#define SBM_CV_COMPAT_HPP

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <fstream>
#include <initializer_list>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include <zlib.h>

typedef unsigned char uchar;
typedef unsigned short ushort;

#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_CN_SHIFT 3
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAT_CN(t) ((((t) >> CV_CN_SHIFT) & 511) + 1)
#define CV_MAKETYPE(depth, cn) (CV_MAT_DEPTH(depth) + (((cn)-1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_8UC4 CV_MAKETYPE(CV_8U, 4)
#define CV_16UC1 CV_MAKETYPE(CV_16U, 1)
#define CV_16SC1 CV_MAKETYPE(CV_16S, 1)
#define CV_16SC3 CV_MAKETYPE(CV_16S, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_PI 3.1415926535897932384626433832795
#define CV_DECL_ALIGNED(x) __attribute__((aligned(x)))

namespace cv
{
typedef std::string String;

namespace Error
{
enum Code { StsError = -2, StsBadArg = -5, StsAssert = -215, StsParseError = -212 };
}

class Exception : public std::exception
{
public:
    Exception(int _code, const std::string &_err, const char *_func, const char *_file, int _line)
        : code(_code), err(_err), line(_line)
    {
        std::ostringstream os;
        os << "sbm/cv_compat(" << _file << ":" << _line << ") error (" << code << ") in " << _func << ": " << err;
        msg = os.str();
    }
    const char *what() const noexcept override { return msg.c_str(); }
    int code;
    std::string err, msg;
    int line;
};

#define CV_Error(code, message) throw cv::Exception(code, message, __func__, __FILE__, __LINE__)
#define CV_Assert(expr)                                                                      \
    do                                                                                       \
    {                                                                                        \
        if (!(expr))                                                                         \
            throw cv::Exception(cv::Error::StsAssert, #expr, __func__, __FILE__, __LINE__); \
    } while (0)
#define CV_DbgAssert(expr) ((void)0)

inline std::string format(const char *fmt, ...)
{
    char buf[4096];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    return std::string(buf);
}

template <typename T> using Ptr = std::shared_ptr<T>;
template <typename T, typename... A> inline Ptr<T> makePtr(A &&... a)
{
    return std::make_shared<T>(std::forward<A>(a)...);
}

struct Size
{
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
    bool operator==(const Size &o) const { return width == o.width && height == o.height; }
    bool operator!=(const Size &o) const { return !(*this == o); }
    int area() const { return width * height; }
};

template <typename T> struct Point_
{
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    template <typename U> Point_(const Point_<U> &o) : x(T(o.x)), y(T(o.y)) {}
    Point_ operator+(const Point_ &o) const { return Point_(x + o.x, y + o.y); }
    Point_ operator-(const Point_ &o) const { return Point_(x - o.x, y - o.y); }
    Point_ &operator/=(double d)
    {
        x = T(x / d);
        y = T(y / d);
        return *this;
    }
    Point_ &operator*=(double d)
    {
        x = T(x * d);
        y = T(y * d);
        return *this;
    }
    bool operator==(const Point_ &o) const { return x == o.x && y == o.y; }
};
typedef Point_<int> Point;
typedef Point_<float> Point2f;

struct Rect
{
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
    int area() const { return width * height; }
    Point tl() const { return Point(x, y); }
    Point br() const { return Point(x + width, y + height); }
};
inline Rect operator&(const Rect &a, const Rect &b)
{
    int x1 = std::max(a.x, b.x), y1 = std::max(a.y, b.y);
    int x2 = std::min(a.x + a.width, b.x + b.width), y2 = std::min(a.y + a.height, b.y + b.height);
    if (x2 <= x1 || y2 <= y1)
        return Rect();
    return Rect(x1, y1, x2 - x1, y2 - y1);
}

struct Scalar
{
    double val[4];
    Scalar(double a = 0, double b = 0, double c = 0, double d = 0) { val[0] = a, val[1] = b, val[2] = c, val[3] = d; }
    static Scalar all(double v) { return Scalar(v, v, v, v); }
    double operator[](int i) const { return val[i]; }
};

template <typename T> static inline T saturate_cast(double v);
template <> inline uchar saturate_cast<uchar>(double v)
{
    long i = lrint(v);
    return (uchar)(i < 0 ? 0 : i > 255 ? 255 : i);
}
template <> inline ushort saturate_cast<ushort>(double v)
{
    long i = lrint(v);
    return (ushort)(i < 0 ? 0 : i > 65535 ? 65535 : i);
}
template <> inline short saturate_cast<short>(double v)
{
    long i = lrint(v);
    return (short)(i < -32768 ? -32768 : i > 32767 ? 32767 : i);
}
template <> inline int saturate_cast<int>(double v) { return (int)lrint(v); }
template <> inline float saturate_cast<float>(double v) { return (float)v; }
template <> inline double saturate_cast<double>(double v) { return v; }

// Dense 2-D matrix, interleaved channels, always continuous, reference counted like cv::Mat.
class Mat
{
public:
    int rows, cols, flags;
    size_t step; // bytes per row
    uchar *data;

    Mat() : rows(0), cols(0), flags(0), step(0), data(nullptr) {}
    Mat(int r, int c, int type) : Mat() { create(r, c, type); }
    Mat(Size s, int type) : Mat() { create(s.height, s.width, type); }
    Mat(int r, int c, int type, const Scalar &s) : Mat()
    {
        create(r, c, type);
        setTo(s);
    }
    Mat(Size sz, int type, const Scalar &s) : Mat()
    {
        create(sz.height, sz.width, type);
        setTo(s);
    }
    // Header over caller-owned memory (no copy, no ownership), as cv::Mat(rows, cols, type, void*).
    Mat(int r, int c, int type, void *ext, size_t ext_step = 0) : rows(r), cols(c), flags(type), data((uchar *)ext)
    {
        step = ext_step ? ext_step : (size_t)c * elemSizeOf(type);
    }

    static size_t depthSize(int depth)
    {
        static const size_t s[8] = {1, 1, 2, 2, 4, 4, 8, 2};
        return s[depth & 7];
    }
    static size_t elemSizeOf(int type) { return depthSize(CV_MAT_DEPTH(type)) * CV_MAT_CN(type); }

    int type() const { return flags; }
    int depth() const { return CV_MAT_DEPTH(flags); }
    int channels() const { return CV_MAT_CN(flags); }
    size_t elemSize() const { return elemSizeOf(flags); }
    size_t elemSize1() const { return depthSize(depth()); }
    size_t step1() const { return step / elemSize1(); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }
    Size size() const { return Size(cols, rows); }
    size_t total() const { return (size_t)rows * cols; }

    void create(int r, int c, int type)
    {
        if (data && r == rows && c == cols && type == flags)
            return;
        rows = r;
        cols = c;
        flags = type;
        step = (size_t)c * elemSizeOf(type);
        size_t bytes = step * (size_t)r;
        // slack after the payload: the reference's SIMD loops read one full register past the last
        // element they use (/root/reference/line2Dup.cpp:842-853).
        holder = std::shared_ptr<uchar>((uchar *)aligned_alloc(64, ((bytes + 256 + 63) / 64) * 64), free);
        data = holder.get();
        memset(data + bytes, 0, 256);
    }
    void create(Size s, int type) { create(s.height, s.width, type); }
    void release()
    {
        holder.reset();
        data = nullptr;
        rows = cols = 0;
    }

    static Mat zeros(int r, int c, int type)
    {
        Mat m(r, c, type);
        memset(m.data, 0, m.step * (size_t)r);
        return m;
    }
    static Mat zeros(Size s, int type) { return zeros(s.height, s.width, type); }

    uchar *ptr(int r = 0) { return data + step * (size_t)r; }
    const uchar *ptr(int r = 0) const { return data + step * (size_t)r; }
    template <typename T> T *ptr(int r = 0) { return (T *)(data + step * (size_t)r); }
    template <typename T> const T *ptr(int r = 0) const { return (const T *)(data + step * (size_t)r); }
    template <typename T> T &at(int r, int c) { return ((T *)(data + step * (size_t)r))[c]; }
    template <typename T> const T &at(int r, int c) const { return ((const T *)(data + step * (size_t)r))[c]; }

    Mat clone() const
    {
        Mat m;
        copyTo(m);
        return m;
    }
    void copyTo(Mat &dst) const
    {
        if (empty())
        {
            dst.release();
            return;
        }
        if (dst.data == data)
            return;
        dst.create(rows, cols, flags);
        for (int r = 0; r < rows; ++r)
            memcpy(dst.ptr(r), ptr(r), (size_t)cols * elemSize());
    }
    void copyTo(Mat &dst, const Mat &mask) const
    {
        if (mask.empty())
        {
            copyTo(dst);
            return;
        }
        CV_Assert(mask.depth() == CV_8U && mask.channels() == 1 && mask.size() == size());
        if (dst.empty() || dst.size() != size() || dst.type() != type())
            dst = zeros(rows, cols, flags);
        size_t es = elemSize();
        for (int r = 0; r < rows; ++r)
        {
            const uchar *m = mask.ptr(r);
            const uchar *s = ptr(r);
            uchar *d = dst.ptr(r);
            for (int c = 0; c < cols; ++c)
                if (m[c])
                    memcpy(d + c * es, s + c * es, es);
        }
    }
    Mat &setTo(const Scalar &s)
    {
        int cn = channels();
        for (int r = 0; r < rows; ++r)
            for (int c = 0; c < cols; ++c)
                for (int k = 0; k < cn; ++k)
                    storeElem(r, c * cn + k, s.val[k < 4 ? k : 3]);
        return *this;
    }
    // dst = saturate(src * alpha + beta); the multiply-add is done in float for 8U/16U/16S/32F like
    // OpenCV's cvt kernels, rounding half-to-even (lrintf).
    void convertTo(Mat &dst, int rtype, double alpha = 1.0, double beta = 0.0) const
    {
        int ddepth = CV_MAT_DEPTH(rtype);
        int cn = channels();
        Mat out(rows, cols, CV_MAKETYPE(ddepth, cn));
        const bool ident = (alpha == 1.0 && beta == 0.0);
        const float a = (float)alpha, b = (float)beta;
        if (depth() == CV_32F && ddepth == CV_8U && beta == 0.0)
        {
            // the conversion on the match() path (line2Dup.cpp:225): a plain loop the compiler vectorises;
            // same arithmetic as the generic path below (float multiply, round half to even, saturate)
            const float *s = ptr<float>(0);
            uchar *d = out.ptr(0);
            const size_t n = (size_t)rows * cols * cn;
            for (size_t i = 0; i < n; ++i)
            {
                const float v = ident ? s[i] : s[i] * a;
                const float cl = v < 0.f ? 0.f : (v > 255.f ? 255.f : v); // NaN falls through to lrintf like before
                d[i] = (uchar)lrintf(cl);
            }
            dst = out;
            return;
        }
        for (int r = 0; r < rows; ++r)
            for (int c = 0; c < cols * cn; ++c)
            {
                double v = loadElem(r, c);
                if (!ident)
                    v = (double)((float)v * a + b);
                out.storeElem(r, c, v);
            }
        dst = out;
    }
    Mat mul(const Mat &o) const
    {
        CV_Assert(o.size() == size() && o.type() == type());
        Mat out(rows, cols, flags);
        int cn = channels();
        for (int r = 0; r < rows; ++r)
            for (int c = 0; c < cols * cn; ++c)
            {
                if (depth() == CV_32F)
                    out.at<float>(r, c) = at<float>(r, c) * o.at<float>(r, c);
                else
                    out.storeElem(r, c, loadElem(r, c) * o.loadElem(r, c));
            }
        return out;
    }
    // Region of interest: copies (this Mat class is always continuous).
    Mat operator()(const Rect &roi) const
    {
        CV_Assert(roi.x >= 0 && roi.y >= 0 && roi.x + roi.width <= cols && roi.y + roi.height <= rows);
        Mat out(roi.height, roi.width, flags);
        size_t es = elemSize();
        for (int r = 0; r < roi.height; ++r)
            memcpy(out.ptr(r), ptr(r + roi.y) + (size_t)roi.x * es, (size_t)roi.width * es);
        return out;
    }

    double loadElem(int r, int c1) const
    {
        switch (depth())
        {
        case CV_8U: return at<uchar>(r, c1);
        case CV_8S: return at<signed char>(r, c1);
        case CV_16U: return at<ushort>(r, c1);
        case CV_16S: return at<short>(r, c1);
        case CV_32S: return at<int>(r, c1);
        case CV_32F: return at<float>(r, c1);
        default: return at<double>(r, c1);
        }
    }
    void storeElem(int r, int c1, double v)
    {
        switch (depth())
        {
        case CV_8U: at<uchar>(r, c1) = saturate_cast<uchar>(v); break;
        case CV_16U: at<ushort>(r, c1) = saturate_cast<ushort>(v); break;
        case CV_16S: at<short>(r, c1) = saturate_cast<short>(v); break;
        case CV_32S: at<int>(r, c1) = saturate_cast<int>(v); break;
        case CV_32F: at<float>(r, c1) = (float)v; break;
        default: at<double>(r, c1) = v; break;
        }
    }

private:
    std::shared_ptr<uchar> holder;
};

inline Mat operator+(const Mat &a, const Mat &b)
{
    CV_Assert(a.size() == b.size() && a.type() == b.type());
    Mat out(a.rows, a.cols, a.type());
    int cn = a.channels();
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols * cn; ++c)
        {
            if (a.depth() == CV_32F)
                out.at<float>(r, c) = a.at<float>(r, c) + b.at<float>(r, c);
            else
                out.storeElem(r, c, a.loadElem(r, c) + b.loadElem(r, c));
        }
    return out;
}
inline Mat operator>(const Mat &a, double v)
{
    Mat out(a.rows, a.cols, CV_MAKETYPE(CV_8U, a.channels()));
    int cn = a.channels();
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols * cn; ++c)
            out.at<uchar>(r, c) = a.loadElem(r, c) > v ? 255 : 0;
    return out;
}

template <typename T> class Mat_ : public Mat
{
public:
    Mat_() : Mat() {}
    Mat_(const Mat &m) : Mat(m) {}
    Mat_ &operator=(const Mat &m)
    {
        Mat::operator=(m);
        return *this;
    }
    T &operator()(int r, int c) { return this->template at<T>(r, c); }
    const T &operator()(int r, int c) const { return this->template at<T>(r, c); }
};

// ------------------------------------------------------------------------------------------------
// Persistence: OpenCV FileStorage "YAML:1.0" subset (maps, block/flow sequences, int/real/string).
// ------------------------------------------------------------------------------------------------
namespace detail
{
struct Node
{
    enum Kind { NONE, INT, REAL, STR, SEQ, MAP } kind = NONE;
    long long i = 0;
    double d = 0;
    std::string s;
    std::vector<std::shared_ptr<Node>> seq;
    std::vector<std::pair<std::string, std::shared_ptr<Node>>> map;
};
typedef std::shared_ptr<Node> NodePtr;

inline std::string trim(const std::string &s)
{
    size_t a = s.find_first_not_of(" \t\r\n");
    if (a == std::string::npos)
        return "";
    size_t b = s.find_last_not_of(" \t\r\n");
    return s.substr(a, b - a + 1);
}

inline NodePtr parseScalar(const std::string &tok0)
{
    std::string tok = trim(tok0);
    NodePtr n = std::make_shared<Node>();
    if (tok.empty())
        return n;
    if (tok[0] == '"')
    {
        n->kind = Node::STR;
        std::string out;
        for (size_t k = 1; k < tok.size() && tok[k] != '"'; ++k)
        {
            if (tok[k] == '\\' && k + 1 < tok.size())
            {
                ++k;
                out += tok[k] == 'n' ? '\n' : tok[k] == 't' ? '\t' : tok[k];
            }
            else
                out += tok[k];
        }
        n->s = out;
        return n;
    }
    char *end = nullptr;
    long long iv = strtoll(tok.c_str(), &end, 10);
    if (end && *end == 0)
    {
        n->kind = Node::INT;
        n->i = iv;
        n->d = (double)iv;
        return n;
    }
    std::string t2 = tok;
    if (t2 == ".Inf" || t2 == ".inf")
        t2 = "inf";
    else if (t2 == "-.Inf" || t2 == "-.inf")
        t2 = "-inf";
    else if (t2 == ".Nan" || t2 == ".NaN" || t2 == ".nan")
        t2 = "nan";
    double dv = strtod(t2.c_str(), &end);
    if (end && *end == 0 && end != t2.c_str())
    {
        n->kind = Node::REAL;
        n->d = dv;
        n->i = (long long)lrint(dv);
        return n;
    }
    n->kind = Node::STR;
    n->s = tok;
    return n;
}

// flow collection starting at text[pos] == '[' or '{'
inline NodePtr parseFlow(const std::string &text, size_t &pos)
{
    NodePtr n = std::make_shared<Node>();
    char open = text[pos];
    char close = open == '[' ? ']' : '}';
    n->kind = open == '[' ? Node::SEQ : Node::MAP;
    ++pos;
    if (pos < text.size() && text[pos] == ':')
        ++pos; // "[:" compact marker is only written, but be lenient
    std::string tok;
    std::string key;
    bool have_child = false;
    NodePtr child;
    auto flush = [&]() {
        if (!have_child)
        {
            if (trim(tok).empty())
            {
                tok.clear();
                return;
            }
            child = parseScalar(tok);
        }
        if (n->kind == Node::SEQ)
            n->seq.push_back(child);
        else
            n->map.push_back(std::make_pair(key, child));
        tok.clear();
        key.clear();
        have_child = false;
    };
    while (pos < text.size())
    {
        char ch = text[pos];
        if (ch == '"')
        {
            size_t e = pos + 1;
            while (e < text.size() && text[e] != '"')
                e += (text[e] == '\\') ? 2 : 1;
            tok += text.substr(pos, e - pos + 1);
            pos = e + 1;
            continue;
        }
        if (ch == '[' || ch == '{')
        {
            child = parseFlow(text, pos);
            have_child = true;
            continue;
        }
        if (ch == close)
        {
            flush();
            ++pos;
            return n;
        }
        if (ch == ',')
        {
            flush();
            ++pos;
            continue;
        }
        if (ch == ':' && n->kind == Node::MAP && key.empty())
        {
            key = trim(tok);
            tok.clear();
            ++pos;
            continue;
        }
        tok += ch;
        ++pos;
    }
    CV_Error(Error::StsParseError, "unterminated flow collection");
    return n;
}

struct Line
{
    int indent;
    std::string text; // without indentation
};

inline NodePtr parseValueText(const std::string &v, std::vector<Line> &lines, size_t &li);

// parse a block (map or sequence) whose entries sit at exactly `indent`
inline NodePtr parseBlock(std::vector<Line> &lines, size_t &li, int indent)
{
    NodePtr n = std::make_shared<Node>();
    if (li >= lines.size())
        return n;
    bool is_seq = lines[li].text[0] == '-' && (lines[li].text.size() == 1 || lines[li].text[1] == ' ');
    n->kind = is_seq ? Node::SEQ : Node::MAP;
    while (li < lines.size() && lines[li].indent == indent)
    {
        std::string &t = lines[li].text;
        if (is_seq)
        {
            if (!(t[0] == '-' && (t.size() == 1 || t[1] == ' ')))
                break;
            std::string rest = trim(t.substr(1));
            if (rest.empty())
            {
                ++li;
                if (li < lines.size() && lines[li].indent > indent)
                    n->seq.push_back(parseBlock(lines, li, lines[li].indent));
                else
                    n->seq.push_back(std::make_shared<Node>());
            }
            else if (rest[0] == '[' || rest[0] == '{')
            {
                // flow collection, possibly continued on following lines
                std::string acc = rest;
                ++li;
                auto depth = [](const std::string &s) {
                    int d = 0;
                    bool q = false;
                    for (size_t k = 0; k < s.size(); ++k)
                    {
                        if (s[k] == '"')
                            q = !q;
                        if (q)
                            continue;
                        if (s[k] == '[' || s[k] == '{')
                            ++d;
                        if (s[k] == ']' || s[k] == '}')
                            --d;
                    }
                    return d;
                };
                while (depth(acc) > 0 && li < lines.size())
                    acc += " " + lines[li++].text;
                size_t p = 0;
                n->seq.push_back(parseFlow(acc, p));
            }
            else
            {
                // "- key: value" (inline map start) or plain scalar
                size_t colon = rest.find(':');
                bool looks_map = colon != std::string::npos && rest[0] != '"' &&
                                 (colon + 1 == rest.size() || rest[colon + 1] == ' ');
                if (looks_map)
                {
                    // rewrite as a nested block at indent+2
                    int sub = indent + 2;
                    lines[li].indent = sub;
                    lines[li].text = rest;
                    n->seq.push_back(parseBlock(lines, li, sub));
                }
                else
                {
                    n->seq.push_back(parseScalar(rest));
                    ++li;
                }
            }
        }
        else
        {
            if (t[0] == '-' && (t.size() == 1 || t[1] == ' '))
                break;
            size_t colon = std::string::npos;
            {
                bool q = false;
                for (size_t k = 0; k < t.size(); ++k)
                {
                    if (t[k] == '"')
                        q = !q;
                    if (!q && t[k] == ':' && (k + 1 == t.size() || t[k + 1] == ' '))
                    {
                        colon = k;
                        break;
                    }
                }
            }
            if (colon == std::string::npos)
                CV_Error(Error::StsParseError, "expected 'key: value' but got: " + t);
            std::string key = trim(t.substr(0, colon));
            if (!key.empty() && key[0] == '"')
                key = parseScalar(key)->s;
            std::string val = trim(t.substr(colon + 1));
            ++li;
            n->map.push_back(std::make_pair(key, parseValueText(val, lines, li)));
        }
    }
    return n;
}

inline NodePtr parseValueText(const std::string &v0, std::vector<Line> &lines, size_t &li)
{
    std::string v = v0;
    // drop OpenCV type tags such as !!opencv-matrix
    if (v.size() > 1 && v[0] == '!' && v[1] == '!')
    {
        size_t sp = v.find(' ');
        v = sp == std::string::npos ? "" : trim(v.substr(sp));
    }
    if (v.empty())
    {
        // nested block: deeper indentation, or a sequence at the same indentation as the key
        if (li < lines.size())
        {
            int prev_indent = li > 0 ? lines[li - 1].indent : 0;
            const std::string &t = lines[li].text;
            bool seq_here = t[0] == '-' && (t.size() == 1 || t[1] == ' ');
            if (lines[li].indent > prev_indent || (lines[li].indent == prev_indent && seq_here))
                return parseBlock(lines, li, lines[li].indent);
        }
        return std::make_shared<Node>();
    }
    if (v[0] == '[' || v[0] == '{')
    {
        std::string acc = v;
        auto depth = [](const std::string &s) {
            int d = 0;
            bool q = false;
            for (size_t k = 0; k < s.size(); ++k)
            {
                if (s[k] == '"')
                    q = !q;
                if (q)
                    continue;
                if (s[k] == '[' || s[k] == '{')
                    ++d;
                if (s[k] == ']' || s[k] == '}')
                    --d;
            }
            return d;
        };
        while (depth(acc) > 0 && li < lines.size())
            acc += " " + lines[li++].text;
        size_t p = 0;
        return parseFlow(acc, p);
    }
    return parseScalar(v);
}

inline NodePtr parseYaml(const std::string &content)
{
    std::vector<Line> lines;
    std::istringstream is(content);
    std::string raw;
    while (std::getline(is, raw))
    {
        if (!raw.empty() && raw.back() == '\r')
            raw.pop_back();
        size_t a = raw.find_first_not_of(' ');
        if (a == std::string::npos)
            continue;
        std::string t = raw.substr(a);
        if (t[0] == '%' || t[0] == '#')
            continue;
        if (t.compare(0, 3, "---") == 0 || t.compare(0, 3, "...") == 0)
            continue;
        size_t b = t.find_last_not_of(" \t");
        lines.push_back(Line{(int)a, t.substr(0, b + 1)});
    }
    size_t li = 0;
    if (lines.empty())
    {
        NodePtr n = std::make_shared<Node>();
        n->kind = Node::MAP;
        return n;
    }
    return parseBlock(lines, li, lines[0].indent);
}

inline bool endsWith(const std::string &s, const char *suf)
{
    size_t n = strlen(suf);
    return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

inline bool readWholeFile(const std::string &path, std::string &out)
{
    // gzopen transparently reads plain files as well
    gzFile f = gzopen(path.c_str(), "rb");
    if (!f)
        return false;
    char buf[1 << 16];
    int n;
    out.clear();
    while ((n = gzread(f, buf, sizeof(buf))) > 0)
        out.append(buf, (size_t)n);
    gzclose(f);
    return true;
}
} // namespace detail

class FileNode;
class FileNodeIterator
{
public:
    FileNodeIterator() : idx(0) {}
    FileNodeIterator(detail::NodePtr c, size_t i) : container(c), idx(i) {}
    FileNode operator*() const;
    FileNodeIterator &operator++()
    {
        ++idx;
        return *this;
    }
    FileNodeIterator operator++(int)
    {
        FileNodeIterator t = *this;
        ++idx;
        return t;
    }
    bool operator!=(const FileNodeIterator &o) const { return idx != o.idx || container != o.container; }
    bool operator==(const FileNodeIterator &o) const { return !(*this != o); }
    template <typename T> FileNodeIterator &operator>>(T &v);

private:
    detail::NodePtr container;
    size_t idx;
};

class FileNode
{
public:
    FileNode() {}
    explicit FileNode(detail::NodePtr n) : node(n) {}
    bool empty() const { return !node || node->kind == detail::Node::NONE; }
    bool isNone() const { return empty(); }
    bool isSeq() const { return node && node->kind == detail::Node::SEQ; }
    bool isMap() const { return node && node->kind == detail::Node::MAP; }
    bool isInt() const { return node && node->kind == detail::Node::INT; }
    bool isReal() const { return node && node->kind == detail::Node::REAL; }
    bool isString() const { return node && node->kind == detail::Node::STR; }
    size_t size() const
    {
        if (!node)
            return 0;
        if (node->kind == detail::Node::SEQ)
            return node->seq.size();
        if (node->kind == detail::Node::MAP)
            return node->map.size();
        return node->kind == detail::Node::NONE ? 0 : 1;
    }
    FileNode operator[](const std::string &key) const
    {
        if (node && node->kind == detail::Node::MAP)
            for (auto &kv : node->map)
                if (kv.first == key)
                    return FileNode(kv.second);
        return FileNode();
    }
    FileNode operator[](const char *key) const { return (*this)[std::string(key)]; }
    FileNode operator[](int i) const
    {
        if (node && node->kind == detail::Node::SEQ && (size_t)i < node->seq.size())
            return FileNode(node->seq[i]);
        return FileNode();
    }
    // A missing node converts to 0 / "" exactly like cv::FileNode (the fixtures lack the fork's extra
    // keys "scale", "orientation", "tagFieldID", "fiducial_src": /root/reference/line2Dup.cpp:60-70).
    operator int() const { return node ? (node->kind == detail::Node::REAL ? (int)lrint(node->d) : (int)node->i) : 0; }
    operator float() const { return node ? (float)numeric() : 0.f; }
    operator double() const { return node ? numeric() : 0.0; }
    operator std::string() const { return node && node->kind == detail::Node::STR ? node->s : std::string(); }
    std::string name() const { return std::string(); }
    FileNodeIterator begin() const { return FileNodeIterator(containerNode(), 0); }
    FileNodeIterator end() const { return FileNodeIterator(containerNode(), size()); }
    detail::NodePtr raw() const { return node; }

private:
    double numeric() const { return node->kind == detail::Node::INT ? (double)node->i : node->d; }
    detail::NodePtr containerNode() const { return node; }
    detail::NodePtr node;
};

inline FileNode FileNodeIterator::operator*() const
{
    if (!container)
        return FileNode();
    if (container->kind == detail::Node::SEQ)
        return idx < container->seq.size() ? FileNode(container->seq[idx]) : FileNode();
    if (container->kind == detail::Node::MAP)
        return idx < container->map.size() ? FileNode(container->map[idx].second) : FileNode();
    return FileNode(container);
}

inline void operator>>(const FileNode &n, int &v) { v = (int)n; }
inline void operator>>(const FileNode &n, float &v) { v = (float)n; }
inline void operator>>(const FileNode &n, double &v) { v = (double)n; }
inline void operator>>(const FileNode &n, std::string &v) { v = (std::string)n; }
template <typename T> inline void operator>>(const FileNode &n, std::vector<T> &v)
{
    v.clear();
    for (FileNodeIterator it = n.begin(), e = n.end(); it != e; ++it)
    {
        T t;
        (*it) >> t;
        v.push_back(t);
    }
}
template <typename T> inline FileNodeIterator &FileNodeIterator::operator>>(T &v)
{
    (**this) >> v;
    ++idx;
    return *this;
}

class FileStorage
{
public:
    enum Mode { READ = 0, WRITE = 1, APPEND = 2, MEMORY = 4 };
    FileStorage() : mode_(READ), opened(false) {}
    FileStorage(const std::string &filename, int flags) : mode_(READ), opened(false) { open(filename, flags); }
    ~FileStorage() { release(); }

    bool open(const std::string &filename, int flags)
    {
        release();
        path = filename;
        mode_ = flags & 3;
        if (mode_ == READ)
        {
            std::string content;
            if (flags & MEMORY)
                content = filename;
            else if (!detail::readWholeFile(filename, content))
            {
                rootNode = std::make_shared<detail::Node>();
                return false;
            }
            rootNode = detail::parseYaml(content);
            opened = true;
        }
        else
        {
            out.str("");
            out << "%YAML:1.0\n---\n";
            stack.clear();
            stack.push_back(Frame{false, false, 0, 0});
            pendingKey.clear();
            haveKey = false;
            opened = true;
        }
        return opened;
    }
    bool isOpened() const { return opened; }
    void release()
    {
        if (opened && mode_ == WRITE)
        {
            std::string text = out.str();
            if (detail::endsWith(path, ".gz"))
            {
                gzFile f = gzopen(path.c_str(), "wb");
                if (f)
                {
                    gzwrite(f, text.data(), (unsigned)text.size());
                    gzclose(f);
                }
            }
            else
            {
                std::ofstream f(path.c_str(), std::ios::binary);
                f << text;
            }
        }
        opened = false;
    }
    std::string releaseAndGetString()
    {
        std::string t = out.str();
        opened = false;
        return t;
    }
    FileNode root() const { return FileNode(rootNode); }
    FileNode operator[](const std::string &key) const { return root()[key]; }
    FileNode operator[](const char *key) const { return root()[std::string(key)]; }

    // ---- writer state machine (mirrors cv::FileStorage's operator<< protocol) ----
    void writeToken(const std::string &s)
    {
        Frame &top = stack.back();
        bool in_map = !top.is_seq;
        if (in_map && !haveKey)
        {
            if (s == "}")
            {
                endCollection();
                return;
            }
            pendingKey = s;
            haveKey = true;
            return;
        }
        if (!in_map && s == "]")
        {
            endCollection();
            return;
        }
        if (s == "[" || s == "[:" || s == "{" || s == "{:")
        {
            beginCollection(s[0] == '[', s.size() > 1);
            return;
        }
        writeScalarText(quoteIfNeeded(s));
    }
    void writeInt(long long v)
    {
        char b[32];
        snprintf(b, sizeof(b), "%lld", v);
        writeScalarText(b);
    }
    void writeReal(double v, bool is_float)
    {
        writeScalarText(realToString(v, is_float));
    }

private:
    struct Frame
    {
        bool is_seq, flow;
        int indent;
        int count;
    };
    static std::string realToString(double v, bool is_float)
    {
        char b[64];
        if (std::isnan(v))
            return ".Nan";
        if (std::isinf(v))
            return v < 0 ? "-.Inf" : ".Inf";
        long long iv = (long long)v;
        if ((double)iv == v && std::fabs(v) < 1e15)
        {
            snprintf(b, sizeof(b), "%lld.", iv);
            return b;
        }
        snprintf(b, sizeof(b), is_float ? "%.8e" : "%.16e", v);
        return b;
    }
    static std::string quoteIfNeeded(const std::string &s)
    {
        bool need = s.empty();
        if (!need)
        {
            if (!(isalpha((unsigned char)s[0]) || s[0] == '_'))
                need = true;
            for (char ch : s)
                if (!(isalnum((unsigned char)ch) || ch == '_' || ch == '-' || ch == '.' || ch == '/'))
                    need = true;
        }
        if (!need)
            return s;
        std::string o = "\"";
        for (char ch : s)
        {
            if (ch == '"' || ch == '\\')
                o += '\\';
            o += ch;
        }
        return o + "\"";
    }
    void emitPrefix(bool opening_block)
    {
        Frame &top = stack.back();
        if (top.flow)
        {
            out << (top.count ? ", " : " ");
            if (!top.is_seq)
                out << pendingKey << ":";
        }
        else
        {
            out << std::string(top.indent, ' ');
            if (top.is_seq)
                out << "-";
            else
                out << pendingKey << ":";
            if (!opening_block)
                out << " ";
        }
        top.count++;
        haveKey = false;
    }
    void writeScalarText(const std::string &txt)
    {
        Frame &top = stack.back();
        if (!top.is_seq && !haveKey)
            CV_Error(Error::StsError, "FileStorage: value without a key inside a mapping");
        emitPrefix(false);
        if (top.flow)
        {
            if (!top.is_seq)
                out << " ";
            out << txt;
        }
        else
            out << txt << "\n";
    }
    void beginCollection(bool is_seq, bool flow)
    {
        Frame &top = stack.back();
        if (!top.is_seq && !haveKey)
            CV_Error(Error::StsError, "FileStorage: collection without a key inside a mapping");
        bool parent_flow = top.flow;
        int parent_indent = top.indent;
        bool parent_root = stack.size() == 1;
        emitPrefix(!(flow || parent_flow));
        Frame f;
        f.is_seq = is_seq;
        f.flow = flow || parent_flow;
        f.count = 0;
        f.indent = parent_root ? 3 : parent_indent + 3;
        if (f.flow)
            out << (is_seq ? "[" : "{");
        else
            out << "\n";
        stack.push_back(f);
    }
    void endCollection()
    {
        Frame f = stack.back();
        stack.pop_back();
        if (f.flow)
        {
            out << (f.is_seq ? " ]" : " }");
            if (!stack.back().flow)
                out << "\n";
        }
        haveKey = false;
    }

    int mode_;
    bool opened;
    std::string path;
    detail::NodePtr rootNode;
    std::ostringstream out;
    std::vector<Frame> stack;
    std::string pendingKey;
    bool haveKey;
};

inline FileStorage &operator<<(FileStorage &fs, const std::string &s)
{
    fs.writeToken(s);
    return fs;
}
inline FileStorage &operator<<(FileStorage &fs, const char *s)
{
    fs.writeToken(std::string(s));
    return fs;
}
inline FileStorage &operator<<(FileStorage &fs, int v)
{
    fs.writeInt(v);
    return fs;
}
inline FileStorage &operator<<(FileStorage &fs, float v)
{
    fs.writeReal(v, true);
    return fs;
}
inline FileStorage &operator<<(FileStorage &fs, double v)
{
    fs.writeReal(v, false);
    return fs;
}
template <typename T> inline FileStorage &operator<<(FileStorage &fs, const std::vector<T> &v)
{
    fs << "[:";
    for (const T &t : v)
        fs << t;
    fs << "]";
    return fs;
}

} // namespace cv

#endif // SBM_CV_COMPAT_HPP
