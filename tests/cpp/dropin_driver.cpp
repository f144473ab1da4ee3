// dropin_driver.cpp -- exercises the drop-in C++ API (include/line2Dup.h) the way the reference's own drivers do
// (test.cpp:162-524: readClasses / match; addTemplate / addTemplate_rotate / writeClasses).  Driven by
// tests/test_cpp_dropin.py, which compares the printed results with the reference's.
//
//   dropin_driver yaml   <in_fmt> <class_id> <out_fmt>                 read a class, write it back, print a digest
//   dropin_driver match  <fmt> <class_id> <img.bin> <w> <h> <c> <thr> <num_features> <T0> [T1 ...]
//   dropin_driver batch  (same arguments as match; the frame is matched 5 times through matchBatch)
//   dropin_driver train  <img.bin> <w> <h> <c> <mask.bin|-> <num_features> <out_fmt> <n_rot> <step_deg> <T0> [T1 ...]
#include "line2Dup.h"

#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>

static cv::Mat loadRaw(const char *path, int w, int h, int c)
{
    cv::Mat m(h, w, CV_MAKETYPE(CV_8U, c));
    std::ifstream f(path, std::ios::binary);
    if (!f.read((char *)m.data, (std::streamsize)w * h * c))
    {
        fprintf(stderr, "cannot read %s\n", path);
        exit(2);
    }
    return m;
}

static void printTemplates(const line2Dup::Detector &det, const std::string &cid)
{
    const int n = det.numTemplates(cid);
    printf("class %s templates %d levels %d\n", cid.c_str(), n, det.pyramidLevels());
    for (int t = 0; t < n; ++t)
    {
        const std::vector<line2Dup::Template> &tp = det.getTemplates(cid, t);
        for (const line2Dup::Template &tl : tp)
        {
            printf("T %d L %d %d %d %d %d nf %zu :", t, tl.pyramid_level, tl.width, tl.height, tl.tl_x, tl.tl_y, tl.features.size());
            for (const line2Dup::Feature &f : tl.features)
                printf(" %d,%d,%d", f.x, f.y, f.label);
            printf("\n");
        }
    }
}

static void printMatches(const std::vector<line2Dup::Match> &ms)
{
    printf("matches %zu\n", ms.size());
    for (const line2Dup::Match &m : ms)
    {
        uint32_t bits;
        memcpy(&bits, &m.similarity, 4);
        printf("M %d %d %08x %s %d\n", m.x, m.y, bits, m.class_id.c_str(), m.template_id);
    }
}

int main(int argc, char **argv)
{
    try
    {
        const std::string mode = argc > 1 ? argv[1] : "";
        if (mode == "yaml" && argc == 5)
        {
            line2Dup::Detector det;
            det.readClasses({argv[3]}, argv[2]);
            printTemplates(det, argv[3]);
            det.writeClasses(argv[4]);
            line2Dup::Detector again;
            again.readClasses({argv[3]}, argv[4]);
            printf("reread %d\n", again.numTemplates(argv[3]));
            return 0;
        }
        if ((mode == "match" || mode == "batch") && argc >= 11)
        {
            std::vector<int> T;
            for (int i = 10; i < argc; ++i)
                T.push_back(atoi(argv[i]));
            line2Dup::Detector det(atoi(argv[9]), T);
            det.readClasses({argv[3]}, argv[2]);
            cv::Mat img = loadRaw(argv[4], atoi(argv[5]), atoi(argv[6]), atoi(argv[7]));
            const float thr = (float)atof(argv[8]);
            if (mode == "match")
            {
                printMatches(det.match(img, thr, {argv[3]}));
                printMatches(det.match(img, thr)); // all classes
                printMatches(det.match(img, thr, {"no_such_class"}));
            }
            else
            {
                std::vector<cv::Mat> frames(5, img);
                for (const auto &ms : det.matchBatch(frames, thr, {argv[3]}))
                    printMatches(ms);
            }
            return 0;
        }
        if (mode == "train" && argc >= 12)
        {
            std::vector<int> T;
            for (int i = 11; i < argc; ++i)
                T.push_back(atoi(argv[i]));
            const int w = atoi(argv[3]), h = atoi(argv[4]), c = atoi(argv[5]);
            cv::Mat img = loadRaw(argv[2], w, h, c);
            cv::Mat mask;
            if (strcmp(argv[6], "-"))
                mask = loadRaw(argv[6], w, h, 1);
            line2Dup::Detector det(atoi(argv[7]), T);
            int id = det.addTemplate(img, "trained", mask);
            printf("addTemplate %d\n", id);
            if (id >= 0)
            {
                const int n_rot = atoi(argv[9]);
                const float step = (float)atof(argv[10]);
                for (int k = 1; k <= n_rot; ++k)
                    printf("rotate %d\n", det.addTemplate_rotate("trained", 0, step * k, cv::Point2f(w / 2.0f, h / 2.0f)));
                printTemplates(det, "trained");
                det.writeClasses(argv[8]);
            }
            return 0;
        }
        fprintf(stderr, "usage: see the header of tests/cpp/dropin_driver.cpp\n");
        return 2;
    }
    catch (const std::exception &e)
    {
        printf("EXCEPTION %s\n", e.what());
        return 3;
    }
}
