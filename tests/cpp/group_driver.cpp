// group_driver.cpp -- a C++ caller of the drop-in API using several GPUs (Detector::setGroup, include/line2Dup.h): no
// Python, no torch, no MPI.  One process per GPU; rank 0 writes the 128-byte NCCL id to <idfile>, the other ranks wait
// for it.  Driven by tests/test_cpp_group.py, which starts the ranks and compares the printed lists with the reference's.
//
//   group_driver <rank> <n_ranks> <mode 0|1|2> <idfile> <fmt> <class_id> <img.bin> <w> <h> <c> <thr> <num_features> <T0> [T1 ...]
//
// mode 0 / 1 (template sharding): every rank reads the full class and matches the same frame; every rank prints the
// merged list.  mode 2 (frame sharding): rank r matches the frame (even r) or its mirror image (odd r) through
// matchBatch; every rank prints the lists of all ranks.
#include "line2Dup.h"

#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <thread>

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
    if (argc < 14)
    {
        fprintf(stderr, "usage: see the header of tests/cpp/group_driver.cpp\n");
        return 2;
    }
    try
    {
        const int rank = atoi(argv[1]), n_ranks = atoi(argv[2]), mode = atoi(argv[3]);
        const std::string idfile = argv[4];
        std::vector<unsigned char> id;
        if (rank == 0)
        {
            id = line2Dup::Detector::groupUniqueId();
            std::ofstream((idfile + ".tmp").c_str(), std::ios::binary).write((const char *)id.data(), (std::streamsize)id.size());
            rename((idfile + ".tmp").c_str(), idfile.c_str());
        }
        else
        {
            for (int tries = 0; tries < 6000 && id.size() != 128; ++tries)
            {
                std::ifstream f(idfile.c_str(), std::ios::binary);
                std::vector<unsigned char> buf(128);
                if (f.read((char *)buf.data(), 128))
                    id = buf;
                else
                    std::this_thread::sleep_for(std::chrono::milliseconds(10));
            }
            if (id.size() != 128)
            {
                fprintf(stderr, "rank %d: no id file\n", rank);
                return 2;
            }
        }
        std::vector<int> T;
        for (int i = 13; i < argc; ++i)
            T.push_back(atoi(argv[i]));
        line2Dup::Detector det(atoi(argv[12]), T);
        det.setDevice(rank);
        det.setGroup(n_ranks, rank, id, (line2Dup::Detector::GroupMode)mode);
        det.readClasses({argv[6]}, argv[5]);
        cv::Mat img = loadRaw(argv[7], atoi(argv[8]), atoi(argv[9]), atoi(argv[10]));
        const float thr = (float)atof(argv[11]);
        printf("rank %d of %d templates %d\n", rank, det.groupSize(), det.numTemplates(argv[6]));
        if (mode != 2)
        {
            printMatches(det.match(img, thr, {argv[6]}));
            printMatches(det.match(img, thr)); // all classes
            for (const auto &ms : det.matchBatch({img, img}, thr, {argv[6]}))
                printMatches(ms);
        }
        else
        {
            cv::Mat mirrored(img.rows, img.cols, img.type());
            const int cn = img.channels();
            for (int y = 0; y < img.rows; ++y)
                for (int x = 0; x < img.cols; ++x)
                    memcpy(mirrored.ptr(y) + (size_t)x * cn, img.ptr(y) + (size_t)(img.cols - 1 - x) * cn, cn);
            for (const auto &ms : det.matchBatch({rank % 2 ? mirrored : img}, thr, {argv[6]}))
                printMatches(ms);
        }
        return 0;
    }
    catch (const std::exception &e)
    {
        printf("EXCEPTION %s\n", e.what());
        return 3;
    }
}
