// sbm_capi.cu -- host side of the C ABI declared in include/sbm_capi.h.
//
// One sbm_ctx = one Detector on one GPU.  It owns
//   * the template set as device SoA (per pyramid level) and, per frame shape, the coarse-level feature
//     offsets into the linear memories;
//   * SBM_SLOTS frame slots.  A slot is everything one in-flight frame needs -- a stream, the pyramid sources,
//     quantised images, linear memories, candidate ping-pong buffers, pinned result staging -- so that
//     sbm_match_batch can keep several frames in flight: the H2D copy of frame i+1 and the kernels of frame i
//     overlap, and the small kernels of neighbouring frames share the GPU.  Slot 0 serves the single-frame calls.
// The only CPU work per frame is ordering the (tiny) candidate list the way the reference's serial loop
// produces it and running the same std::sort / std::unique (line2Dup.cpp:1143-1145).  There is no CPU
// implementation of any pipeline stage in this library.
#include "sbm_capi.h"
#include "sbm_kernels.cuh"
#include "sbm_nccl.hpp"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
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

constexpr int SBM_SLOTS = 8; // frame slots a context can hold
// frame slots the batch entry points keep in flight (SBM_FRAME_SLOTS=1..8 overrides; see profiles/ for the sweep)
// Eight frame graphs of two branches each want more than the default 8 hardware queues (false dependencies between unrelated
// kernels otherwise: M1 in flight 19.3 -> 18.8 us per frame).  Only takes effect if CUDA is not initialised yet in this
// process and the caller has not set the variable; a host that initialises CUDA first sets it itself.
const int g_max_connections_set = setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);

int frameSlots()
{
    static const int n = [] {
        const char *e = getenv("SBM_FRAME_SLOTS");
        if (!e)
            e = getenv("SBM_DEVICE_SLOTS"); // older name
        const int v = e ? atoi(e) : 8; // sweep on M1 (round 1): 2 -> 41.5, 4 -> 33.2, 6 -> 31.8, 8 -> 32.1 us per frame; now 6 -> 19.1, 8 -> 18.8
                                       // (with CUDA_DEVICE_MAX_CONNECTIONS = 32: a frame graph has two branches, 8 slots want 16 queues)
        return v < 1 ? 1 : (v > SBM_SLOTS ? SBM_SLOTS : v);
    }();
    return n;
}
constexpr int COUNT_WORDS = 16; // [0..levels-1] stage counts, [10..11] u64 local evaluations
constexpr uint32_t HEAD_RECS = 1024;

// K1 runs on the tensor cores (k_orient_mma: blur and Sobel as banded int8 matrix products) unless SBM_ORIENT_MMA=0
// asks for the integer-pipe kernel k_orient; both produce the same bytes (the parity suite runs both).
bool orientMma()
{
    static const bool on = !(getenv("SBM_ORIENT_MMA") && atoi(getenv("SBM_ORIENT_MMA")) == 0);
    return on;
}

// K1 tile height for a w x h level.  k_orient: 64x32 tiles unless that leaves the GPU under two CTAs per SM, then 64x16.
// k_orient_mma: 64 x (8 NT - 4) = 36, 20 or 12 rows, see below.
// SBM_ORIENT_TH=8|16|28|32 (k_orient) / 12|20|28|36 (k_orient_mma) forces one height (experiments).
// sms: multiprocessors of the context's device.
int orientTileH(int w, int h, int sms)
{
    static const int forced = getenv("SBM_ORIENT_TH") ? atoi(getenv("SBM_ORIENT_TH")) : 0;
    const int cols = (w + K1_TW - 1) / K1_TW;
    if (orientMma())
    {
        if (forced == 12 || forced == 20 || forced == 28 || forced == 36)
            return forced;
        // With several frames in flight a kernel costs its instruction count (tools/inflight_shares.py), and tall tiles
        // have the smaller halo: M1 in flight 20.6 us per frame with 28 / 12 rows (levels 0 / 1), 20.1 with 28 / 28,
        // 19.6 with 36 / 36, 20.2 with 44 / 44 (three CTAs per SM; that height is gone since a CTA holds two windows); alone, level 0 takes 9.6 us with 28 rows (592 CTAs =
        // 4 per SM) and 11.3 with 36.  Small levels keep shorter tiles so that half the SMs have a CTA.
        for (int th : {36, 20})
            if (2 * cols * ((h + th - 1) / th) >= sms)
                return th;
        return 12;
    }
    if (forced == 8 || forced == 16 || forced == 28 || forced == 32)
        return forced;
    if (cols * ((h + 31) / 32) >= 2 * sms)
        return 32;
    // 64x28 tiles fill the 148 SMs evenly at 1024 rows (592 CTAs = 4 per SM; the kernel alone 13.6 -> 12.5 us) but the
    // frame as a whole got no faster (graph 52.2 -> 52.8 us, 23.2 -> 23.5 us per frame in flight): not used
    return 16; // 64x8 tiles (512 CTAs at level 1 of a 1024^2 frame) measured no faster than 64x16: 8.5 vs 8.1 us
}

void *orientFunc(int ch, int th)
{
    if (orientMma())
    {
        if (ch == 3)
            return th == 36 ? (void *)k_orient_mma<3, 5> : th == 28 ? (void *)k_orient_mma<3, 4> : th == 20 ? (void *)k_orient_mma<3, 3> : (void *)k_orient_mma<3, 2>;
        return th == 36 ? (void *)k_orient_mma<1, 5> : th == 28 ? (void *)k_orient_mma<1, 4> : th == 20 ? (void *)k_orient_mma<1, 3> : (void *)k_orient_mma<1, 2>;
    }
    if (ch == 3)
        return th == 32 ? (void *)k_orient<3, 32> : th == 28 ? (void *)k_orient<3, 28> : th == 16 ? (void *)k_orient<3, 16> : (void *)k_orient<3, 8>;
    return th == 32 ? (void *)k_orient<1, 32> : th == 28 ? (void *)k_orient<1, 28> : th == 16 ? (void *)k_orient<1, 16> : (void *)k_orient<1, 8>;
}

// Measurement aid (tools/inflight_shares.py): SBM_DEBUG_SHRINK is a bit mask of graph kernels that are launched with ONE CTA
// instead of their grid -- bit 0 orient L0, 1 spread L0, 2 pyrDown, 3 orient L1, 4 spread L1, 5 similarity, 6 refine.  The
// results are then garbage; what the frame rate gains is what that kernel costs with several frames in flight.
dim3 dbgGrid(int bit, dim3 g)
{
    static const int mask = getenv("SBM_DEBUG_SHRINK") ? atoi(getenv("SBM_DEBUG_SHRINK")) : 0;
    return (mask >> bit) & 1 ? dim3(1, 1, 1) : g;
}

// k_orient_mma's stage A through the copy engine: the source image as a 2-D tensor of bytes [h][w * ch] whose box is the
// tile's staged window (K1M<C, NT>: SROW bytes x NR rows).  cuTensorMapEncodeTiled is looked up through the runtime (no link
// against libcuda).  Returns 1 when the kernel may use the map: 16-byte aligned rows (else its vector / byte loads serve),
// SBM_ORIENT_TMA not 0.
typedef CUresult (*TensorMapEncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                      const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
TensorMapEncodeFn tensorMapEncoder()
{
    static const TensorMapEncodeFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (TensorMapEncodeFn)p;
    }();
    return fn;
}
int orientMap(CUtensorMap *map, const uint8_t *src, int w, int h, int ch, int th, int sms)
{
    static const bool on = orientMma() && !(getenv("SBM_ORIENT_TMA") && atoi(getenv("SBM_ORIENT_TMA")) == 0);
    memset(map, 0, sizeof *map);
    const size_t rowbytes = (size_t)w * ch;
    if (!on || !tensorMapEncoder() || (rowbytes & 15) || (reinterpret_cast<uintptr_t>(src) & 15))
        return 0;
    const int nt = (th + 4) / 8, npc = (ch * (K1_TW + 4) + 15) / 16; // as K1M<C, NT>
    const cuuint64_t gdim[2] = {(cuuint64_t)rowbytes, (cuuint64_t)h}, gstride[1] = {(cuuint64_t)rowbytes};
    const cuuint32_t box[2] = {(cuuint32_t)(16 * (npc + 2)), (cuuint32_t)(8 * (nt + 1))}, estr[2] = {1, 1};
    if (tensorMapEncoder()(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t *>(src), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return 0;
    // 2: a CTA takes two tiles and asks for both windows at once, where that still leaves every SM a CTA (M1 level 0: 464
    // tiles -> 240 CTAs; alone 11.0 -> 13.5 us, with frames in flight 18.35 -> 18.1 us per frame: the second window arrives
    // while the first is worked on, and waiting for the window is 1.5 of a CTA's 6.7 us).  SBM_ORIENT_TILES=1 / 2 forces it.
    static const int forced = getenv("SBM_ORIENT_TILES") ? atoi(getenv("SBM_ORIENT_TILES")) : 0;
    const int rows = (h + th - 1) / th, cols = (w + K1_TW - 1) / K1_TW;
    if (forced == 1 || rows < 2)
        return 1;
    return forced == 2 || cols * ((rows + 1) / 2) >= sms ? 2 : 1;
}
// grid of a K1 launch: one CTA per 64 x th tile, or per two tiles one above the other (use_tma == 2)
dim3 orientGrid(int w, int h, int th, int use_tma)
{
    const int rows = (h + th - 1) / th;
    return dim3((w + K1_TW - 1) / K1_TW, use_tma == 2 ? (rows + 1) / 2 : rows);
}

// one K1 launch on stream st (match path: no dump outputs)
cudaError_t launchOrient(int sms, int ch, const uint8_t *src, int w, int h, const uint8_t *mask, int mask_shift, int mask_w, float weak_sq,
                         uint8_t *quant, cudaStream_t st)
{
    const int th = orientTileH(w, h, sms);
    float *nf = nullptr;
    CUtensorMap map;
    int use_tma = orientMap(&map, src, w, h, ch, th, sms);
    void *argv[] = {&src, &w, &h, &mask, &mask_shift, &mask_w, &weak_sq, &quant, &nf, &nf, &map, &use_tma};
    return cudaLaunchKernel(orientFunc(ch, th), orientGrid(w, h, th, use_tma), dim3(K1_THREADS), argv, 0, st);
}

struct ClassSet
{
    int n = 0;
    std::vector<sbm_template_desc> desc; // [n][levels]
    std::vector<sbm_feature> feats;
};

// geometry of one pyramid level for the current frame shape
struct LevelGeom
{
    int T = 0, w = 0, h = 0, W = 0, H = 0;
    size_t lm_bytes = 0, lm_alloc = 0;
    bool nib = false; // memories packed two responses per byte (the coarsest level when W is even): k_similarity_nib reads them
    bool bits = false; // memories as hit / any bit planes (the coarsest level when W % 32 == 0): k_similarity_bits reads them
    bool tiled = false; // memories tiled 8 rows x 16 bytes per 128-byte line (refined levels with W % 16 == 0): k_refine_tiled
};

// templates of one pyramid level on the device
struct LevelTempl
{
    TemplLevel *d_tl = nullptr;
    Feat *d_feats = nullptr;
};

struct LevelBuf
{
    uint8_t *d_src = nullptr; // level >= 1 (level 0 is the frame itself)
    uint8_t *d_quant = nullptr;
    uint8_t *d_lm = nullptr;       // the level's linear memories (bit planes: K5B_FRONT bytes into the allocation)
    uint8_t *d_lm_alloc = nullptr; // what cudaMalloc returned
    uint8_t *d_spread = nullptr, *d_resp = nullptr; // debug taps
};

template <typename T> void freeDev(T *&p)
{
    if (p)
        cudaFree(p);
    p = nullptr;
}

struct Slot
{
    cudaStream_t stream = nullptr;
    uint64_t shape_gen = 0; // generation of ctx geometry these buffers were allocated for
    bool debug_bufs = false;
    std::vector<LevelBuf> lv;
    uint8_t *d_frame = nullptr; // staging for host frames
    size_t frame_alloc = 0;
    uint8_t *d_mask = nullptr;
    size_t mask_alloc = 0;
    const uint8_t *last_src0 = nullptr;
    // candidates
    uint32_t cap = 1u << 16;
    Cand *d_buf[2] = {nullptr, nullptr};
    uint32_t *d_counts = nullptr;
    Cand *d_final = nullptr; // where the last stage of the most recent frame wrote
    uint32_t *d_final_count = nullptr;
    uint32_t final_cap = 0;
    // pinned host staging
    uint32_t *h_counts = nullptr;
    Cand *h_recs = nullptr;
    uint32_t h_recs_cap = 0;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    bool have_frame = false;
    int launches = 0;
    // CUDA graph of the whole per-frame pipeline on this slot's buffers (see buildGraph)
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t gexec = nullptr;
    uint64_t graph_epoch = 0;
    float graph_threshold = 0.f;
    bool graph_fetch = false;
    const uint8_t *graph_src0 = nullptr;
    // the two kernel nodes that read the level-0 frame, with what is needed to re-issue their parameters
    cudaGraphNode_t n_orient0 = nullptr, n_pyr1 = nullptr;
    void *f_orient0 = nullptr, *f_pyr1 = nullptr;
    dim3 g_orient0, g_pyr1;
    int graph_launches = 0;
    Cand *graph_final = nullptr; // where the graph's last stage writes
    uint32_t *graph_final_count = nullptr;
    uint32_t graph_final_cap = 0;
};
} // namespace

struct sbm_ctx
{
    int device = 0;
    int sms = 148;                 // multiprocessors of `device`
    int refine_grid[2] = {0, 0};   // K5b grid (flat / tiled kernel) = CTAs resident at once on `device`
    int bits_split_ctas[3][5] = {}; // k_similarity_bits<NH, SPLIT>: CTAs of 64 * K threads resident per SM, [NH variant][K]
    std::vector<int> T;
    float weak = 30.f, strong = 60.f;
    int num_features = 63;
    int levels = 0;
    std::mutex mu;

    std::vector<ClassSet> classes;
    std::vector<uint32_t> gid_begin; // per class slot (+ total)
    bool templates_dirty = true;
    uint32_t n_templates = 0;
    std::vector<LevelTempl> tl;

    // frame shape
    int w = 0, h = 0, c = 0;
    uint64_t shape_gen = 0;
    uint64_t epoch = 1; // bumped whenever anything a captured graph depends on changes
    std::vector<LevelGeom> geom;
    Slot slot[SBM_SLOTS];

    // coarse level, per (templates, shape)
    bool coarse_dirty = true;
    TemplCoarse *d_tcoarse = nullptr;
    uint32_t *d_coff = nullptr;
    std::vector<TemplCoarse> h_tcoarse;

    // k_similarity_bits: one SimItem per active template for the current threshold (see ensureSimItems)
    SimItem *d_sim_items = nullptr;
    uint32_t sim_items_cap = 0;
    uint2 *d_sim_offs = nullptr; // [n_active][sim_stride] pre-decoded feature offsets of the active templates
    size_t sim_offs_cap = 0;
    uint32_t sim_stride = 8;
    uint64_t sim_offs_epoch = 0;
    std::vector<uint32_t> h_coff; // host copy of d_coff
    uint64_t sim_epoch = 0;
    float sim_threshold = 0.f;
    int max_coarse_valid = 0; // most in-image coarse features any template has (selects the counter width)

    // active template list
    std::vector<int> active_classes; // as requested (may repeat)
    bool active_dirty = true;
    uint32_t *d_active = nullptr;
    uint32_t n_active = 0;
    uint64_t coarse_evals = 0;

    bool debug = false, timing = false;
    float last_ms[4] = {0, 0, 0, 0};
    uint64_t stats[8] = {0};

    // multi-GPU group (sbm_group_*): this context is rank `rank` of `n` contexts, one per GPU, one NCCL communicator
    struct Group
    {
        bool on = false;
        int n = 1, rank = 0, mode = 0;
        NcclApi::comm_t comm = nullptr;
        std::vector<int> global_n;  // template modes: size of the FULL class per class slot (this rank keeps a slice)
        uint8_t *d_send = nullptr, *d_recv = nullptr; // [n_frames] / [n][n_frames] result blocks (header + records)
        size_t send_alloc = 0, recv_alloc = 0;
        uint8_t *h_recv = nullptr; // pinned copy of d_recv
        size_t h_alloc = 0;
        uint32_t cap = 256; // records per frame block of the host-side group calls (grown on overflow)
        // host-side group calls: the exchange + D2H of a chunk of frames run on their own stream while the next chunk
        // uploads, and the host orders chunk k's lists meanwhile
        cudaStream_t xstream = nullptr;
        std::vector<cudaEvent_t> chunk_done;
        std::vector<sbm_match_rec> pool; // final records of the chunks, before they are laid out list by list
    } group;
};

namespace
{
void freeSlotShape(Slot &s)
{
    for (auto &l : s.lv)
    {
        freeDev(l.d_src);
        freeDev(l.d_quant);
        freeDev(l.d_lm_alloc);
        l.d_lm = nullptr;
        freeDev(l.d_spread);
        freeDev(l.d_resp);
    }
    s.have_frame = false;
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

// frame shape -> level geometry (no allocation; slots follow lazily through shape_gen)
// SBM_COARSE_NIBBLES=0 keeps one byte per response at the coarsest level too (the layout of every other level)
bool coarseNibbles()
{
    static const bool on = [] {
        const char *e = getenv("SBM_COARSE_NIBBLES");
        return !(e && atoi(e) == 0);
    }();
    return on;
}

// SBM_COARSE_BITS=0 keeps the byte / nibble layouts at the coarsest level whatever its width
bool coarseBits()
{
    static const bool on = [] {
        const char *e = getenv("SBM_COARSE_BITS");
        return !(e && atoi(e) == 0);
    }();
    return on;
}

// SBM_REFINE_TILED=0 keeps the refined levels flat (k_refine) whatever their shape, =1 tiles them where the shape allows
#ifndef SBM_REFINE_TILED_DEFAULT
#define SBM_REFINE_TILED_DEFAULT 1
#endif
bool refineTiled()
{
    static const bool on = [] {
        const char *e = getenv("SBM_REFINE_TILED");
        return e ? atoi(e) != 0 : SBM_REFINE_TILED_DEFAULT;
    }();
    return on;
}

int ensureShape(sbm_ctx *c, int w, int h, int ch)
{
    if (c->w == w && c->h == h && c->c == ch && !c->geom.empty())
        return 0;
    int rc = checkShape(c, w, h, ch);
    if (rc)
        return rc;
    c->geom.assign(c->levels, LevelGeom());
    int lw = w, lh = h;
    for (int l = 0; l < c->levels; ++l)
    {
        LevelGeom &G = c->geom[l];
        if (l > 0)
            lw /= 2, lh /= 2;
        G.T = c->T[l], G.w = lw, G.h = lh, G.W = lw / G.T, G.H = lh / G.T;
        const size_t px = (size_t)lw * lh, WH = (size_t)G.W * G.H;
        G.lm_bytes = 8 * px;
        // tail slack: flat addressing lets the last memory be over-read: a whole memory for the
        // coarse scan past template_positions, 16 rows + 16 for the local patch, + the realigning vector loads
        // (K5a runs whole warps: up to 31 lanes x 16 positions beyond the last reported one)
        G.lm_alloc = G.lm_bytes + std::max(WH, (size_t)16 * G.W + 16) + 1024;
        G.bits = l == c->levels - 1 && (G.W & 31) == 0 && coarseBits();
        G.nib = !G.bits && l == c->levels - 1 && (G.W & 1) == 0 && coarseNibbles();
        if (G.bits)
        {
            // 2 bits per position; behind the memories: the zero words the padded offsets read (two parity classes), a
            // whole map for the scan past template_positions, and a tile + a warp of positions for the whole-warp loads
            const size_t positions = (G.lm_bytes + 63) / 64 * 64 + 64 + (WH + K5B_TILE - 1) / K5B_TILE * K5B_TILE + 2 * K5B_TILE;
            G.lm_alloc = K5B_FRONT + positions / 4 + 256;
        }
        // K5b's levels: tiles need whole 16-byte segments per row and an orientation stride of whole tile rows
        G.tiled = l < c->levels - 1 && (G.W & 15) == 0 && (((size_t)G.T * G.T * G.H) & 7) == 0 && refineTiled();
        if (G.tiled) // rows up to 16 (+1 for the wrap) behind the last memory, rounded up to a tile row
            G.lm_alloc += (size_t)24 * G.W;
    }
    c->w = w, c->h = h, c->c = ch;
    ++c->shape_gen;
    ++c->epoch;
    c->coarse_dirty = true;
    return 0;
}

// K5b walks the candidate list grid-stride with a static split, so the grid is exactly the number of CTAs that are
// resident at once on the context's device (a CTA that starts after the others have finished would do its whole share
// alone).  Computed once per context in sbm_create, for c->device.
int computeRefineGrid(int sms, bool tiled)
{
    int per_sm = 0;
    const cudaError_t e = tiled ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_refine_tiled, K6_THREADS, 0)
                                : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_refine, K6_THREADS, 0);
    if (e != cudaSuccess || per_sm < 1)
        per_sm = 1;
    if (const char *ev = getenv("SBM_REFINE_CTAS_PER_SM"))
        per_sm = std::max(1, atoi(ev));
    return sms * per_sm;
}
int refineGrid(const sbm_ctx *c, bool tiled) { return c->refine_grid[tiled ? 1 : 0]; }

// the slot's stream and its events are created together (the timers rely on the events)
int ensureStream(Slot &s)
{
    if (!s.stream)
        CK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    for (int i = 0; i < 4; ++i)
        if (!s.ev[i])
            CK(cudaEventCreate(&s.ev[i]));
    return 0;
}

int ensureSlot(sbm_ctx *c, Slot &s)
{
    int rc0;
    if ((rc0 = ensureStream(s)))
        return rc0;
    if (s.shape_gen != c->shape_gen || (c->debug && !s.debug_bufs))
    {
        freeSlotShape(s);
        s.lv.assign(c->levels, LevelBuf());
        for (int l = 0; l < c->levels; ++l)
        {
            const LevelGeom &G = c->geom[l];
            LevelBuf &L = s.lv[l];
            const size_t px = (size_t)G.w * G.h;
            if (l > 0)
                CK(cudaMalloc(&L.d_src, px * c->c));
            CK(cudaMalloc(&L.d_quant, px));
            CK(cudaMalloc(&L.d_lm_alloc, G.lm_alloc));
            CK(cudaMemset(L.d_lm_alloc, 0, G.lm_alloc));
            L.d_lm = L.d_lm_alloc + (G.bits ? K5B_FRONT : 0);
            if (c->debug)
            {
                CK(cudaMalloc(&L.d_spread, px));
                CK(cudaMalloc(&L.d_resp, 8 * px));
            }
        }
        s.shape_gen = c->shape_gen;
        s.debug_bufs = c->debug;
        s.graph_epoch = 0;
    }
    if (!s.d_buf[0])
    {
        CK(cudaMalloc(&s.d_buf[0], (size_t)s.cap * sizeof(Cand)));
        CK(cudaMalloc(&s.d_buf[1], (size_t)s.cap * sizeof(Cand)));
    }
    if (!s.d_counts)
    {
        CK(cudaMalloc(&s.d_counts, COUNT_WORDS * sizeof(uint32_t)));
        CK(cudaMallocHost(&s.h_counts, COUNT_WORDS * sizeof(uint32_t)));
    }
    if (!s.h_recs)
    {
        s.h_recs_cap = HEAD_RECS;
        CK(cudaMallocHost(&s.h_recs, (size_t)s.h_recs_cap * sizeof(Cand)));
    }
    return 0;
}

void destroyGraph(Slot &s)
{
    if (s.gexec)
        cudaGraphExecDestroy(s.gexec);
    if (s.graph)
        cudaGraphDestroy(s.graph);
    s.gexec = nullptr, s.graph = nullptr, s.graph_epoch = 0;
}

void destroySlot(Slot &s)
{
    destroyGraph(s);
    freeSlotShape(s);
    freeDev(s.d_frame);
    freeDev(s.d_mask);
    freeDev(s.d_buf[0]);
    freeDev(s.d_buf[1]);
    freeDev(s.d_counts);
    if (s.h_counts)
        cudaFreeHost(s.h_counts);
    if (s.h_recs)
        cudaFreeHost(s.h_recs);
    for (int i = 0; i < 4; ++i)
        if (s.ev[i])
            cudaEventDestroy(s.ev[i]);
    if (s.stream)
        cudaStreamDestroy(s.stream);
    s = Slot();
}

int ensureTemplates(sbm_ctx *c)
{
    if (c->templates_dirty)
    {
        CK(cudaDeviceSynchronize()); // frames in flight may still read the old arrays
        const int L = c->levels;
        c->gid_begin.assign(c->classes.size() + 1, 0);
        for (size_t k = 0; k < c->classes.size(); ++k)
            c->gid_begin[k + 1] = c->gid_begin[k] + (uint32_t)c->classes[k].n;
        c->n_templates = c->gid_begin.back();
        c->tl.resize(L);
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
            LevelTempl &LT = c->tl[l];
            freeDev(LT.d_tl);
            freeDev(LT.d_feats);
            CK(cudaMalloc(&LT.d_tl, std::max<size_t>(1, tl.size()) * sizeof(TemplLevel)));
            CK(cudaMalloc(&LT.d_feats, std::max<size_t>(1, ff.size()) * sizeof(Feat)));
            if (!tl.empty())
                CK(cudaMemcpy(LT.d_tl, tl.data(), tl.size() * sizeof(TemplLevel), cudaMemcpyHostToDevice));
            if (!ff.empty())
                CK(cudaMemcpy(LT.d_feats, ff.data(), ff.size() * sizeof(Feat), cudaMemcpyHostToDevice));
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
        const LevelGeom &G = c->geom[L - 1];
        const int T = G.T, W = G.W, H = G.H;
        const uint32_t WH = (uint32_t)W * H;
        c->h_tcoarse.assign(c->n_templates, TemplCoarse());
        c->max_coarse_valid = 0;
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
                            if (f.x < 0 || f.x >= G.w || f.y < 0 || f.y >= G.h)
                                continue;
                            if (f.label < 0 || f.label > 7)
                                return fail(SBM_ERR_BAD_ARG, "feature label %d out of range", f.label);
                            uint32_t a = (uint32_t)f.label * T * T * WH + (uint32_t)((f.y % T) * T + f.x % T) * WH +
                                         (uint32_t)(f.y / T) * W + (uint32_t)(f.x / T);
                            // k_similarity_nib has two shift classes (bit 3 of the offset): buckets 1 and 3 only, half the padding
                            // k_similarity_bits: the parity of the stream's first word pair (bit 5), buckets 1 and 3 as well
                            // k_similarity_bits shifts by off & 31 whatever it is: one bucket (3), padded once
                            cls[G.bits ? 3 : G.nib ? ((a >> 3) & 1) * 2 + 1 : (a >> 2) & 3].push_back(a);
                        }
                        int nv = 0, stored = 0;
                        for (int k = 0; k < 4; ++k)
                        {
                            nv += (int)cls[k].size();
                            // pad with reads of the zeroed tail (same word shift): adds 0, keeps batches full
                            // (bit planes: the zero words behind the memories, first pair even for bucket 1, odd for bucket 3)
                            // (bit planes: the zero words behind the memories, first pair even for bucket 1, odd for bucket 3)
                            const uint32_t zero_off = G.bits ? (uint32_t)((G.lm_bytes + 63) / 64 * 64) : (uint32_t)G.lm_bytes + 4u * k;
                            while (cls[k].size() % K5_U)
                                cls[k].push_back(zero_off);
                            off.insert(off.end(), cls[k].begin(), cls[k].end());
                            stored += (int)cls[k].size();
                            o.cls_end[k] = (uint16_t)stored;
                        }
                        o.n_valid = nv;
                        c->max_coarse_valid = std::max(c->max_coarse_valid, nv);
                    }
                    fo += d.num_features;
                }
        }
        CK(cudaDeviceSynchronize());
        freeDev(c->d_tcoarse);
        freeDev(c->d_coff);
        CK(cudaMalloc(&c->d_tcoarse, std::max<size_t>(1, c->h_tcoarse.size()) * sizeof(TemplCoarse)));
        CK(cudaMalloc(&c->d_coff, std::max<size_t>(1, off.size()) * sizeof(uint32_t)));
        if (!c->h_tcoarse.empty())
            CK(cudaMemcpy(c->d_tcoarse, c->h_tcoarse.data(), c->h_tcoarse.size() * sizeof(TemplCoarse), cudaMemcpyHostToDevice));
        if (!off.empty())
            CK(cudaMemcpy(c->d_coff, off.data(), off.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        if (G.bits)
            c->h_coff.swap(off);
        else
            std::vector<uint32_t>().swap(c->h_coff);
        c->coarse_dirty = false;
        ++c->epoch;
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
    // the previous list may still be read by frames in flight on the slot streams
    CK(cudaDeviceSynchronize());
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
            if (g < c->h_tcoarse.size()) // empty until a frame shape is known
                c->coarse_evals += (uint64_t)c->h_tcoarse[g].n_valid * (uint64_t)std::max(c->h_tcoarse[g].tp, 0);
        }
    freeDev(c->d_active);
    CK(cudaMalloc(&c->d_active, std::max<size_t>(1, act.size()) * sizeof(uint32_t)));
    if (!act.empty())
        CK(cudaMemcpy(c->d_active, act.data(), act.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    c->n_active = (uint32_t)act.size();
    c->active_dirty = false;
    ++c->epoch;
    return 0;
}

// everything a match call needs before it can enqueue on slot s
cudaError_t launchSpread(const LevelGeom &G, const uint8_t *quant, uint8_t *lm, uint8_t *dbg_spread, uint8_t *dbg_resp, cudaStream_t st);
cudaError_t launchRefine(bool tiled, int grid, const uint8_t *lm, const TemplLevel *tl, const Feat *feats, const Cand *in, const uint32_t *in_count,
                         uint32_t in_cap, Cand *out, uint32_t *out_count, uint32_t out_cap, int w, int h, int T, float threshold, uint16_t *dump,
                         int raw_center, unsigned long long *local_evals, cudaStream_t st);

// ---- template sharding arithmetic ----
// templates of a class of n owned by `rank` of `world`: interleaved ids rank, rank + world, ...; or the contiguous block
// that follows the blocks of the lower ranks (blocks differ by at most one template)
int shardCount(int n, int rank, int world, int mode)
{
    if (mode == SBM_SHARD_INTERLEAVED)
        return n > rank ? (n - rank + world - 1) / world : 0;
    return n / world + (rank < n % world ? 1 : 0);
}
int shardBegin(int n, int rank, int world) { return rank * (n / world) + std::min(rank, n % world); } // contiguous mode
bool shardOwns(int t, int n, int rank, int world, int mode)
{
    if (mode == SBM_SHARD_INTERLEAVED)
        return t % world == rank;
    const int b = shardBegin(n, rank, world);
    return t >= b && t < b + shardCount(n, rank, world, mode);
}

// smallest raw sum whose score passes:  (raw*100.f)/(4*nf) > threshold  (line2Dup.cpp:1206-1208), in the same IEEE single
// operations as the kernels' score_of (min_passing_raw in sbm_kernels.cuh computes the same number on the device)
int minPassingRawHost(int nf, float threshold)
{
    auto score = [nf](int raw) {
        volatile float a = (float)raw * 100.f;
        volatile float b = a / (float)(4 * nf);
        return (float)b;
    };
    const int top = 4 * nf + 1;
    if (!(score(top) > threshold))
        return 0x7fffffff;
    int lo = 0, hi = top; // score is monotone in raw
    while (lo < hi)
    {
        const int mid = (lo + hi) / 2;
        if (score(mid) > threshold)
            hi = mid;
        else
            lo = mid + 1;
    }
    return lo;
}

SimItem makeSimItem(const sbm_ctx *c, uint32_t gid, float threshold)
{
    const TemplCoarse &t = c->h_tcoarse[gid];
    SimItem it = {};
    it.tgid = gid;
    it.nf = t.nf;
    it.tp = t.tp;
    it.rawmin = minPassingRawHost(t.nf, threshold);
    it.n_off = t.cls_end[3];
    return it;
}

// offsets of template gid as k_similarity_bits stages them: .x = byte offset of the first word pair, .y = bit shift
void decodeOffsets(const sbm_ctx *c, uint32_t gid, uint2 *dst)
{
    const TemplCoarse &t = c->h_tcoarse[gid];
    for (int i = 0; i < (int)t.cls_end[3]; ++i)
    {
        // stream of position j = bit (o + j) of the flat bit sequence: first pair o >> 5, shift o & 31
        const uint32_t o = c->h_coff[t.off_begin + i];
        dst[i] = make_uint2(K5B_FRONT + (o >> 5) * 8u, o & 31u);
    }
}

// what k_similarity_bits reads per active template: the offsets (fixed stride; depend on the active list and the frame
// shape) and the SimItem records (depend on the threshold as well)
int ensureSimItems(sbm_ctx *c, float threshold)
{
    if (c->geom.empty() || !c->geom[c->levels - 1].bits)
        return 0;
    const bool offs_stale = c->sim_offs_epoch != c->epoch;
    if (!offs_stale && c->sim_epoch == c->epoch && memcmp(&c->sim_threshold, &threshold, sizeof(float)) == 0)
        return 0;
    std::vector<uint32_t> act;
    {
        std::vector<int> uniq = c->active_classes;
        std::sort(uniq.begin(), uniq.end());
        uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
        for (int k : uniq)
            for (uint32_t g = c->gid_begin[k]; g < c->gid_begin[k + 1]; ++g)
                act.push_back(g);
    }
    CK(cudaDeviceSynchronize()); // frames in flight may still read the previous arrays
    bool realloc = false;
    if (offs_stale)
    {
        uint32_t stride = 8;
        for (uint32_t g : act)
            stride = std::max<uint32_t>(stride, c->h_tcoarse[g].cls_end[3]);
        std::vector<uint2> offs((size_t)act.size() * stride, make_uint2(0u, 0u));
        for (size_t i = 0; i < act.size(); ++i)
            decodeOffsets(c, act[i], offs.data() + i * stride);
        if (c->sim_offs_cap < offs.size() || !c->d_sim_offs)
        {
            freeDev(c->d_sim_offs);
            c->sim_offs_cap = std::max<size_t>(8, offs.size());
            CK(cudaMalloc(&c->d_sim_offs, c->sim_offs_cap * sizeof(uint2)));
            realloc = true;
        }
        if (!offs.empty())
            CK(cudaMemcpy(c->d_sim_offs, offs.data(), offs.size() * sizeof(uint2), cudaMemcpyHostToDevice));
        if (stride != c->sim_stride)
            realloc = true; // the stride is a kernel argument baked into the graphs
        c->sim_stride = stride;
    }
    std::vector<SimItem> items(act.size());
    for (size_t i = 0; i < act.size(); ++i)
        items[i] = makeSimItem(c, act[i], threshold);
    if (c->sim_items_cap < items.size() || !c->d_sim_items)
    {
        freeDev(c->d_sim_items);
        c->sim_items_cap = (uint32_t)std::max<size_t>(1, items.size());
        CK(cudaMalloc(&c->d_sim_items, c->sim_items_cap * sizeof(SimItem)));
        realloc = true;
    }
    if (!items.empty())
        CK(cudaMemcpy(c->d_sim_items, items.data(), items.size() * sizeof(SimItem), cudaMemcpyHostToDevice));
    if (realloc)
        ++c->epoch; // pointers / stride are baked into the slots' graphs
    c->sim_epoch = c->sim_offs_epoch = c->epoch;
    c->sim_threshold = threshold;
    return 0;
}

int prepare(sbm_ctx *c, Slot &s, int w, int h, int ch, const int *class_idx, int n_classes, float threshold)
{
    int rc;
    if ((rc = ensureShape(c, w, h, ch)) || (rc = ensureTemplates(c)) || (rc = ensureActive(c, class_idx, n_classes)) ||
        (rc = ensureSimItems(c, threshold)) || (rc = ensureSlot(c, s)))
        return rc;
    return 0;
}

// The coarse similarity launch of a level in whichever layout it has: function, shape and an argument block that
// stays valid until the launch / graph node is created.  items: NULL = the context's active list.
struct CoarseLaunch
{
    void *func = nullptr;
    dim3 grid, block;
    const uint8_t *lm;
    const uint32_t *coff, *active;
    const TemplCoarse *tc;
    const SimItem *items;
    const uint2 *offs;
    uint32_t stride;
    int W, HW, T;
    float threshold;
    Cand *out;
    uint32_t *counter;
    uint32_t cap;
    uint16_t *dump;
    void *argv[12];
};

// items / offs: NULL = the context's active list (ensureSimItems); else one-off arrays with their own stride
void makeCoarse(CoarseLaunch &L, const sbm_ctx *c, const uint8_t *lm, const SimItem *items, const uint2 *offs, uint32_t stride, const uint32_t *active,
                uint32_t n_active, float threshold, Cand *out, uint32_t *counter, uint32_t cap, uint16_t *dump)
{
    const LevelGeom &CG = c->geom[c->levels - 1];
    // (the bit-plane kernel addresses from the start of the allocation: its staged offsets carry K5B_FRONT)
    L.lm = CG.bits ? lm - K5B_FRONT : lm, L.coff = c->d_coff, L.active = active, L.tc = c->d_tcoarse, L.items = items ? items : c->d_sim_items;
    L.offs = items ? offs : c->d_sim_offs, L.stride = items ? stride : c->sim_stride;
    L.W = CG.W, L.HW = CG.W * CG.H, L.T = CG.T, L.threshold = threshold, L.out = out, L.counter = counter, L.cap = cap, L.dump = dump;
    if (CG.bits)
    {
        // counter planes: 127 features fit 7 planes, 1023 fit 10, 8191 fit 13
        const int nh = c->max_coarse_valid <= 127 ? 0 : c->max_coarse_valid <= 1023 ? 1 : 2;
        L.grid = dim3(n_active, (L.HW + K5B_TILE - 1) / K5B_TILE);
        // What bounds the kernel is issue latency per warp, so small workloads want more warps: split each tile's features
        // over K warp groups, K the smallest that gives every scheduler three warps, as long as the whole grid still
        // fits the SMs at once (sbm_create asked the occupancy calculator)
        const long ctas = (long)L.grid.x * L.grid.y;
        int K = 1;
        for (int k = 2; k <= K5B_SPLIT_MAX && ctas * (K5B_THREADS / 32) * K < 12L * c->sms; ++k)
            if (ctas <= (long)c->sms * c->bits_split_ctas[nh][k])
                K = k;
        if (const char *ev = getenv("SBM_BITS_SPLIT"))
            K = std::max(1, std::min(K5B_SPLIT_MAX, atoi(ev)));
        static void *const plain[3] = {(void *)k_similarity_bits<3, false>, (void *)k_similarity_bits<6, false>, (void *)k_similarity_bits<9, false>};
        static void *const split[3] = {(void *)k_similarity_bits<3, true>, (void *)k_similarity_bits<6, true>, (void *)k_similarity_bits<9, true>};
        L.func = K > 1 ? split[nh] : plain[nh];
        L.block = dim3(K5B_THREADS * K);
        void *a[] = {&L.lm, &L.offs, &L.stride, &L.items, &L.W, &L.HW, &L.T, &L.out, &L.counter, &L.cap, &L.dump};
        memcpy(L.argv, a, sizeof(a));
    }
    else
    {
        L.func = CG.nib ? (void *)k_similarity_nib : (void *)k_similarity;
        L.grid = dim3(n_active, (L.HW + K5_TILE - 1) / K5_TILE), L.block = dim3(K5_THREADS);
        void *a[] = {&L.lm, &L.coff, &L.tc, &L.active, &L.W, &L.HW, &L.T, &L.threshold, &L.out, &L.counter, &L.cap, &L.dump};
        memcpy(L.argv, a, sizeof(a));
    }
}

cudaError_t launchCoarse(const sbm_ctx *c, const uint8_t *lm, const SimItem *items, const uint2 *offs, uint32_t stride, const uint32_t *active,
                         uint32_t n_active, float threshold, Cand *out, uint32_t *counter, uint32_t cap, uint16_t *dump, cudaStream_t st)
{
    CoarseLaunch L;
    makeCoarse(L, c, lm, items, offs, stride, active, n_active, threshold, out, counter, cap, dump);
    return cudaLaunchKernel(L.func, L.grid, L.block, L.argv, 0, st);
}

// Enqueue the whole pipeline for one frame resident at d_img on stream st, using slot s's scratch.
// user_recs/user_count: optional caller buffers for the last stage.
int enqueueFrame(sbm_ctx *c, Slot &s, const uint8_t *d_img, const uint8_t *d_mask, float threshold, cudaStream_t st, Cand *user_recs,
                 uint32_t user_cap, uint32_t *user_count)
{
    const int L = c->levels;
    int launches = 0;
    CK(cudaMemsetAsync(s.d_counts, 0, COUNT_WORDS * sizeof(uint32_t), st));
    if (user_count)
        CK(cudaMemsetAsync(user_count, 0, 2 * sizeof(uint32_t), st));
    if (c->timing)
        CK(cudaEventRecord(s.ev[0], st));
    const float weak_sq = c->weak * c->weak;
    const uint8_t *src = d_img;
    for (int l = 0; l < L; ++l)
    {
        const LevelGeom &G = c->geom[l];
        LevelBuf &LB = s.lv[l];
        if (l > 0)
        {
            const LevelGeom &P = c->geom[l - 1];
            dim3 g((G.w + 31) / 32, (G.h + K1B_ROWS - 1) / K1B_ROWS);
            if (c->c == 3)
                k_pyrdown<3><<<g, K1B_THREADS, 0, st>>>(src, P.w, P.h, LB.d_src, G.w, G.h);
            else
                k_pyrdown<1><<<g, K1B_THREADS, 0, st>>>(src, P.w, P.h, LB.d_src, G.w, G.h);
            src = LB.d_src;
            ++launches;
        }
        CK(launchOrient(c->sms, c->c, src, G.w, G.h, d_mask, l, c->w, weak_sq, LB.d_quant, st));
        uint8_t *ds = c->debug ? LB.d_spread : nullptr, *dr = c->debug ? LB.d_resp : nullptr;
        CK(launchSpread(G, LB.d_quant, LB.d_lm, ds, dr, st));
        launches += 2;
    }
    if (c->timing)
        CK(cudaEventRecord(s.ev[1], st));

    // coarse similarity + threshold at the last level
    const LevelGeom &CG = c->geom[L - 1];
    const int HW = CG.W * CG.H;
    Cand *out = (L == 1 && user_recs) ? user_recs : s.d_buf[0];
    uint32_t out_cap = (L == 1 && user_recs) ? user_cap : s.cap;
    uint32_t *out_cnt = (L == 1 && user_count) ? user_count : s.d_counts + 0;
    if (c->n_active > 0)
    {
        CK(launchCoarse(c, s.lv[L - 1].d_lm, nullptr, nullptr, 0, c->d_active, c->n_active, threshold, out, out_cnt, out_cap, nullptr, st));
        ++launches;
    }
    if (c->timing)
        CK(cudaEventRecord(s.ev[3], st));
    Cand *in = out;
    uint32_t in_cap = out_cap;
    uint32_t *in_cnt = out_cnt;
    int stage = 1;
    for (int l = L - 2; l >= 0; --l, ++stage)
    {
        const LevelGeom &G = c->geom[l];
        bool last = (l == 0);
        out = (last && user_recs) ? user_recs : s.d_buf[stage & 1];
        out_cap = (last && user_recs) ? user_cap : s.cap;
        out_cnt = (last && user_count) ? user_count : s.d_counts + stage;
        if (c->n_active > 0)
        {
            CK(launchRefine(G.tiled, refineGrid(c, G.tiled), s.lv[l].d_lm, c->tl[l].d_tl, c->tl[l].d_feats, in, in_cnt, in_cap, out, out_cnt, out_cap, G.w,
                            G.h, G.T, threshold, nullptr, 0, reinterpret_cast<unsigned long long *>(s.d_counts + 10), st));
            ++launches;
        }
        in = out, in_cap = out_cap, in_cnt = out_cnt;
    }
    if (user_count)
    {
        // stages 0 .. L-2 counted in the slot's counters (the last one in user_count[0] itself)
        k_overflow_flag<<<1, 1, 0, st>>>(user_count, user_cap, s.d_counts, L - 1, s.cap);
        ++launches;
    }
    if (c->timing)
        CK(cudaEventRecord(s.ev[2], st));
    s.d_final = in;
    s.d_final_count = in_cnt;
    s.final_cap = in_cap;
    s.launches = launches;
    CK(cudaGetLastError());
    s.have_frame = true;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// The per-frame pipeline as a CUDA graph.  All launches of a frame (1 memset + 3L+1+(L-1) kernels, and for the
// host path the two D2H copies of the result head) become ONE cudaGraphLaunch: the CPU cost per frame drops from
// ~12 driver calls to 2, and the two pyramid branches (orient/spread of level 0 vs pyrDown/orient/spread of
// level 1) are independent nodes that the GPU may overlap.  The graph is rebuilt when anything baked into it
// changes (frame shape, template set, active classes, threshold, buffer reallocation) -- tracked by ctx->epoch.
// Only the level-0 source pointer is patched per launch (cudaGraphExecKernelNodeSetParams) when the caller's
// frame lives somewhere else than last time.  Masks, debug taps and stage timing use the eager path.
// ------------------------------------------------------------------------------------------------
template <typename... A> int addKernel(cudaGraph_t g, cudaGraphNode_t *node, const std::vector<cudaGraphNode_t> &deps, void *func, dim3 grid,
                                       dim3 block, A... args)
{
    void *argv[] = {(void *)&args...};
    cudaKernelNodeParams p = {};
    p.func = func;
    p.gridDim = grid;
    p.blockDim = block;
    p.sharedMemBytes = 0;
    p.kernelParams = argv;
    p.extra = nullptr;
    CK(cudaGraphAddKernelNode(node, g, deps.data(), deps.size(), &p));
    return 0;
}

void *spreadFunc(int T, bool nib, bool tiled, bool bits)
{
    if (bits)
        return T == 4 ? (void *)k_spread_linearize<4, false, false, true> : T == 8 ? (void *)k_spread_linearize<8, false, false, true>
               : T == 2 ? (void *)k_spread_linearize<2, false, false, true> : (void *)k_spread_linearize<16, false, false, true>;
    if (nib)
        return T == 4 ? (void *)k_spread_linearize<4, true> : T == 8 ? (void *)k_spread_linearize<8, true>
               : T == 2 ? (void *)k_spread_linearize<2, true> : (void *)k_spread_linearize<16, true>;
    if (tiled)
        return T == 4 ? (void *)k_spread_linearize<4, false, true> : T == 8 ? (void *)k_spread_linearize<8, false, true>
               : T == 2 ? (void *)k_spread_linearize<2, false, true> : (void *)k_spread_linearize<16, false, true>;
    return T == 4 ? (void *)k_spread_linearize<4> : T == 8 ? (void *)k_spread_linearize<8> : T == 2 ? (void *)k_spread_linearize<2>
                                                                                                     : (void *)k_spread_linearize<16>;
}

void *refineFunc(bool tiled) { return tiled ? (void *)k_refine_tiled : (void *)k_refine; }

cudaError_t launchRefine(bool tiled, int grid, const uint8_t *lm, const TemplLevel *tl, const Feat *feats, const Cand *in, const uint32_t *in_count,
                         uint32_t in_cap, Cand *out, uint32_t *out_count, uint32_t out_cap, int w, int h, int T, float threshold, uint16_t *dump,
                         int raw_center, unsigned long long *local_evals, cudaStream_t st)
{
    void *argv[] = {(void *)&lm,  (void *)&tl,        (void *)&feats,   (void *)&in, (void *)&in_count, (void *)&in_cap,
                    (void *)&out, (void *)&out_count, (void *)&out_cap, (void *)&w,  (void *)&h,        (void *)&T,
                    (void *)&threshold, (void *)&dump, (void *)&raw_center, (void *)&local_evals};
    return cudaLaunchKernel(refineFunc(tiled), dim3(grid), dim3(K6_THREADS), argv, 0, st);
}

cudaError_t launchSpread(const LevelGeom &G, const uint8_t *quant, uint8_t *lm, uint8_t *dbg_spread, uint8_t *dbg_resp, cudaStream_t st)
{
    dim3 g2((G.w + K2_PW - 1) / K2_PW, (G.h + K2_PH - 1) / K2_PH);
    int w = G.w, h = G.h;
    void *argv[] = {(void *)&quant, (void *)&w, (void *)&h, (void *)&lm, (void *)&dbg_spread, (void *)&dbg_resp};
    return cudaLaunchKernel(spreadFunc(G.T, G.nib, G.tiled, G.bits), g2, dim3(G.bits ? K2_THREADS_BITS : K2_THREADS), argv, 0, st);
}

int buildGraph(sbm_ctx *c, Slot &s, const uint8_t *src0, float threshold, bool with_fetch)
{
    destroyGraph(s);
    const int L = c->levels;
    CK(cudaGraphCreate(&s.graph, 0));
    cudaGraph_t g = s.graph;
    int launches = 0;
    s.n_orient0 = s.n_pyr1 = nullptr;

    cudaGraphNode_t n_memset;
    {
        cudaMemsetParams mp = {};
        mp.dst = s.d_counts;
        mp.value = 0;
        mp.elementSize = 4;
        mp.width = COUNT_WORDS;
        mp.height = 1;
        mp.pitch = COUNT_WORDS * 4;
        CK(cudaGraphAddMemsetNode(&n_memset, g, nullptr, 0, &mp));
    }
    const float weak_sq = c->weak * c->weak;
    const uint8_t *null_mask = nullptr;
    uint8_t *null_dbg = nullptr;
    float *null_f = nullptr;
    std::vector<cudaGraphNode_t> spread_node(L);
    cudaGraphNode_t prev_src_node = nullptr; // producer of this level's source image
    const uint8_t *src = src0;
    int rc;
    for (int l = 0; l < L; ++l)
    {
        const LevelGeom &G = c->geom[l];
        LevelBuf &LB = s.lv[l];
        std::vector<cudaGraphNode_t> dep_src;
        if (l > 0)
        {
            const LevelGeom &P = c->geom[l - 1];
            if (prev_src_node)
                dep_src.push_back(prev_src_node);
            cudaGraphNode_t n_pyr;
            dim3 gp((G.w + 31) / 32, (G.h + K1B_ROWS - 1) / K1B_ROWS);
            void *f = c->c == 3 ? (void *)k_pyrdown<3> : (void *)k_pyrdown<1>;
            if ((rc = addKernel(g, &n_pyr, dep_src, f, dbgGrid(2, gp), dim3(K1B_THREADS), src, P.w, P.h, LB.d_src, G.w, G.h)))
                return rc;
            if (l == 1)
                s.n_pyr1 = n_pyr, s.f_pyr1 = f, s.g_pyr1 = dbgGrid(2, gp);
            src = LB.d_src;
            prev_src_node = n_pyr;
            dep_src.assign(1, n_pyr);
            ++launches;
        }
        const int th = orientTileH(G.w, G.h, c->sms);
        CUtensorMap map;
        const int use_tma = orientMap(&map, src, G.w, G.h, c->c, th, c->sms);
        dim3 g1 = orientGrid(G.w, G.h, th, use_tma);
        cudaGraphNode_t n_or, n_sp;
        int mask_shift = l, mask_w = c->w;
        if ((rc = addKernel(g, &n_or, dep_src, orientFunc(c->c, th), dbgGrid(l == 0 ? 0 : 3, g1), dim3(K1_THREADS), src, G.w, G.h, null_mask, mask_shift, mask_w, weak_sq,
                            LB.d_quant, null_f, null_f, map, use_tma)))
            return rc;
        if (l == 0)
            s.n_orient0 = n_or, s.f_orient0 = orientFunc(c->c, th), s.g_orient0 = dbgGrid(0, g1);
        dim3 g2((G.w + K2_PW - 1) / K2_PW, (G.h + K2_PH - 1) / K2_PH);
        const uint8_t *qin = LB.d_quant;
        if ((rc = addKernel(g, &n_sp, {n_or}, spreadFunc(G.T, G.nib, G.tiled, G.bits), dbgGrid(l == 0 ? 1 : 4, g2), dim3(G.bits ? K2_THREADS_BITS : K2_THREADS), qin, G.w, G.h, LB.d_lm, null_dbg, null_dbg)))
            return rc;
        spread_node[l] = n_sp;
        launches += 2;
    }
    // coarse similarity
    const LevelGeom &CG = c->geom[L - 1];
    const int HW = CG.W * CG.H;
    cudaGraphNode_t tail = n_memset;
    Cand *in = s.d_buf[0];
    uint32_t in_cap = s.cap;
    uint32_t *in_cnt = s.d_counts + 0;
    if (c->n_active > 0)
    {
        CoarseLaunch CL;
        makeCoarse(CL, c, s.lv[L - 1].d_lm, nullptr, nullptr, 0, c->d_active, c->n_active, threshold, in, in_cnt, in_cap, nullptr);
        cudaGraphNode_t n_sim;
        {
            const cudaGraphNode_t deps[2] = {n_memset, spread_node[L - 1]};
            cudaKernelNodeParams kp = {};
            kp.func = CL.func, kp.gridDim = dbgGrid(5, CL.grid), kp.blockDim = CL.block, kp.sharedMemBytes = 0, kp.kernelParams = CL.argv, kp.extra = nullptr;
            CK(cudaGraphAddKernelNode(&n_sim, g, deps, 2, &kp));
        }
        tail = n_sim;
        ++launches;
    }
    int stage = 1;
    for (int l = L - 2; l >= 0; --l, ++stage)
    {
        const LevelGeom &G = c->geom[l];
        Cand *out = s.d_buf[stage & 1];
        uint32_t out_cap = s.cap;
        uint32_t *out_cnt = s.d_counts + stage;
        if (c->n_active > 0)
        {
            const uint8_t *lm = s.lv[l].d_lm;
            const TemplLevel *tl = c->tl[l].d_tl;
            const Feat *ff = c->tl[l].d_feats;
            const Cand *cin = in;
            const uint32_t *cin_cnt = in_cnt;
            uint16_t *nodump = nullptr;
            int raw_center = 0;
            unsigned long long *ev = reinterpret_cast<unsigned long long *>(s.d_counts + 10);
            cudaGraphNode_t n_ref;
            if ((rc = addKernel(g, &n_ref, {tail, spread_node[l]}, refineFunc(G.tiled), dbgGrid(6, dim3(refineGrid(c, G.tiled))), dim3(K6_THREADS), lm, tl, ff, cin, cin_cnt, in_cap,
                                out, out_cnt, out_cap, G.w, G.h, G.T, threshold, nodump, raw_center, ev)))
                return rc;
            tail = n_ref;
            ++launches;
        }
        in = out, in_cap = out_cap, in_cnt = out_cnt;
    }
    s.d_final = in;
    s.d_final_count = in_cnt;
    s.final_cap = in_cap;
    s.graph_final = in, s.graph_final_count = in_cnt, s.graph_final_cap = in_cap;
    if (with_fetch)
    {
        // tail depends on every other node transitively except possibly finer-level spreads when n_active == 0
        std::vector<cudaGraphNode_t> deps{tail};
        for (int l = 0; l < L; ++l)
            deps.push_back(spread_node[l]);
        cudaGraphNode_t n_c, n_r;
        CK(cudaGraphAddMemcpyNode1D(&n_c, g, deps.data(), deps.size(), s.h_counts, s.d_counts, COUNT_WORDS * sizeof(uint32_t), cudaMemcpyDeviceToHost));
        const uint32_t head = std::min(HEAD_RECS, s.final_cap);
        CK(cudaGraphAddMemcpyNode1D(&n_r, g, deps.data(), deps.size(), s.h_recs, s.d_final, (size_t)head * sizeof(Cand), cudaMemcpyDeviceToHost));
    }
    CK(cudaGraphInstantiate(&s.gexec, g, 0));
    s.graph_epoch = c->epoch;
    s.graph_threshold = threshold;
    s.graph_fetch = with_fetch;
    s.graph_src0 = src0;
    s.graph_launches = launches;
    return 0;
}

bool graphEligible(const sbm_ctx *c, const uint8_t *d_mask) { return !d_mask && !c->debug && !c->timing; }

// one frame through the slot's graph on stream st (results in the slot's own buffers)
int launchFrameGraph(sbm_ctx *c, Slot &s, const uint8_t *d_img, float threshold, bool with_fetch, cudaStream_t st)
{
    int rc;
    if (!s.gexec || s.graph_epoch != c->epoch || s.graph_threshold != threshold || s.graph_fetch != with_fetch)
    {
        CK(cudaStreamSynchronize(s.stream));
        if ((rc = buildGraph(c, s, d_img, threshold, with_fetch)))
            return rc;
    }
    else if (s.graph_src0 != d_img)
    {
        // patch the two kernels that read the level-0 frame
        const uint8_t *src = d_img, *null_mask = nullptr;
        const float weak_sq = c->weak * c->weak;
        {
            const LevelGeom &G = c->geom[0];
            int w = G.w, h = G.h, shift = 0, mask_w = c->w;
            uint8_t *q = s.lv[0].d_quant;
            float *null_f = nullptr;
            CUtensorMap map;
            const int th = orientTileH(w, h, c->sms);
            int use_tma = orientMap(&map, src, w, h, c->c, th, c->sms);
            void *argv[] = {&src, &w, &h, &null_mask, &shift, &mask_w, (void *)&weak_sq, &q, &null_f, &null_f, &map, &use_tma};
            cudaKernelNodeParams p = {};
            // (the grid follows the frame: a frame whose address rules the tensor copy out gets one CTA per tile)
            p.func = s.f_orient0, p.gridDim = dbgGrid(0, orientGrid(w, h, th, use_tma)), p.blockDim = dim3(K1_THREADS), p.kernelParams = argv;
            CK(cudaGraphExecKernelNodeSetParams(s.gexec, s.n_orient0, &p));
        }
        if (s.n_pyr1)
        {
            const LevelGeom &P = c->geom[0], &G = c->geom[1];
            int pw = P.w, ph = P.h, gw = G.w, gh = G.h;
            uint8_t *dst = s.lv[1].d_src;
            void *argv[] = {&src, &pw, &ph, &dst, &gw, &gh};
            cudaKernelNodeParams p = {};
            p.func = s.f_pyr1, p.gridDim = s.g_pyr1, p.blockDim = dim3(K1B_THREADS), p.kernelParams = argv;
            CK(cudaGraphExecKernelNodeSetParams(s.gexec, s.n_pyr1, &p));
        }
        s.graph_src0 = d_img;
    }
    CK(cudaGraphLaunch(s.gexec, st));
    // where THIS graph leaves the frame's records (an eager frame in between -- mask, debug, timing -- may have pointed
    // the slot at the caller's own buffers)
    s.d_final = s.graph_final, s.d_final_count = s.graph_final_count, s.final_cap = s.graph_final_cap;
    s.launches = s.graph_launches;
    s.last_src0 = d_img;
    s.have_frame = true;
    return 0;
}

int enqueueFetch(Slot &s, cudaStream_t st);

// one frame on slot s: through the graph when nothing frame-specific but the source pointer varies, else eagerly
int runFrame(sbm_ctx *c, Slot &s, const uint8_t *d_img, const uint8_t *d_mask, float threshold, cudaStream_t st, bool with_fetch)
{
    int rc;
    if (graphEligible(c, d_mask))
        return launchFrameGraph(c, s, d_img, threshold, with_fetch, st);
    if ((rc = enqueueFrame(c, s, d_img, d_mask, threshold, st, nullptr, 0, nullptr)))
        return rc;
    return with_fetch ? enqueueFetch(s, st) : 0;
}

int exportResults(sbm_ctx *c, Slot &s, cudaStream_t st, void *d_records, uint32_t cap, void *d_count)
{
    if (!d_records && !d_count)
        return 0;
    if (!d_records || !d_count)
        return fail(SBM_ERR_BAD_ARG, "d_records and d_count must be given together");
    k_export<<<8, 256, 0, st>>>(s.d_final, s.d_final_count, s.final_cap, (Cand *)d_records, cap, (uint32_t *)d_count, s.d_counts, c->levels,
                                s.cap);
    ++s.launches;
    CK(cudaGetLastError());
    return 0;
}

int growCap(sbm_ctx *c, Slot &s, uint32_t need)
{
    uint32_t nc = s.cap;
    while (nc < need)
        nc *= 2;
    CK(cudaStreamSynchronize(s.stream));
    freeDev(s.d_buf[0]);
    freeDev(s.d_buf[1]);
    s.cap = nc;
    s.graph_epoch = 0;
    return ensureSlot(c, s);
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

// queue the D2H of the counters and the head of the record buffer behind the frame's kernels
int enqueueFetch(Slot &s, cudaStream_t st)
{
    const uint32_t head = std::min(HEAD_RECS, s.final_cap);
    CK(cudaMemcpyAsync(s.h_counts, s.d_counts, COUNT_WORDS * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(s.h_recs, s.d_final, (size_t)head * sizeof(Cand), cudaMemcpyDeviceToHost, st));
    return 0;
}

// wait for the slot's frame; returns SBM_ERR_CAPACITY (without error text) when a stage overflowed
int finishFetch(sbm_ctx *c, Slot &s, cudaStream_t st, uint32_t *n_out)
{
    CK(cudaStreamSynchronize(st));
    const uint32_t n = s.h_counts[s.d_final_count - s.d_counts];
    *n_out = n;
    for (int k = 0; k < c->levels; ++k)
        if (s.h_counts[k] > s.cap)
            return SBM_ERR_CAPACITY;
    uint32_t head = std::min(HEAD_RECS, s.final_cap);
    if (n > head)
    {
        if (n > s.h_recs_cap)
        {
            cudaFreeHost(s.h_recs);
            s.h_recs = nullptr;
            s.h_recs_cap = std::max(n, 2 * s.h_recs_cap);
            CK(cudaMallocHost(&s.h_recs, (size_t)s.h_recs_cap * sizeof(Cand)));
            head = 0;
            s.graph_epoch = 0; // the slot's graph holds the old host pointer in its fetch node: rebuild before the next launch
        }
        CK(cudaMemcpyAsync(s.h_recs + head, s.d_final + head, (size_t)(n - head) * sizeof(Cand), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    return 0;
}

int uploadFrame(Slot &s, const uint8_t *img, int w, int h, int ch, size_t stride, const uint8_t *mask, const uint8_t **d_mask_out)
{
    cudaStream_t st = s.stream;
    const size_t row = (size_t)w * ch, bytes = row * h;
    if (s.frame_alloc < bytes)
    {
        CK(cudaStreamSynchronize(st));
        freeDev(s.d_frame);
        CK(cudaMalloc(&s.d_frame, bytes));
        s.frame_alloc = bytes;
    }
    if (stride == 0 || stride == row)
        CK(cudaMemcpyAsync(s.d_frame, img, bytes, cudaMemcpyHostToDevice, st));
    else
        CK(cudaMemcpy2DAsync(s.d_frame, row, img, stride, row, h, cudaMemcpyHostToDevice, st));
    *d_mask_out = nullptr;
    if (mask)
    {
        const size_t mb = (size_t)w * h;
        if (s.mask_alloc < mb)
        {
            CK(cudaStreamSynchronize(st));
            freeDev(s.d_mask);
            CK(cudaMalloc(&s.d_mask, mb));
            s.mask_alloc = mb;
        }
        CK(cudaMemcpyAsync(s.d_mask, mask, mb, cudaMemcpyHostToDevice, st));
        *d_mask_out = s.d_mask;
    }
    s.last_src0 = s.d_frame;
    return 0;
}

void recordStats(sbm_ctx *c, const Slot &s, uint32_t n)
{
    c->stats[0] = c->coarse_evals;
    c->stats[1] = s.h_counts[0];
    memcpy(&c->stats[2], &s.h_counts[10], sizeof(uint64_t));
    c->stats[3] = n;
    c->stats[4] = (uint64_t)s.launches;
}

int matchHost(sbm_ctx *c, const uint8_t *img, int w, int h, int ch, size_t stride, const uint8_t *mask, float threshold,
              const int *class_idx, int n_classes, bool finalize, sbm_match_rec *out, int cap, int *n_out)
{
    if (!c || !img || !n_out)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    Slot &s = c->slot[0];
    int rc;
    if ((rc = prepare(c, s, w, h, ch, class_idx, n_classes, threshold)))
        return rc;
    const uint8_t *dm = nullptr;
    if ((rc = uploadFrame(s, img, w, h, ch, stride, mask, &dm)))
        return rc;
    for (int attempt = 0; attempt < 8; ++attempt)
    {
        if ((rc = runFrame(c, s, s.d_frame, dm, threshold, s.stream, true)))
            return rc;
        uint32_t n = 0;
        rc = finishFetch(c, s, s.stream, &n);
        if (rc == SBM_ERR_CAPACITY)
        {
            uint32_t need = 0;
            for (int k = 0; k < c->levels; ++k)
                need = std::max(need, s.h_counts[k]);
            if ((rc = growCap(c, s, need)))
                return rc;
            continue;
        }
        if (rc)
            return rc;
        recordStats(c, s, n);
        return orderRecords(c, s.h_recs, n, finalize, out, cap, n_out);
    }
    return fail(SBM_ERR_CAPACITY, "candidate buffers keep overflowing");
}

// finish the frame in flight on slot s (re-running it with larger buffers if a stage overflowed) and append its
// ordered records to out[*used ..)
int finishBatchFrame(sbm_ctx *c, Slot &s, const uint8_t *dm, float threshold, bool finalize, sbm_match_rec *out, int cap, int *used,
                     bool *overflowed)
{
    int rc;
    for (int attempt = 0; attempt < 8; ++attempt)
    {
        uint32_t n = 0;
        rc = finishFetch(c, s, s.stream, &n);
        if (rc == SBM_ERR_CAPACITY)
        {
            uint32_t need = 0;
            for (int k = 0; k < c->levels; ++k)
                need = std::max(need, s.h_counts[k]);
            if ((rc = growCap(c, s, need)) || (rc = runFrame(c, s, s.d_frame, dm, threshold, s.stream, true)))
                return rc;
            continue;
        }
        if (rc)
            return rc;
        recordStats(c, s, n);
        int got = 0;
        const int room = *overflowed ? 0 : std::max(0, cap - *used);
        rc = orderRecords(c, s.h_recs, n, finalize, out ? out + *used : nullptr, room, &got);
        if (rc == SBM_ERR_CAPACITY)
            *overflowed = true, rc = 0;
        *used += got;
        return rc;
    }
    return fail(SBM_ERR_CAPACITY, "candidate buffers keep overflowing");
}
// Template sharding: the merge step after the allgather.  Pure host code, no context.  The reference appends
// candidates class by class in the caller's order (repeats included, line2Dup.cpp:1124-1140), within a class by
// template_id ascending and coarse scan order (:1170, :1200-1216); so the ranks' records are re-labelled with GLOBAL
// template ids, ordered that way and fed to the same std::sort + std::unique as Detector::match (:1143-1145).
// class_order == NULL: every class slot once, ascending (the "all classes" request).
int mergeRecords(const sbm_device_rec *recs, const int *n_per_rank, int n_ranks, const int *class_sizes, int n_classes, int mode,
                 const int *class_order, int n_order, bool finalize, std::vector<sbm_match_rec> &m)
{
    struct G
    {
        int32_t tid;
        uint32_t key;
        const sbm_device_rec *r;
    };
    std::vector<std::vector<G>> per_class(n_classes);
    size_t at = 0;
    for (int rk = 0; rk < n_ranks; ++rk)
    {
        const int *cs = class_sizes + (size_t)rk * n_classes;
        std::vector<uint32_t> begin(n_classes + 1, 0);
        for (int k = 0; k < n_classes; ++k)
        {
            if (cs[k] < 0)
                return fail(SBM_ERR_BAD_ARG, "negative class size");
            begin[k + 1] = begin[k] + (uint32_t)cs[k];
        }
        if (n_per_rank[rk] < 0 || (n_per_rank[rk] > 0 && !recs))
            return fail(SBM_ERR_BAD_ARG, "bad record count of rank %d", rk);
        for (int i = 0; i < n_per_rank[rk]; ++i, ++at)
        {
            const sbm_device_rec &r = recs[at];
            if (r.template_gid >= begin[n_classes])
                return fail(SBM_ERR_BAD_ARG, "rank %d record %d names local template %u of %u", rk, i, r.template_gid, begin[n_classes]);
            const int k = (int)(std::upper_bound(begin.begin(), begin.end(), r.template_gid) - begin.begin()) - 1;
            const int local = (int)(r.template_gid - begin[k]);
            int tid;
            if (mode == SBM_SHARD_INTERLEAVED)
                tid = rk + local * n_ranks; // rank rk owns ids rk, rk + G, rk + 2G, ...
            else
            {
                tid = local; // rank rk owns a contiguous block after the blocks of the lower ranks
                for (int q = 0; q < rk; ++q)
                    tid += class_sizes[(size_t)q * n_classes + k];
            }
            per_class[k].push_back(G{tid, r.order_key, &r});
        }
    }
    for (auto &v : per_class)
        std::sort(v.begin(), v.end(), [](const G &a, const G &b) { return a.tid != b.tid ? a.tid < b.tid : a.key < b.key; });
    m.clear();
    m.reserve(at);
    std::vector<int> all;
    if (!class_order || n_order <= 0)
    {
        for (int k = 0; k < n_classes; ++k)
            all.push_back(k);
        class_order = all.data(), n_order = n_classes;
    }
    for (int i = 0; i < n_order; ++i)
    {
        const int k = class_order[i];
        if (k < 0 || k >= n_classes) // unknown ids are skipped (:1136-1138)
            continue;
        for (const G &e : per_class[k])
            m.push_back(sbm_match_rec{e.r->x, e.r->y, e.r->similarity, k, e.tid});
    }
    if (finalize)
    {
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
    return 0;
}

// ------------------------------------------------------------------------------------------------
// Multi-GPU group: result blocks and the one exchange.
// A frame's result travels as one block: {uint32 count, uint32 overflow, uint32 0, uint32 0} + cap records of 20 bytes.
// Every rank fills one block per frame (k_export writes the header and the records), then ONE ncclAllGather moves all
// blocks of all ranks -- counts and records in the same buffer, no second collective.
// ------------------------------------------------------------------------------------------------
void destroyGroup(sbm_ctx *c)
{
    sbm_ctx::Group &G = c->group;
    if (!G.on)
        return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    if (G.comm)
        if (NcclApi *N = NcclApi::get())
            N->CommDestroy(G.comm);
    freeDev(G.d_send);
    freeDev(G.d_recv);
    if (G.h_recv)
        cudaFreeHost(G.h_recv);
    for (cudaEvent_t e : G.chunk_done)
        cudaEventDestroy(e);
    if (G.xstream)
        cudaStreamDestroy(G.xstream);
    if (G.mode != SBM_GROUP_FRAMES) // the context holds a slice of every class: meaningless without the group
    {
        c->classes.clear();
        c->templates_dirty = true;
    }
    G = sbm_ctx::Group();
}

constexpr size_t GROUP_HEAD = 16;
size_t groupBlockBytes(uint32_t cap) { return GROUP_HEAD + (size_t)cap * sizeof(Cand); }

int ncclFail(NcclApi *N, int r, const char *what) { return fail(SBM_ERR_CUDA, "%s failed: %s", what, N->GetErrorString(r)); }

int ensureGroupBuffers(sbm_ctx *c, size_t send_bytes, bool own_recv)
{
    sbm_ctx::Group &G = c->group;
    if (G.send_alloc < send_bytes)
    {
        CK(cudaDeviceSynchronize());
        freeDev(G.d_send);
        CK(cudaMalloc(&G.d_send, send_bytes));
        G.send_alloc = send_bytes;
    }
    const size_t recv_bytes = send_bytes * G.n;
    if (own_recv && G.recv_alloc < recv_bytes)
    {
        CK(cudaDeviceSynchronize());
        freeDev(G.d_recv);
        CK(cudaMalloc(&G.d_recv, recv_bytes));
        G.recv_alloc = recv_bytes;
    }
    if (own_recv && G.h_alloc < recv_bytes)
    {
        if (G.h_recv)
            cudaFreeHost(G.h_recv);
        G.h_recv = nullptr;
        CK(cudaMallocHost(&G.h_recv, recv_bytes));
        G.h_alloc = recv_bytes;
    }
    return 0;
}

// Enqueue n_frames frames (host pointers `imgs`, or resident at d_imgs + i * frame_stride) over the frame slots with their
// results exported into the send blocks, join the slot streams on `user`, and allgather the blocks into d_gathered on `user`.
// A caller that feeds a batch in chunks passes the chunk's position (`first` of `total` frames): the chunk's send blocks
// and slots continue where the previous chunk stopped, so chunks in flight never share memory; wait_user = false when
// nothing queued on `user` has to finish before the frames may start.
int groupEnqueue(sbm_ctx *c, const uint8_t *const *imgs, const uint8_t *const *masks, const void *d_imgs, size_t frame_stride, int n_frames,
                 int w, int h, int ch, size_t stride, float threshold, const int *class_idx, int n_classes, uint32_t cap, uint8_t *d_gathered,
                 cudaStream_t user, int first = 0, int total = -1, bool wait_user = true)
{
    sbm_ctx::Group &G = c->group;
    NcclApi *N = NcclApi::get();
    if (!G.on || !N)
        return fail(SBM_ERR_STATE, "no group: call sbm_group_create first");
    int rc;
    const int NS = frameSlots();
    if (total < 0)
        total = n_frames;
    const int ns = std::min(total, NS); // (a chunk of a longer batch may wrap around the slots)
    for (int k = 0; k < ns; ++k)
        if ((rc = prepare(c, c->slot[k], w, h, ch, class_idx, n_classes, threshold)))
            return rc;
    const size_t blk = groupBlockBytes(cap);
    if ((rc = ensureGroupBuffers(c, blk * total, d_gathered == nullptr)))
        return rc;
    if (!d_gathered)
        d_gathered = G.d_recv;
    Slot &s0 = c->slot[0];
    if (ns > 0 && wait_user)
    {
        CK(cudaEventRecord(s0.ev[0], user));
        for (int k = 0; k < ns; ++k)
            CK(cudaStreamWaitEvent(c->slot[k].stream, s0.ev[0], 0));
    }
    const bool timing = c->timing, debug = c->debug;
    c->timing = false, c->debug = false;
    rc = 0;
    uint8_t *const send = G.d_send + (size_t)first * blk;
    for (int i = 0; i < n_frames && !rc; ++i)
    {
        Slot &s = c->slot[(first + i) % NS];
        uint8_t *block = send + (size_t)i * blk;
        const uint8_t *img = nullptr, *dm = nullptr;
        if (imgs)
        {
            if ((rc = uploadFrame(s, imgs[i], w, h, ch, stride, masks ? masks[i] : nullptr, &dm)))
                break;
            img = s.d_frame;
        }
        else
            img = (const uint8_t *)d_imgs + (size_t)i * frame_stride;
        if (dm) // masks take the eager path, straight into the block
            rc = enqueueFrame(c, s, img, dm, threshold, s.stream, (Cand *)(block + GROUP_HEAD), cap, (uint32_t *)block);
        else if (!(rc = launchFrameGraph(c, s, img, threshold, false, s.stream)))
            rc = exportResults(c, s, s.stream, block + GROUP_HEAD, cap, block);
    }
    c->timing = timing, c->debug = debug;
    if (rc)
        return rc;
    for (int k = 0; k < std::min(ns, n_frames); ++k)
    {
        Slot &s = c->slot[(first + k) % NS];
        CK(cudaEventRecord(s.ev[1], s.stream));
        CK(cudaStreamWaitEvent(user, s.ev[1], 0));
    }
    if (n_frames > 0)
    {
        const int r = N->AllGather(send, d_gathered, blk * n_frames, NcclApi::kUint8, G.comm, user);
        if (r)
            return ncclFail(N, r, "ncclAllGather");
    }
    return 0;
}

// gathered host blocks -> final lists.  Template modes: frame f's list = merge of the n ranks' blocks of frame f (n_frames
// lists).  Frame mode: every (rank, frame) block is a complete list of its own frame (n * n_frames lists, rank-major).
// *worst = largest record count any block reported, *overflow = a stage or block overflowed somewhere.
int groupFinalize(sbm_ctx *c, const uint8_t *h, int n_frames, uint32_t cap, const int *class_idx, int n_classes, bool finalize,
                  sbm_match_rec *out, int out_cap, int *frame_begin, uint32_t *worst, bool *overflow)
{
    sbm_ctx::Group &G = c->group;
    const size_t blk = groupBlockBytes(cap);
    *worst = 0, *overflow = false;
    for (int r = 0; r < G.n; ++r)
        for (int f = 0; f < n_frames; ++f)
        {
            const uint32_t *hd = (const uint32_t *)(h + ((size_t)r * n_frames + f) * blk);
            *worst = std::max(*worst, hd[0]);
            if (hd[1] || hd[0] > cap)
                *overflow = true;
        }
    if (*overflow)
        return 0;
    int used = 0;
    bool full = false;
    frame_begin[0] = 0;
    std::vector<sbm_match_rec> m;
    if (G.mode == SBM_GROUP_FRAMES)
    {
        for (int r = 0; r < G.n; ++r)
            for (int f = 0; f < n_frames; ++f)
            {
                const uint8_t *b = h + ((size_t)r * n_frames + f) * blk;
                const uint32_t n = *(const uint32_t *)b;
                int got = 0;
                const int room = full ? 0 : std::max(0, out_cap - used);
                int rc = orderRecords(c, (const Cand *)(b + GROUP_HEAD), n, finalize, out ? out + used : nullptr, room, &got);
                if (rc == SBM_ERR_CAPACITY)
                    full = true, rc = 0;
                if (rc)
                    return rc;
                used += got;
                frame_begin[r * n_frames + f + 1] = used;
            }
    }
    else
    {
        const int n_slots = (int)c->classes.size();
        std::vector<int> sizes((size_t)G.n * std::max(1, n_slots), 0);
        for (int r = 0; r < G.n; ++r)
            for (int k = 0; k < n_slots; ++k)
                sizes[(size_t)r * n_slots + k] = shardCount((size_t)k < G.global_n.size() ? G.global_n[k] : 0, r, G.n, G.mode);
        // the caller's class order against the GLOBAL class sizes (a rank whose slice of a class is empty must still
        // produce the same merged list as the others)
        std::vector<int> order;
        if (!class_idx || n_classes <= 0)
        {
            for (int k = 0; k < n_slots; ++k)
                if ((size_t)k < G.global_n.size() && G.global_n[k] > 0)
                    order.push_back(k);
        }
        else
            for (int i = 0; i < n_classes; ++i)
                if (class_idx[i] >= 0 && (size_t)class_idx[i] < G.global_n.size() && G.global_n[class_idx[i]] > 0)
                    order.push_back(class_idx[i]);
        std::vector<sbm_device_rec> recs;
        std::vector<int> per_rank(G.n);
        for (int f = 0; f < n_frames; ++f)
        {
            recs.clear();
            for (int r = 0; r < G.n; ++r)
            {
                const uint8_t *b = h + ((size_t)r * n_frames + f) * blk;
                const uint32_t n = *(const uint32_t *)b;
                per_rank[r] = (int)n;
                const sbm_device_rec *p = (const sbm_device_rec *)(b + GROUP_HEAD);
                recs.insert(recs.end(), p, p + n);
            }
            int rc = 0;
            m.clear();
            if (n_slots > 0 && !order.empty())
                rc = mergeRecords(recs.data(), per_rank.data(), G.n, sizes.data(), n_slots, G.mode, order.data(), (int)order.size(), finalize, m);
            if (rc)
                return rc;
            if (!full && used + (int)m.size() <= out_cap)
            {
                if (!m.empty() && out)
                    memcpy(out + used, m.data(), m.size() * sizeof(sbm_match_rec));
            }
            else
                full = true;
            used += (int)m.size();
            frame_begin[f + 1] = used;
        }
    }
    if (full)
        return fail(SBM_ERR_CAPACITY, "output holds %d matches, %d needed", out_cap, used);
    return 0;
}

// frames per chunk of the host-side group calls (SBM_GROUP_CHUNK; at least one round of the frame slots)
int groupChunk()
{
    static const int v = [] {
        const char *e = getenv("SBM_GROUP_CHUNK");
        return std::max(frameSlots(), e ? atoi(e) : 32);
    }();
    return v;
}

// host frames -> final host lists through the group (see sbm_group_match_batch).  The batch goes through in chunks:
// chunk k's frames upload and run on the frame slots, its blocks are exchanged and copied to the host on the group's
// own stream, and while chunk k+1 is on its way the host orders chunk k's lists (with n ranks that is n times the
// host work of a single GPU, and it would otherwise sit behind the last copy).
int groupMatchHost(sbm_ctx *c, const uint8_t *const *imgs, const uint8_t *const *masks, int n_frames, int w, int h, int ch, size_t stride,
                   float threshold, const int *class_idx, int n_classes, bool finalize, sbm_match_rec *out, int cap, int *frame_begin)
{
    sbm_ctx::Group &G = c->group;
    int rc;
    const int CH = groupChunk(), n_chunks = (n_frames + CH - 1) / CH;
    const int per_frame = G.mode == SBM_GROUP_FRAMES ? G.n : 1, lists = n_frames * per_frame;
    if (!G.xstream)
        CK(cudaStreamCreateWithFlags(&G.xstream, cudaStreamNonBlocking));
    while ((int)G.chunk_done.size() < n_chunks)
    {
        cudaEvent_t e;
        CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        G.chunk_done.push_back(e);
    }
    std::vector<size_t> pos(lists);
    std::vector<int> cnt(lists), fb;
    for (int attempt = 0; attempt < 8; ++attempt)
    {
        const size_t blk = groupBlockBytes(G.cap);
        if ((rc = ensureGroupBuffers(c, blk * n_frames, true)))
            return rc;
        uint32_t worst = 0;
        bool overflow = false;
        size_t pool_used = 0;
        // the lists of chunk k: wait for its copy, order them into the pool, remember where each global list lies
        auto finish_chunk = [&](int k) -> int {
            const int c0 = k * CH, cn = std::min(CH, n_frames - c0);
            CK(cudaEventSynchronize(G.chunk_done[k]));
            const size_t need = (size_t)G.n * cn * G.cap;
            if (G.pool.size() < pool_used + need)
                G.pool.resize(std::max(pool_used + need, 2 * G.pool.size()));
            fb.assign((size_t)cn * per_frame + 1, 0);
            uint32_t wk = 0;
            bool ov = false;
            const int r = groupFinalize(c, G.h_recv + (size_t)G.n * blk * c0, cn, G.cap, class_idx, n_classes, finalize, G.pool.data() + pool_used,
                                        (int)std::min<size_t>(need, 0x7fffffff), fb.data(), &wk, &ov);
            if (r)
                return r;
            worst = std::max(worst, wk);
            if (ov)
            {
                overflow = true;
                return 0;
            }
            for (int j = 0; j < cn * per_frame; ++j)
            {
                // frame mode: chunk lists are rank-major within the chunk; the batch's are rank-major over all frames
                const int g = G.mode == SBM_GROUP_FRAMES ? (j / cn) * n_frames + c0 + j % cn : c0 + j;
                pos[g] = pool_used + fb[j], cnt[g] = fb[j + 1] - fb[j];
            }
            pool_used += fb[(size_t)cn * per_frame];
            return 0;
        };
        for (int k = 0; k < n_chunks; ++k)
        {
            const int c0 = k * CH, cn = std::min(CH, n_frames - c0);
            uint8_t *d_dst = G.d_recv + (size_t)G.n * blk * c0;
            if ((rc = groupEnqueue(c, imgs + c0, masks ? masks + c0 : nullptr, nullptr, 0, cn, w, h, ch, stride, threshold, class_idx, n_classes, G.cap,
                                   d_dst, G.xstream, c0, n_frames, false)))
                return rc;
            CK(cudaMemcpyAsync(G.h_recv + (size_t)G.n * blk * c0, d_dst, (size_t)G.n * blk * cn, cudaMemcpyDeviceToHost, G.xstream));
            CK(cudaEventRecord(G.chunk_done[k], G.xstream));
            if (k > 0 && (rc = finish_chunk(k - 1)))
                return rc;
        }
        if (n_chunks > 0 && (rc = finish_chunk(n_chunks - 1)))
            return rc;
        if (!overflow)
        {
            int used = 0;
            bool full = false;
            frame_begin[0] = 0;
            for (int g = 0; g < lists; ++g)
            {
                if (!full && used + cnt[g] <= cap)
                {
                    if (cnt[g] && out)
                        memcpy(out + used, G.pool.data() + pos[g], (size_t)cnt[g] * sizeof(sbm_match_rec));
                }
                else
                    full = true;
                used += cnt[g];
                frame_begin[g + 1] = used;
            }
            if (full)
                return fail(SBM_ERR_CAPACITY, "output holds %d matches, %d needed", cap, used);
            return 0;
        }
        // every rank sees the same gathered headers, so every rank takes the same decision: larger blocks, larger stage
        // buffers, and the whole batch again (the collective stays matched across the ranks)
        uint32_t nc = G.cap;
        while (nc < worst || nc == G.cap)
            nc *= 2;
        G.cap = nc;
        CK(cudaDeviceSynchronize());
        for (Slot &s : c->slot)
            if (s.cap < 4 * nc || s.cap < (1u << 18))
            {
                freeDev(s.d_buf[0]);
                freeDev(s.d_buf[1]);
                s.cap = std::max(4 * nc, 1u << 18);
                s.graph_epoch = 0;
            }
    }
    return fail(SBM_ERR_CAPACITY, "group result blocks keep overflowing");
}
} // namespace

extern "C"
{
const char *sbm_last_error(void) { return g_err.c_str(); }

#ifdef SBM_TRACE
// -DSBM_TRACE builds only (tools/timeline.py): copies the CTA records out (dst: cap records of 32 bytes) and resets the count
int sbm_debug_timeline(void *dst, unsigned int cap, unsigned int *n)
{
    unsigned int have = 0;
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpyFromSymbol(&have, sbm::g_timeline_n, sizeof have));
    const unsigned int take = std::min(std::min(have, cap), sbm::TIMELINE_CAP);
    if (dst && take)
        CK(cudaMemcpyFromSymbol(dst, sbm::g_timeline, (size_t)take * sizeof(sbm::TimelineRec)));
    const unsigned int zero = 0;
    CK(cudaMemcpyToSymbol(sbm::g_timeline_n, &zero, sizeof zero));
    if (n)
        *n = take;
    return 0;
}
#endif

const char *sbm_build_info(void)
{
#define SBM_STR2(x) #x
#define SBM_STR(x) SBM_STR2(x)
    return "shape_based_matching_b200 csrc: nvcc " SBM_STR(__CUDACC_VER_MAJOR__) "." SBM_STR(__CUDACC_VER_MINOR__)
           " -gencode arch=compute_100a,code=sm_100a; kernels k_orient_mma k_orient k_pyrdown k_spread_linearize k_similarity_bits k_similarity_nib k_similarity k_refine_tiled k_refine";
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
    if (cudaDeviceGetAttribute(&c->sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || c->sms < 1)
        c->sms = 148;
    c->refine_grid[0] = computeRefineGrid(c->sms, false);
    c->refine_grid[1] = computeRefineGrid(c->sms, true);
    for (int k = 2; k <= K5B_SPLIT_MAX; ++k)
    {
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c->bits_split_ctas[0][k], k_similarity_bits<3, true>, K5B_THREADS * k, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c->bits_split_ctas[1][k], k_similarity_bits<6, true>, K5B_THREADS * k, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c->bits_split_ctas[2][k], k_similarity_bits<9, true>, K5B_THREADS * k, 0);
    }
    c->T.assign(T, T + levels);
    c->levels = levels;
    c->weak = weak, c->strong = strong, c->num_features = num_features;
    c->tl.resize(levels);
    *out = c;
    return 0;
}

int sbm_destroy(sbm_ctx *c)
{
    if (!c)
        return 0;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    destroyGroup(c);
    for (auto &s : c->slot)
        destroySlot(s);
    for (auto &l : c->tl)
    {
        freeDev(l.d_tl);
        freeDev(l.d_feats);
    }
    freeDev(c->d_tcoarse);
    freeDev(c->d_coff);
    freeDev(c->d_active);
    freeDev(c->d_sim_items);
    freeDev(c->d_sim_offs);
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
    sbm_ctx::Group &G = c->group;
    if (G.on && G.mode != SBM_GROUP_FRAMES)
    {
        // template sharding: the caller hands every rank the FULL class, this rank keeps its slice (ids rank, rank + n,
        // ... or its contiguous block) and remembers the size of the whole class for the merge after the allgather
        if ((size_t)class_idx >= G.global_n.size())
            G.global_n.resize(class_idx + 1, 0);
        G.global_n[class_idx] = n_templates;
        cs.n = 0;
        cs.desc.clear();
        cs.feats.clear();
        size_t fo = 0;
        for (int t = 0; t < n_templates; ++t)
        {
            size_t tf = 0;
            for (int l = 0; l < c->levels; ++l)
                tf += desc[(size_t)t * c->levels + l].num_features;
            if (shardOwns(t, n_templates, G.rank, G.n, G.mode))
            {
                cs.desc.insert(cs.desc.end(), desc + (size_t)t * c->levels, desc + (size_t)(t + 1) * c->levels);
                cs.feats.insert(cs.feats.end(), feats + fo, feats + fo + tf);
                ++cs.n;
            }
            fo += tf;
        }
    }
    else
    {
        cs.n = n_templates;
        cs.desc.assign(desc, desc + nd);
        cs.feats.assign(feats, feats + nf);
    }
    c->templates_dirty = true;
    return 0;
}

int sbm_clear_templates(sbm_ctx *c)
{
    if (!c)
        return fail(SBM_ERR_BAD_ARG, "null context");
    std::lock_guard<std::mutex> lock(c->mu);
    c->classes.clear();
    c->group.global_n.clear();
    c->templates_dirty = true;
    return 0;
}

int sbm_set_candidate_capacity(sbm_ctx *c, int cap)
{
    if (!c || cap < 1)
        return fail(SBM_ERR_BAD_ARG, "sbm_set_candidate_capacity: bad arguments");
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    CK(cudaDeviceSynchronize()); // frames in flight still write the old buffers
    const uint32_t want = std::max<uint32_t>(1024u, (uint32_t)cap);
    for (Slot &s : c->slot)
    {
        if (s.cap == want)
            continue;
        freeDev(s.d_buf[0]);
        freeDev(s.d_buf[1]);
        s.cap = want;
        s.graph_epoch = 0; // the capacities are baked into the slot's graph
    }
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

int sbm_match_batch(sbm_ctx *c, const uint8_t *const *imgs, const uint8_t *const *masks, int n_frames, int w, int h, int ch,
                    size_t stride, float threshold, const int *class_idx, int n_classes, sbm_match_rec *out, int cap,
                    int *frame_begin)
{
    if (!c || !imgs || n_frames < 0 || !frame_begin)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    for (int i = 0; i < n_frames; ++i)
        if (!imgs[i])
            return fail(SBM_ERR_BAD_ARG, "frame %d is null", i);
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    int rc;
    frame_begin[0] = 0;
    if (n_frames == 0)
        return 0;
    const int NS = frameSlots();
    for (int k = 0; k < std::min(n_frames, NS); ++k)
        if ((rc = prepare(c, c->slot[k], w, h, ch, class_idx, n_classes, threshold)))
            return rc;
    const uint8_t *dm[SBM_SLOTS] = {nullptr};
    int used = 0;
    bool overflowed = false;
    for (int i = 0; i < n_frames + NS; ++i)
    {
        const int done = i - NS; // the frame that occupied this slot before
        Slot &s = c->slot[i % NS];
        if (done >= 0 && done < n_frames)
        {
            if ((rc = finishBatchFrame(c, s, dm[i % NS], threshold, true, out, cap, &used, &overflowed)))
                return rc;
            frame_begin[done + 1] = used;
        }
        if (i < n_frames)
        {
            if ((rc = uploadFrame(s, imgs[i], w, h, ch, stride, masks ? masks[i] : nullptr, &dm[i % NS])) ||
                (rc = runFrame(c, s, s.d_frame, dm[i % NS], threshold, s.stream, true)))
                return rc;
        }
    }
    if (overflowed)
        return fail(SBM_ERR_CAPACITY, "output holds %d matches, %d needed", cap, used);
    return 0;
}

int sbm_match_device(sbm_ctx *c, const void *d_img, int w, int h, int ch, const void *d_mask, float threshold,
                     const int *class_idx, int n_classes, void *d_records, int cap, void *d_count, void *stream)
{
    if (!c || !d_img)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    if (!d_records != !d_count)
        return fail(SBM_ERR_BAD_ARG, "d_records and d_count must be given together");
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    Slot &s = c->slot[0];
    int rc;
    if ((rc = prepare(c, s, w, h, ch, class_idx, n_classes, threshold)))
        return rc;
    cudaStream_t st = stream ? (cudaStream_t)stream : s.stream;
    s.last_src0 = (const uint8_t *)d_img;
    if (!graphEligible(c, (const uint8_t *)d_mask))
        return enqueueFrame(c, s, (const uint8_t *)d_img, (const uint8_t *)d_mask, threshold, st, (Cand *)d_records,
                            d_records ? (uint32_t)cap : 0, (uint32_t *)d_count);
    if ((rc = launchFrameGraph(c, s, (const uint8_t *)d_img, threshold, false, st)))
        return rc;
    return exportResults(c, s, st, d_records, (uint32_t)cap, d_count);
}

int sbm_match_batch_device(sbm_ctx *c, const void *d_imgs, size_t frame_stride, int n_frames, int w, int h, int ch, float threshold,
                           const int *class_idx, int n_classes, void *d_records, int cap_per_frame, void *d_counts, void *stream)
{
    if (!c || !d_imgs || n_frames < 0 || !d_records || !d_counts || cap_per_frame <= 0)
        return fail(SBM_ERR_BAD_ARG, "bad argument");
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    int rc;
    const int NS = frameSlots();
    const int ns = std::min(n_frames, NS);
    for (int k = 0; k < ns; ++k)
        if ((rc = prepare(c, c->slot[k], w, h, ch, class_idx, n_classes, threshold)))
            return rc;
    cudaStream_t user = (cudaStream_t)stream;
    // fork: every slot stream starts after the work already queued on the caller's stream
    Slot &s0 = c->slot[0];
    if (ns > 0)
    {
        CK(cudaEventRecord(s0.ev[0], user));
        for (int k = 0; k < ns; ++k)
            CK(cudaStreamWaitEvent(c->slot[k].stream, s0.ev[0], 0));
    }
    const bool timing = c->timing, debug = c->debug;
    c->timing = false, c->debug = false; // stage events and debug taps belong to the single-frame path
    for (int i = 0; i < n_frames; ++i)
    {
        Slot &s = c->slot[i % NS];
        const uint8_t *img = (const uint8_t *)d_imgs + (size_t)i * frame_stride;
        if ((rc = launchFrameGraph(c, s, img, threshold, false, s.stream)) ||
            (rc = exportResults(c, s, s.stream, (Cand *)d_records + (size_t)i * cap_per_frame, (uint32_t)cap_per_frame,
                                (uint32_t *)d_counts + 2 * (size_t)i)))
        {
            c->timing = timing, c->debug = debug;
            return rc;
        }
    }
    c->timing = timing, c->debug = debug;
    // join: the caller's stream continues after the last frame of every slot
    for (int k = 0; k < ns; ++k)
    {
        CK(cudaEventRecord(c->slot[k].ev[1], c->slot[k].stream));
        CK(cudaStreamWaitEvent(user, c->slot[k].ev[1], 0));
    }
    return 0;
}

int sbm_device_results(sbm_ctx *c, void **d_records, void **d_count, int *cap)
{
    if (!c || !c->slot[0].have_frame)
        return fail(SBM_ERR_STATE, "no frame has been processed");
    const Slot &s = c->slot[0];
    if (d_records)
        *d_records = s.d_final;
    if (d_count)
        *d_count = s.d_final_count;
    if (cap)
        *cap = (int)s.final_cap;
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

// Template-sharded matching: the merge step after the allgather.  Pure host code, no context: the
// reference's candidate order is (class order, template_id ascending, coarse scan order) (line2Dup.cpp:1124-1140,
// :1170, :1200-1216), so the ranks' records are re-labelled with GLOBAL template ids, ordered that way and fed
// to the same std::sort + std::unique as Detector::match (:1143-1145).
int sbm_merge_shard_records(const sbm_device_rec *recs, const int *n_per_rank, int n_ranks, const int *class_sizes, int n_classes,
                            int mode, const int *class_order, int n_order, int finalize, sbm_match_rec *out, int cap, int *n_out)
{
    if (!n_per_rank || !class_sizes || !n_out || n_ranks < 1 || n_classes < 1 || (mode != SBM_SHARD_INTERLEAVED && mode != SBM_SHARD_CONTIGUOUS))
        return fail(SBM_ERR_BAD_ARG, "sbm_merge_shard_records: bad arguments");
    std::vector<sbm_match_rec> m;
    int rc = mergeRecords(recs, n_per_rank, n_ranks, class_sizes, n_classes, mode, class_order, n_order, finalize != 0, m);
    if (rc)
        return rc;
    *n_out = (int)m.size();
    if ((int)m.size() > cap)
        return fail(SBM_ERR_CAPACITY, "output holds %d matches, %zu needed", cap, m.size());
    if (!m.empty() && out)
        memcpy(out, m.data(), m.size() * sizeof(sbm_match_rec));
    return 0;
}

// ---------------------------------------------------------------- multi-GPU group

int sbm_group_unique_id(void *id128)
{
    if (!id128)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    NcclApi *N = NcclApi::get();
    if (!N)
        return fail(SBM_ERR_STATE, "%s", NcclApi::loadError().c_str());
    NcclApi::unique_id id;
    const int r = N->GetUniqueId(&id);
    if (r)
        return ncclFail(N, r, "ncclGetUniqueId");
    memcpy(id128, &id, sizeof(id));
    return 0;
}

int sbm_group_create(sbm_ctx *c, int n_ranks, int rank, const void *id128, int mode)
{
    if (!c || !id128 || n_ranks < 1 || rank < 0 || rank >= n_ranks ||
        (mode != SBM_GROUP_TEMPLATES && mode != SBM_GROUP_TEMPLATES_CONTIGUOUS && mode != SBM_GROUP_FRAMES))
        return fail(SBM_ERR_BAD_ARG, "sbm_group_create: bad arguments");
    NcclApi *N = NcclApi::get();
    if (!N)
        return fail(SBM_ERR_STATE, "%s", NcclApi::loadError().c_str());
    std::lock_guard<std::mutex> lock(c->mu);
    if (c->group.on)
        return fail(SBM_ERR_STATE, "the context already belongs to a group");
    for (const ClassSet &cs : c->classes)
        if (cs.n > 0 && mode != SBM_GROUP_FRAMES)
            return fail(SBM_ERR_STATE, "template sharding: create the group before installing templates (every rank is then handed the full classes)");
    CK(cudaSetDevice(c->device));
    NcclApi::unique_id id;
    memcpy(&id, id128, sizeof(id));
    NcclApi::comm_t comm = nullptr;
    const int r = N->CommInitRank(&comm, n_ranks, id, rank); // collective: returns once every rank has called it
    if (r)
        return ncclFail(N, r, "ncclCommInitRank");
    c->group.on = true;
    c->group.n = n_ranks, c->group.rank = rank, c->group.mode = mode;
    c->group.comm = comm;
    c->group.global_n.clear();
    return 0;
}

int sbm_group_destroy(sbm_ctx *c)
{
    if (!c)
        return 0;
    std::lock_guard<std::mutex> lock(c->mu);
    destroyGroup(c);
    return 0;
}

int sbm_group_info(sbm_ctx *c, int *n_ranks, int *rank, int *mode)
{
    if (!c)
        return fail(SBM_ERR_BAD_ARG, "null context");
    if (n_ranks)
        *n_ranks = c->group.on ? c->group.n : 1;
    if (rank)
        *rank = c->group.on ? c->group.rank : 0;
    if (mode)
        *mode = c->group.on ? c->group.mode : -1;
    return c->group.on ? 1 : 0;
}

int sbm_group_match_batch(sbm_ctx *c, const uint8_t *const *imgs, const uint8_t *const *masks, int n_frames, int w, int h, int ch, size_t stride,
                          float threshold, const int *class_idx, int n_classes, sbm_match_rec *out, int cap, int *frame_begin)
{
    if (!c || !imgs || n_frames < 0 || !frame_begin)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    for (int i = 0; i < n_frames; ++i)
        if (!imgs[i])
            return fail(SBM_ERR_BAD_ARG, "frame %d is null", i);
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    if (!c->group.on)
        return fail(SBM_ERR_STATE, "no group: call sbm_group_create first");
    return groupMatchHost(c, imgs, masks, n_frames, w, h, ch, stride, threshold, class_idx, n_classes, true, out, cap, frame_begin);
}

int sbm_group_match(sbm_ctx *c, const uint8_t *img, int w, int h, int ch, size_t stride, const uint8_t *mask, float threshold,
                    const int *class_idx, int n_classes, sbm_match_rec *out, int cap, int *n_out)
{
    if (!c || !img || !n_out)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    if (!c->group.on || c->group.mode == SBM_GROUP_FRAMES)
        return fail(SBM_ERR_STATE, "sbm_group_match needs a template-sharding group (frame sharding: sbm_group_match_batch)");
    int begin[2] = {0, 0};
    const uint8_t *imgs[1] = {img}, *masks[1] = {mask};
    const int rc = groupMatchHost(c, imgs, mask ? masks : nullptr, 1, w, h, ch, stride, threshold, class_idx, n_classes, true, out, cap, begin);
    *n_out = begin[1];
    return rc;
}

size_t sbm_group_block_bytes(int cap_per_frame) { return cap_per_frame > 0 ? groupBlockBytes((uint32_t)cap_per_frame) : 0; }

int sbm_group_match_batch_device(sbm_ctx *c, const void *d_imgs, size_t frame_stride, int n_frames, int w, int h, int ch, float threshold,
                                 const int *class_idx, int n_classes, void *d_gathered, int cap_per_frame, void *stream)
{
    if (!c || !d_imgs || n_frames < 0 || !d_gathered || cap_per_frame <= 0)
        return fail(SBM_ERR_BAD_ARG, "bad argument");
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    if (!c->group.on)
        return fail(SBM_ERR_STATE, "no group: call sbm_group_create first");
    return groupEnqueue(c, nullptr, nullptr, d_imgs, frame_stride, n_frames, w, h, ch, 0, threshold, class_idx, n_classes, (uint32_t)cap_per_frame,
                        (uint8_t *)d_gathered, (cudaStream_t)stream);
}

int sbm_group_finalize(sbm_ctx *c, const void *h_gathered, int n_frames, int cap_per_frame, const int *class_idx, int n_classes, int finalize,
                       sbm_match_rec *out, int cap, int *frame_begin)
{
    if (!c || !h_gathered || n_frames < 0 || cap_per_frame <= 0 || !frame_begin)
        return fail(SBM_ERR_BAD_ARG, "bad argument");
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    if (!c->group.on)
        return fail(SBM_ERR_STATE, "no group: call sbm_group_create first");
    int rc;
    if ((rc = ensureTemplates(c)) || (rc = ensureActive(c, class_idx, n_classes)))
        return rc;
    uint32_t worst = 0;
    bool overflow = false;
    if ((rc = groupFinalize(c, (const uint8_t *)h_gathered, n_frames, (uint32_t)cap_per_frame, class_idx, n_classes, finalize != 0, out, cap,
                            frame_begin, &worst, &overflow)))
        return rc;
    if (overflow)
    {
        const int lists = c->group.mode == SBM_GROUP_FRAMES ? c->group.n * n_frames : n_frames;
        frame_begin[lists] = (int)worst;
        return fail(SBM_ERR_CAPACITY, "a result block overflowed (largest count %u, cap_per_frame %d) or a stage exceeded the candidate capacity", worst,
                    cap_per_frame);
    }
    return 0;
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
    const Slot &s = c->slot[0];
    if (!s.have_frame)
        return fail(SBM_ERR_STATE, "no frame has been processed");
    CK(cudaSetDevice(c->device));
    const LevelGeom &G = c->geom[level];
    const LevelBuf &L = s.lv[level];
    size_t px = (size_t)G.w * G.h, need = 0;
    const void *src = nullptr;
    switch (kind)
    {
    case SBM_QUANTIZED: need = px, src = L.d_quant; break;
    case SBM_SPREAD: need = px, src = L.d_spread; break;
    case SBM_RESPONSE: need = 8 * px, src = L.d_resp; break;
    case SBM_LINEAR: need = 8 * px, src = L.d_lm; break;
    case SBM_PYRAMID_SRC: need = px * c->c, src = level == 0 ? s.last_src0 : L.d_src; break;
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
    if (kind == SBM_LINEAR && G.tiled)
    {
        uint8_t *d_bytes = nullptr;
        CK(cudaMalloc(&d_bytes, need));
        k_untile<<<148 * 4, 256>>>(L.d_lm, d_bytes, need, (uint32_t)G.W);
        CK(cudaMemcpy(dst, d_bytes, need, cudaMemcpyDeviceToHost));
        cudaFree(d_bytes);
        return 0;
    }
    if (kind == SBM_LINEAR && G.bits)
    {
        uint8_t *d_bytes = nullptr;
        CK(cudaMalloc(&d_bytes, need));
        k_unpack_bits<<<148 * 4, 256>>>(reinterpret_cast<const uint32_t *>(L.d_lm), d_bytes, need);
        CK(cudaMemcpy(dst, d_bytes, need, cudaMemcpyDeviceToHost));
        cudaFree(d_bytes);
        return 0;
    }
    if (kind == SBM_LINEAR && G.nib)
    {
        // the caller sees one byte per response whatever the device layout
        uint8_t *d_bytes = nullptr;
        CK(cudaMalloc(&d_bytes, need));
        k_unpack_nibbles<<<148 * 4, 256>>>(L.d_lm, d_bytes, need);
        CK(cudaMemcpy(dst, d_bytes, need, cudaMemcpyDeviceToHost));
        cudaFree(d_bytes);
        return 0;
    }
    CK(cudaMemcpy(dst, src, need, cudaMemcpyDeviceToHost));
    return 0;
}

int sbm_similarity_map(sbm_ctx *c, int class_idx, int template_id, uint16_t *dst, size_t bytes)
{
    if (!c || !dst)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lock(c->mu);
    Slot &s = c->slot[0];
    if (!s.have_frame)
        return fail(SBM_ERR_STATE, "no frame has been processed");
    if (class_idx < 0 || (size_t)class_idx >= c->classes.size() || template_id < 0 || template_id >= c->classes[class_idx].n)
        return fail(SBM_ERR_BAD_ARG, "unknown class/template"); // CV_Assert in getTemplates (line2Dup.cpp:1454-1456)
    CK(cudaSetDevice(c->device));
    const LevelGeom &CG = c->geom[c->levels - 1];
    const int HW = CG.W * CG.H;
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
    SimItem *d_item = nullptr;
    uint2 *d_offs = nullptr;
    uint32_t stride = 8;
    if (CG.bits) // the bit-plane kernel takes one record and one offset row per template instead of (active[], tinfo[])
    {
        const SimItem it = makeSimItem(c, gid, 1e30f);
        stride = std::max<uint32_t>(8, it.n_off);
        std::vector<uint2> offs(stride, make_uint2(0u, 0u));
        decodeOffsets(c, gid, offs.data());
        CK(cudaMalloc(&d_item, sizeof(SimItem)));
        CK(cudaMalloc(&d_offs, stride * sizeof(uint2)));
        CK(cudaMemcpy(d_item, &it, sizeof(SimItem), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_offs, offs.data(), stride * sizeof(uint2), cudaMemcpyHostToDevice));
    }
    CK(launchCoarse(c, s.lv[c->levels - 1].d_lm, d_item, d_offs, stride, d_one, 1, 1e30f, d_sink, d_cnt, 0, d_dump, s.stream));
    CK(cudaStreamSynchronize(s.stream));
    CK(cudaMemcpy(dst, d_dump, (size_t)HW * 2, cudaMemcpyDeviceToHost));
    cudaFree(d_one), cudaFree(d_cnt), cudaFree(d_sink), cudaFree(d_dump), cudaFree(d_item), cudaFree(d_offs);
    return 0;
}

int sbm_similarity_local(sbm_ctx *c, int class_idx, int template_id, int level, int x, int y, uint16_t *dst)
{
    if (!c || !dst || level < 0 || level >= c->levels)
        return fail(SBM_ERR_BAD_ARG, "bad argument");
    std::lock_guard<std::mutex> lock(c->mu);
    Slot &s = c->slot[0];
    if (!s.have_frame)
        return fail(SBM_ERR_STATE, "no frame has been processed");
    if (class_idx < 0 || (size_t)class_idx >= c->classes.size() || template_id < 0 || template_id >= c->classes[class_idx].n)
        return fail(SBM_ERR_BAD_ARG, "unknown class/template");
    CK(cudaSetDevice(c->device));
    const LevelGeom &G = c->geom[level];
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
    const uint8_t *lm = s.lv[level].d_lm;
    uint8_t *d_bytes = nullptr;
    if (G.nib || G.bits) // the coarsest level is never refined by the pipeline; the tap unpacks it first
    {
        const size_t bytes = G.lm_bytes + std::max((size_t)G.W * G.H, (size_t)16 * G.W + 16) + 1024;
        CK(cudaMalloc(&d_bytes, bytes));
        CK(cudaMemsetAsync(d_bytes, 0, bytes, s.stream));
        if (G.bits)
            k_unpack_bits<<<148 * 4, 256, 0, s.stream>>>(reinterpret_cast<const uint32_t *>(lm), d_bytes, G.lm_bytes);
        else
            k_unpack_nibbles<<<148 * 4, 256, 0, s.stream>>>(lm, d_bytes, G.lm_bytes);
        lm = d_bytes;
    }
    CK(launchRefine(G.tiled, 1, lm, c->tl[level].d_tl, c->tl[level].d_feats, d_in, d_cnt, 1, d_out, d_cnt + 1, 1, G.w, G.h, G.T, 1e30f, d_dump, 1,
                    d_ev, s.stream));
    CK(cudaStreamSynchronize(s.stream));
    cudaFree(d_bytes);
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
    stats[4] = (uint64_t)c->slot[0].launches;
    stats[5] = c->levels == 0 ? 0 : c->geom[c->levels - 1].bits ? 2 : c->geom[c->levels - 1].nib ? 1 : 0; // coarse layout
    stats[6] = c->levels > 1 && c->geom[0].tiled ? 1 : 0;
    return 0;
}

int sbm_last_timing(sbm_ctx *c, float ms[4])
{
    if (!c || !ms)
        return fail(SBM_ERR_BAD_ARG, "null argument");
    Slot &s = c->slot[0];
    if (c->timing && s.have_frame)
    {
        CK(cudaSetDevice(c->device));
        CK(cudaEventSynchronize(s.ev[2]));
        cudaEventElapsedTime(&c->last_ms[0], s.ev[0], s.ev[2]);
        cudaEventElapsedTime(&c->last_ms[1], s.ev[0], s.ev[1]);
        cudaEventElapsedTime(&c->last_ms[2], s.ev[1], s.ev[3]);
        cudaEventElapsedTime(&c->last_ms[3], s.ev[3], s.ev[2]);
    }
    memcpy(ms, c->last_ms, sizeof(c->last_ms));
    return 0;
}

int sbm_time_similarity(sbm_ctx *c, float threshold, int iters, float *ms_per_launch)
{
    if (!c || !ms_per_launch || iters < 1)
        return fail(SBM_ERR_BAD_ARG, "bad argument");
    std::lock_guard<std::mutex> lock(c->mu);
    Slot &s = c->slot[0];
    if (!s.have_frame || c->n_active == 0)
        return fail(SBM_ERR_STATE, "no frame has been processed");
    CK(cudaSetDevice(c->device));
    const int L = c->levels;
    {
        int rc_i;
        if ((rc_i = ensureSimItems(c, threshold)))
            return rc_i;
    }
    CK(cudaStreamSynchronize(s.stream));
    // `iters` launches as one captured graph (stream launches are CPU-paced on this host, tools/launch_quant.cu)
    cudaGraph_t tg = nullptr;
    cudaGraphExec_t te = nullptr;
    CK(cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal));
    for (int i = 0; i < iters; ++i)
        // same launch as the pipeline's; the candidate counter keeps running, records beyond cap are dropped
        CK(launchCoarse(c, s.lv[L - 1].d_lm, nullptr, nullptr, 0, c->d_active, c->n_active, threshold, s.d_buf[1], s.d_counts + 12, 0u, nullptr, s.stream));
    CK(cudaStreamEndCapture(s.stream, &tg));
    CK(cudaGraphInstantiate(&te, tg, 0));
    CK(cudaGraphLaunch(te, s.stream));
    CK(cudaEventRecord(s.ev[0], s.stream));
    CK(cudaGraphLaunch(te, s.stream));
    CK(cudaEventRecord(s.ev[1], s.stream));
    CK(cudaEventSynchronize(s.ev[1]));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, s.ev[0], s.ev[1]));
    cudaGraphExecDestroy(te);
    cudaGraphDestroy(tg);
    *ms_per_launch = ms / iters;
    return 0;
}

int sbm_time_l2_read(sbm_ctx *c, size_t bytes, int iters, float *gb_per_s)
{
    if (!c || !gb_per_s || iters < 1 || bytes < 4096 || bytes > ((size_t)1 << 30))
        return fail(SBM_ERR_BAD_ARG, "bad argument");
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    Slot &s = c->slot[0];
    int rc_s;
    if ((rc_s = ensureStream(s)))
        return rc_s;
    const uint32_t n_vec = (uint32_t)(bytes / 16 / 1024 * 1024); // whole 1024-vector rounds only
    if (n_vec == 0)
        return fail(SBM_ERR_BAD_ARG, "buffer too small");
    void *d = nullptr;
    uint32_t *sink = nullptr;
    CK(cudaMalloc(&d, (size_t)n_vec * 16));
    CK(cudaMalloc(&sink, 4));
    CK(cudaMemsetAsync(d, 1, (size_t)n_vec * 16, s.stream));
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    const int grid = sms * 4, reps = 8;
    k_l2_read<<<grid, 256, 0, s.stream>>>((const uint4 *)d, n_vec, reps, sink); // warm-up: pulls the buffer into L2
    CK(cudaEventRecord(s.ev[0], s.stream));
    for (int i = 0; i < iters; ++i)
        k_l2_read<<<grid, 256, 0, s.stream>>>((const uint4 *)d, n_vec, reps, sink);
    CK(cudaEventRecord(s.ev[1], s.stream));
    CK(cudaEventSynchronize(s.ev[1]));
    CK(cudaGetLastError());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, s.ev[0], s.ev[1]));
    cudaFree(d);
    cudaFree(sink);
    *gb_per_s = (float)((double)n_vec * 16.0 * reps * grid * iters / (ms * 1e-3) / 1e9);
    return 0;
}

int sbm_time_kernels(sbm_ctx *c, float threshold, int iters, float *ms, int n_ms)
{
    if (!c || !ms || iters < 1)
        return fail(SBM_ERR_BAD_ARG, "bad argument");
    std::lock_guard<std::mutex> lock(c->mu);
    Slot &s = c->slot[0];
    if (!s.have_frame || c->n_active == 0 || !s.last_src0)
        return fail(SBM_ERR_STATE, "no frame has been processed");
    const int L = c->levels;
    if (n_ms < 4 * L + 1)
        return fail(SBM_ERR_BAD_ARG, "ms[] needs %d entries", 4 * L + 1);
    CK(cudaSetDevice(c->device));
    {
        int rc_i;
        if ((rc_i = ensureSimItems(c, threshold)))
            return rc_i;
    }
    cudaStream_t st = s.stream;
    CK(cudaStreamSynchronize(st));
    const float weak_sq = c->weak * c->weak;
    for (int i = 0; i < n_ms; ++i)
        ms[i] = 0.f;
    // run(k, body): `iters` back-to-back launches of one kernel between two events, after 2 warm-up launches
#define SBM_TIME(slot_idx, launch)                                               \
    do                                                                           \
    {                                                                            \
        /* `iters` launches captured into one graph: stream launches on this host are CPU-paced (~5 us each, */ \
        /* 2 us quantum, tools/launch_quant.cu), graph nodes follow each other within ~0.7 us */               \
        cudaGraph_t tg_ = nullptr;                                               \
        cudaGraphExec_t te_ = nullptr;                                           \
        CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));        \
        for (int it_ = 0; it_ < iters; ++it_)                                    \
        {                                                                        \
            launch;                                                              \
        }                                                                        \
        CK(cudaStreamEndCapture(st, &tg_));                                      \
        CK(cudaGraphInstantiate(&te_, tg_, 0));                                  \
        CK(cudaGraphLaunch(te_, st));                                            \
        CK(cudaEventRecord(s.ev[0], st));                                        \
        CK(cudaGraphLaunch(te_, st));                                            \
        CK(cudaEventRecord(s.ev[1], st));                                        \
        CK(cudaEventSynchronize(s.ev[1]));                                       \
        float t_ = 0;                                                            \
        CK(cudaEventElapsedTime(&t_, s.ev[0], s.ev[1]));                         \
        ms[slot_idx] = t_ / iters;                                               \
        cudaGraphExecDestroy(te_);                                               \
        cudaGraphDestroy(tg_);                                                   \
    } while (0)
    const uint8_t *src = s.last_src0;
    for (int l = 0; l < L; ++l)
    {
        const LevelGeom &G = c->geom[l];
        LevelBuf &LB = s.lv[l];
        if (l > 0)
        {
            const LevelGeom &P = c->geom[l - 1];
            dim3 g((G.w + 31) / 32, (G.h + K1B_ROWS - 1) / K1B_ROWS);
            if (c->c == 3)
                SBM_TIME(3 * l + 0, (k_pyrdown<3><<<g, K1B_THREADS, 0, st>>>(src, P.w, P.h, LB.d_src, G.w, G.h)));
            else
                SBM_TIME(3 * l + 0, (k_pyrdown<1><<<g, K1B_THREADS, 0, st>>>(src, P.w, P.h, LB.d_src, G.w, G.h)));
            src = LB.d_src;
        }
        const uint8_t *nm = nullptr;
        SBM_TIME(3 * l + 1, (void)launchOrient(c->sms, c->c, src, G.w, G.h, nm, l, c->w, weak_sq, LB.d_quant, st));
        uint8_t *nd = nullptr;
        SBM_TIME(3 * l + 2, (void)launchSpread(G, LB.d_quant, LB.d_lm, nd, nd, st));
    }
    // scratch counters so that the frame's real candidate counts stay intact
    uint32_t *d_tmp = nullptr;
    CK(cudaMalloc(&d_tmp, COUNT_WORDS * 4));
    CK(cudaMemsetAsync(d_tmp, 0, COUNT_WORDS * 4, st));
    SBM_TIME(3 * L, (void)launchCoarse(c, s.lv[L - 1].d_lm, nullptr, nullptr, 0, c->d_active, c->n_active, threshold, s.d_buf[1], d_tmp, 0u, nullptr, st));
    // refinement of the frame's own coarse candidates (stage counts of the last real run), output discarded
    int stage = 1;
    for (int l = L - 2; l >= 0; --l, ++stage)
    {
        const LevelGeom &G = c->geom[l];
        if (stage == 1)
            SBM_TIME(3 * L + stage, (void)launchRefine(G.tiled, refineGrid(c, G.tiled), s.lv[l].d_lm, c->tl[l].d_tl, c->tl[l].d_feats, s.d_buf[0],
                                                       s.d_counts + 0, s.cap, s.d_buf[1], d_tmp + 1, 0u, G.w, G.h, G.T, threshold, nullptr, 0,
                                                       reinterpret_cast<unsigned long long *>(d_tmp + 10), st));
    }
    if (s.gexec && s.graph_epoch == c->epoch)
    {
        CK(cudaGraphLaunch(s.gexec, st));
        CK(cudaEventRecord(s.ev[0], st));
        for (int it = 0; it < iters; ++it)
            CK(cudaGraphLaunch(s.gexec, st));
        CK(cudaEventRecord(s.ev[1], st));
        CK(cudaEventSynchronize(s.ev[1]));
        float t = 0;
        CK(cudaEventElapsedTime(&t, s.ev[0], s.ev[1]));
        ms[4 * L] = t / iters;
    }
#undef SBM_TIME
    CK(cudaStreamSynchronize(st));
    cudaFree(d_tmp);
    CK(cudaGetLastError());
    return 0;
}

int sbm_phase_degrees(int device, const int *dy, const int *dx, int n, float *angle, uint8_t *label)
{
    if (!dy || !dx || n < 0 || (!angle && !label))
        return fail(SBM_ERR_BAD_ARG, "bad argument");
    CK(cudaSetDevice(device));
    struct DevBuf
    {
        void *p = nullptr;
        ~DevBuf() { cudaFree(p); }
    } d_y, d_x, d_a, d_l;
    const size_t nb = (size_t)std::max(n, 1) * 4;
    CK(cudaMalloc(&d_y.p, nb));
    CK(cudaMalloc(&d_x.p, nb));
    CK(cudaMalloc(&d_a.p, nb));
    CK(cudaMalloc(&d_l.p, (size_t)std::max(n, 1)));
    CK(cudaMemcpy(d_y.p, dy, (size_t)n * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_x.p, dx, (size_t)n * 4, cudaMemcpyHostToDevice));
    k_phase_tap<<<148 * 8, 256>>>((const int *)d_y.p, (const int *)d_x.p, n, (float *)d_a.p, (uint8_t *)d_l.p);
    CK(cudaGetLastError());
    if (angle)
        CK(cudaMemcpy(angle, d_a.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    if (label)
        CK(cudaMemcpy(label, d_l.p, (size_t)n, cudaMemcpyDeviceToHost));
    return 0;
}

int sbm_quantized_orientations(sbm_ctx *c, const uint8_t *img, int w, int h, int ch, size_t stride, int level, float *magnitude,
                               float *angle_ori, uint8_t *angle, uint8_t *src_level, int *w_out, int *h_out)
{
    if (!c || !img || level < 0 || level > 8)
        return fail(SBM_ERR_BAD_ARG, "bad argument");
    if (ch != 1 && ch != 3)
        return fail(SBM_ERR_BAD_ARG, "source must have 1 or 3 channels (got %d)", ch);
    if ((w >> level) < 1 || (h >> level) < 1)
        return fail(SBM_ERR_BAD_ARG, "image too small for pyramid level %d", level);
    std::lock_guard<std::mutex> lock(c->mu);
    CK(cudaSetDevice(c->device));
    Slot &s = c->slot[0];
    int rc_s;
    if ((rc_s = ensureStream(s)))
        return rc_s;
    cudaStream_t st = s.stream;
    const size_t row = (size_t)w * ch;
    // every early return (CK) releases what was allocated so far
    struct DevBuf
    {
        void *p = nullptr;
        ~DevBuf() { cudaFree(p); }
        void swap(DevBuf &o) { std::swap(p, o.p); }
    } a, b, q, mag, ang;
    CK(cudaMalloc(&a.p, row * h));
    if (stride == 0 || stride == row)
        CK(cudaMemcpyAsync(a.p, img, row * h, cudaMemcpyHostToDevice, st));
    else
        CK(cudaMemcpy2DAsync(a.p, row, img, stride, row, h, cudaMemcpyHostToDevice, st));
    int lw = w, lh = h;
    for (int l = 1; l <= level; ++l) // cv::pyrDown(src, src, Size(cols/2, rows/2))  line2Dup.cpp:431-433
    {
        const int dw = lw / 2, dh = lh / 2;
        DevBuf next;
        CK(cudaMalloc(&next.p, (size_t)dw * dh * ch));
        dim3 g((dw + 31) / 32, (dh + K1B_ROWS - 1) / K1B_ROWS);
        if (ch == 3)
            k_pyrdown<3><<<g, K1B_THREADS, 0, st>>>((const uint8_t *)a.p, lw, lh, (uint8_t *)next.p, dw, dh);
        else
            k_pyrdown<1><<<g, K1B_THREADS, 0, st>>>((const uint8_t *)a.p, lw, lh, (uint8_t *)next.p, dw, dh);
        CK(cudaStreamSynchronize(st)); // the previous level is freed when `next` leaves the scope
        a.swap(next);
        lw = dw, lh = dh;
    }
    const size_t px = (size_t)lw * lh;
    CK(cudaMalloc(&q.p, px));
    CK(cudaMalloc(&mag.p, px * sizeof(float)));
    CK(cudaMalloc(&ang.p, px * sizeof(float)));
    dim3 g1((lw + K1_TW - 1) / K1_TW, (lh + 15) / 16);
    const float weak_sq = c->weak * c->weak;
    if (ch == 3)
        k_orient<3, 16, true><<<g1, K1_THREADS, 0, st>>>((const uint8_t *)a.p, lw, lh, nullptr, 0, lw, weak_sq, (uint8_t *)q.p, (float *)mag.p,
                                                          (float *)ang.p);
    else
        k_orient<1, 16, true><<<g1, K1_THREADS, 0, st>>>((const uint8_t *)a.p, lw, lh, nullptr, 0, lw, weak_sq, (uint8_t *)q.p, (float *)mag.p,
                                                          (float *)ang.p);
    CK(cudaGetLastError());
    if (magnitude)
        CK(cudaMemcpyAsync(magnitude, mag.p, px * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (angle_ori)
        CK(cudaMemcpyAsync(angle_ori, ang.p, px * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (angle)
        CK(cudaMemcpyAsync(angle, q.p, px, cudaMemcpyDeviceToHost, st));
    if (src_level)
        CK(cudaMemcpyAsync(src_level, a.p, px * ch, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (w_out)
        *w_out = lw;
    if (h_out)
        *h_out = lh;
    return 0;
}
}

// ------------------------------------------------------------------------------------------------
// sbm_quantized_level1_static: pyramid level 1's quantised orientations for callers that cannot allocate (see the header).
// The scratch -- the pyrDown'ed source that k_orient reads -- is this file-scope array: every GPU that loads the library
// holds SBM_STATIC_SCRATCH_BYTES of it until the process exits.  In host code `g_static_scratch` is only a stand-in for
// the symbol; its device address is asked for once per device (cudaGetSymbolAddress, that device current) and kept.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) uint8_t g_static_scratch[SBM_STATIC_SCRATCH_BYTES];

namespace
{
struct StaticScratch
{
    uint8_t *ptr = nullptr;
    int sms = 0;
};
constexpr int STATIC_SCRATCH_MAX_DEVICES = 64;
std::mutex g_static_mu;
StaticScratch g_static_dev[STATIC_SCRATCH_MAX_DEVICES];
} // namespace

extern "C" int sbm_quantized_level1_static(const uint8_t *d_img, int w, int h, int channels, float weak_threshold, uint8_t *d_angle,
                                           void *stream)
{
    if (!d_img || !d_angle)
        return fail(SBM_ERR_BAD_ARG, "sbm_quantized_level1_static: d_img and d_angle must be device pointers");
    if (channels != 1 && channels != 3)
        return fail(SBM_ERR_BAD_ARG, "source must have 1 or 3 channels (got %d)", channels);
    if (w < 2 || h < 2)
        return fail(SBM_ERR_BAD_ARG, "image too small for pyramid level 1 (%d x %d)", w, h);
    const int dw = w / 2, dh = h / 2;
    const size_t need = (size_t)dw * dh * channels;
    if (need > (size_t)SBM_STATIC_SCRATCH_BYTES)
        return fail(SBM_ERR_CAPACITY, "sbm_quantized_level1_static: a %d x %d x %d source needs %zu bytes of scratch, the library's static array holds %d",
                    w, h, channels, need, (int)SBM_STATIC_SCRATCH_BYTES);
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= STATIC_SCRATCH_MAX_DEVICES)
        return fail(SBM_ERR_BAD_ARG, "device %d: at most %d devices", dev, STATIC_SCRATCH_MAX_DEVICES);
    StaticScratch sc;
    {
        std::lock_guard<std::mutex> lock(g_static_mu);
        if (!g_static_dev[dev].ptr) // set-up, once per device, with that device current
        {
            void *p = nullptr;
            CK(cudaGetSymbolAddress(&p, g_static_scratch));
            int sms = 0;
            if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms < 1)
                sms = 148;
            g_static_dev[dev].sms = sms;
            g_static_dev[dev].ptr = (uint8_t *)p;
        }
        sc = g_static_dev[dev];
    }
    cudaStream_t st = (cudaStream_t)stream;
    // cv::pyrDown(src, src, Size(cols/2, rows/2)) into the static array (every byte k_orient reads below is written here,
    // so nothing of the previous call's contents is seen and nothing needs clearing) ...
    dim3 g((dw + 31) / 32, (dh + K1B_ROWS - 1) / K1B_ROWS);
    if (channels == 3)
        k_pyrdown<3><<<g, K1B_THREADS, 0, st>>>(d_img, w, h, sc.ptr, dw, dh);
    else
        k_pyrdown<1><<<g, K1B_THREADS, 0, st>>>(d_img, w, h, sc.ptr, dw, dh);
    CK(cudaGetLastError());
    // ... and quantizedOrientations + hysteresisGradient on it, straight into the caller's buffer
    CK(launchOrient(sc.sms, channels, sc.ptr, dw, dh, nullptr, 0, dw, weak_threshold * weak_threshold, d_angle, st));
    return 0;
}

#ifdef SBM_TRACE
// experiments only (tools/stage_trace.py): the per-stage clock sums of K5a / K5b, optionally cleared
extern "C" int sbm_trace_read(unsigned long long *out16, int reset)
{
    if (out16 && cudaMemcpyFromSymbol(out16, sbm::g_trace, sizeof(unsigned long long) * 16) != cudaSuccess)
        return -1;
    if (reset)
    {
        unsigned long long z[16] = {0};
        if (cudaMemcpyToSymbol(sbm::g_trace, z, sizeof(z)) != cudaSuccess)
            return -1;
    }
    return 0;
}
#endif

#ifdef K5B_TRACE // experiments only (tools/sim_trace.py)
extern "C" int sbm_debug_k5b_trace(long long *out16)
{
    return (int)cudaMemcpyFromSymbol(out16, sbm::g_k5b_trace, 16 * sizeof(long long));
}
#endif
