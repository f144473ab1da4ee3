/* sbm_capi.h -- C ABI of the B200 (sm_100a) implementation of line2Dup::Detector::match.
 *
 * This is the drop-in boundary: plain C, pointers and sizes only.  The C++ drop-in header
 * (include/line2Dup.h) and the Python mirror (shape_based_matching_b200/detector.py) sit on top of
 * exactly these entry points; nothing below this line runs on the CPU except list ordering.
 *
 * Every function returns 0 on success or a negative sbm_status; sbm_last_error() gives the text
 * (thread-local).  There is no CPU fallback: without a CUDA device sbm_create fails with
 * SBM_ERR_CUDA.
 *
 * Reference interface each entry point replaces (file:line in /root/reference):
 *   sbm_create              Detector::Detector(int, std::vector<int>, float, float)   line2Dup.cpp:1056-1076
 *   sbm_set_class_templates Detector::readClass / addTemplate (template storage)      line2Dup.cpp:1299-1353,1507-1546
 *   sbm_match               Detector::match                                           line2Dup.cpp:1078-1150
 *   sbm_match_candidates    Detector::match without its sort/unique tail              line2Dup.cpp:1124-1140
 *   sbm_match_batch         Detector::match over a stream of frames (config 5)        line2Dup.cpp:1078-1150, test_jabil.cpp:314-374
 *   sbm_match_device        same pipeline, frame already resident in HBM              (bench `value`, multi-GPU)
 *   sbm_match_batch_device  same, a batch of resident frames kept in flight together
 *   sbm_get_intermediate    the locals of match(): quantized / spread / response maps / linear memories
 *                                                                                    line2Dup.cpp:1106-1117
 *   sbm_similarity_map      similarity / similarity_64                                line2Dup.cpp:807-858,924-984
 *   sbm_similarity_local    similarityLocal / similarityLocal_64                      line2Dup.cpp:860-922,986-1048
 *   sbm_merge_shard_records the `matches.insert(...)` reduction over templates + sort/unique, across GPUs   line2Dup.cpp:1124-1145,1166-1170
 *   sbm_group_create / sbm_group_match[_batch][_device]  Detector::match with the template loop (`#pragma omp parallel for`,
 *                           line2Dup.cpp:1166-1170) or the caller's frame loop (test_jabil.cpp:314-374) spread over the GPUs
 *                           of one box: one context per GPU, one ncclAllGather of the compact record lists
 *   sbm_set_candidate_capacity  the `std::vector<Match> candidates` of matchClass (unbounded there)       line2Dup.cpp:1160-1216
 *   sbm_quantized_orientations  quantizedOrientations + hysteresisGradient + pyrDown (training)  line2Dup.cpp:218-444
 *   sbm_quantized_level1_static  (no counterpart in the reference; operators' request) the same at level 1 without any allocation
 */
#ifndef SBM_CAPI_H
#define SBM_CAPI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sbm_ctx sbm_ctx;

typedef enum sbm_status {
    SBM_OK = 0,
    SBM_ERR_BAD_ARG = -1,   /* cv::Error::StsBadArg / CV_Assert in the reference */
    SBM_ERR_CUDA = -2,      /* CUDA runtime failure, or no device */
    SBM_ERR_CAPACITY = -3,  /* caller's output buffer too small; *n_out holds the needed size */
    SBM_ERR_STATE = -4      /* e.g. intermediate requested before any frame was processed */
} sbm_status;

/* One Template of a TemplatePyramid (line2Dup.h:131-153), without its feature vector. */
typedef struct sbm_template_desc {
    int32_t width, height, tl_x, tl_y;
    int32_t num_features; /* features.size(); the features follow each other in the sbm_feature array */
} sbm_template_desc;

/* line2Dup::Feature (line2Dup.h:116-129) without theta (theta is not used by match()). */
typedef struct sbm_feature {
    int32_t x, y, label;
} sbm_feature;

/* line2Dup::Match (line2Dup.h:222-255); class_idx indexes the caller's class list. */
typedef struct sbm_match_rec {
    int32_t x, y;
    float similarity;
    int32_t class_idx;
    int32_t template_id;
} sbm_match_rec;

typedef enum sbm_intermediate {
    SBM_QUANTIZED = 0, /* u8  [h_l][w_l]                      ColorGradientPyramid::quantize  :446-450 */
    SBM_SPREAD = 1,    /* u8  [h_l][w_l]                      spread                          :616-630 */
    SBM_RESPONSE = 2,  /* u8  [8][h_l][w_l]                   computeResponseMaps             :637-747 */
    SBM_LINEAR = 3,    /* u8  [8][T*T][(w_l/T)*(h_l/T)]       linearize                       :749-777 */
    SBM_PYRAMID_SRC = 4 /* u8 [h_l][w_l][channels]            cv::pyrDown chain               :424-444 */
} sbm_intermediate;

const char *sbm_last_error(void);
/* Compile-time facts, so that a loader can verify it got the sm_100a build. */
const char *sbm_build_info(void);

/* T[levels]: spread/stride per pyramid level (default {4,8}); weak/strong: magnitude thresholds
 * (defaults 30/60); num_features: the training-time feature budget (kept for API parity; match() does
 * not use it).  device: CUDA ordinal. */
int sbm_create(const int *T, int levels, float weak, float strong, int num_features, int device, sbm_ctx **out);
int sbm_destroy(sbm_ctx *ctx);

/* Replace the template set of class slot `class_idx` (0..). desc has n_templates*levels entries ordered
 * [template][level]; feats holds sum(num_features) entries in the same order.  n_templates == 0 removes
 * the class.  Templates with >= 8192 features at any level are rejected (line2Dup.cpp:811,1195). */
int sbm_set_class_templates(sbm_ctx *ctx, int class_idx, int n_templates, const sbm_template_desc *desc,
                            const sbm_feature *feats);
int sbm_clear_templates(sbm_ctx *ctx);
int sbm_num_templates(sbm_ctx *ctx, int class_idx);
/* Records each stage of a frame may hold inside the context (default 65536 per frame slot).  The host entry points
 * (sbm_match, sbm_match_batch, sbm_group_match*) grow these buffers and re-run a frame that overflowed them; the
 * device-resident entry points never synchronise, so they raise the overflow flag instead (d_count[1], block header) --
 * the caller then re-runs the frame through sbm_match or after raising the capacity here. */
int sbm_set_candidate_capacity(sbm_ctx *ctx, int records);

/* Detector::match.  img: 8-bit, channels 1 or 3 (interleaved), `stride` bytes per row (0 = packed);
 * w,h must satisfy the reference's asserts at every level (h_l % T_l == 0, w_l % T_l == 0,
 * h_l*w_l % 16 == 0).  mask: NULL or u8 [h][w] (packed).  class_idx[n_classes]: class slots to match in
 * this order (NULL / 0 = all slots in ascending order, i.e. the caller's std::map order).
 * out[cap] receives the final list: sorted by (similarity desc, template_id asc) and de-duplicated by the
 * same std::sort/std::unique calls as the reference, fed in the reference's serial candidate order. */
int sbm_match(sbm_ctx *ctx, const uint8_t *img, int w, int h, int channels, size_t stride, const uint8_t *mask,
              float threshold, const int *class_idx, int n_classes, sbm_match_rec *out, int cap, int *n_out);

/* Same, but stops before sort/unique: the candidate list in the order the reference appends it with one
 * OpenMP thread (class order, template_id ascending, coarse scan order). */
int sbm_match_candidates(sbm_ctx *ctx, const uint8_t *img, int w, int h, int channels, size_t stride,
                         const uint8_t *mask, float threshold, const int *class_idx, int n_classes,
                         sbm_match_rec *out, int cap, int *n_out);

/* Detector::match over n_frames frames of one shape (the streaming configuration; the reference's caller loops
 * over the images of a directory, test_jabil.cpp:314-374).  Several frames are kept in flight: the H2D copy
 * of frame i+1 overlaps the kernels of frame i.  imgs[i]: host pointer of frame i (pinned memory makes the copies
 * truly asynchronous); masks: NULL, or n_frames pointers (entries may be NULL).  The final lists are
 * concatenated in out[cap]; frame i owns out[frame_begin[i] .. frame_begin[i+1]) (frame_begin has n_frames+1
 * entries).  Every list is exactly what sbm_match returns for that frame.  On SBM_ERR_CAPACITY
 * frame_begin[n_frames] is the capacity needed. */
int sbm_match_batch(sbm_ctx *ctx, const uint8_t *const *imgs, const uint8_t *const *masks, int n_frames, int w, int h,
                    int channels, size_t stride, float threshold, const int *class_idx, int n_classes, sbm_match_rec *out,
                    int cap, int *frame_begin);

/* Device-resident variant: d_img is a packed u8 [h][w][channels] frame in HBM.  The whole pipeline is
 * enqueued on `stream` (a cudaStream_t, or NULL for the context's own stream) and nothing is copied to
 * the host.  Results stay on the device:
 *   d_records: sbm_device_rec[cap]   (unordered; see sbm_device_rec)
 *   d_count  : uint32[2]             ([0] = number of records produced (may exceed cap), [1] = overflow flag: the list
 *                                     is incomplete because cap, or the context's own per-stage candidate capacity
 *                                     (sbm_set_candidate_capacity), was too small)
 * Either may be NULL, in which case the context's internal buffers are used (sbm_device_results). */
typedef struct sbm_device_rec {
    int32_t x, y;          /* level-0 top-left of the match */
    float similarity;
    uint32_t template_gid; /* global template index: templates of class slot 0 first, then slot 1, ... */
    uint32_t order_key;    /* coarse scan index j: orders candidates of one template like the reference */
} sbm_device_rec;

int sbm_match_device(sbm_ctx *ctx, const void *d_img, int w, int h, int channels, const void *d_mask, float threshold,
                     const int *class_idx, int n_classes, void *d_records, int cap, void *d_count, void *stream);
/* A batch of resident frames: frame i = d_imgs + i*frame_stride bytes.  Frames are spread over the context's
 * internal streams (forked from / joined to `stream`), so neighbouring frames overlap on the GPU.
 * d_records: sbm_device_rec[n_frames][cap_per_frame]; d_counts: uint32[n_frames][2] (as in sbm_match_device). */
int sbm_match_batch_device(sbm_ctx *ctx, const void *d_imgs, size_t frame_stride, int n_frames, int w, int h, int channels,
                           float threshold, const int *class_idx, int n_classes, void *d_records, int cap_per_frame,
                           void *d_counts, void *stream);
int sbm_device_results(sbm_ctx *ctx, void **d_records, void **d_count, int *cap);
/* Turn device records (copied to the host by the caller, e.g. after an NCCL allgather) into the final
 * list: same ordering + std::sort + std::unique as sbm_match.  finalize != 0 applies sort/unique. */
int sbm_finalize_records(sbm_ctx *ctx, const sbm_device_rec *recs, int n, const int *class_idx, int n_classes,
                         int finalize, sbm_match_rec *out, int cap, int *n_out);

/* Template sharding over several GPUs (templates are independent given the frame's linear memories,
 * line2Dup.cpp:1166-1170): every rank installs a slice of each class and matches the same frame; the ranks'
 * sbm_device_rec lists are exchanged (NCCL allgather) and merged here.  No context and no GPU needed: pure host code.
 *   recs         the ranks' records laid end to end, rank 0 first; n_per_rank[n_ranks] how many each rank produced
 *   class_sizes  int[n_ranks][n_classes]: templates of class slot k installed on rank r (template_gid of a
 *                record is LOCAL to its rank: class slot 0's templates first, then slot 1, ...)
 *   mode         SBM_SHARD_INTERLEAVED: rank r holds global ids r, r+G, r+2G, ... of every class;
 *                SBM_SHARD_CONTIGUOUS : rank r holds the block that follows the blocks of ranks < r
 * Output as sbm_finalize_records: the reference's serial candidate order over GLOBAL template ids, then
 * (finalize != 0) Detector::match's std::sort + std::unique (line2Dup.cpp:1143-1145).
 * class_order[n_order]: the caller's class list (class slots, repeats allowed, entries outside 0..n_classes-1 skipped):
 *                the reference appends class by class in that order (:1124-1140); NULL / 0 = every class slot once, ascending. */
enum { SBM_SHARD_INTERLEAVED = 0, SBM_SHARD_CONTIGUOUS = 1 };
int sbm_merge_shard_records(const sbm_device_rec *recs, const int *n_per_rank, int n_ranks, const int *class_sizes, int n_classes,
                            int mode, const int *class_order, int n_order, int finalize, sbm_match_rec *out, int cap, int *n_out);

/* ---- Multi-GPU group: one context per GPU (one process or thread each), NCCL over NVLink / NVSwitch -------------------
 * The path shards two ways, both with ONE exchange of KBs per batch:
 *   SBM_GROUP_TEMPLATES[_CONTIGUOUS]  every rank is handed the SAME frames and the FULL classes (sbm_set_class_templates
 *        keeps this rank's slice: ids rank, rank + n, ... or its contiguous block); each rank builds the frames' linear
 *        memories and matches its slice; the merged list of a frame is identical, on every rank, to the unsharded sbm_match.
 *   SBM_GROUP_FRAMES  every rank holds all templates and its OWN frames; the per-frame lists of all ranks are gathered.
 * A frame's result travels as one block of sbm_group_block_bytes(cap) bytes: {uint32 count, uint32 overflow, 0, 0} followed
 * by cap sbm_device_rec; the ranks' blocks are exchanged with one ncclAllGather on the stream the kernels ran on (counts
 * and records in one buffer).  libnccl.so.2 is bound at run time on the first sbm_group_* call (the copy already loaded in
 * the process, else the system's; SBM_NCCL_LIB overrides); single-GPU callers never need it.
 *
 * Set-up: rank 0 calls sbm_group_unique_id and hands the 128 bytes to the other ranks by any means (file, pipe, MPI,
 * torch.distributed ...); every rank then calls sbm_group_create (collective: returns when all n ranks have called it),
 * and only then installs its templates.  All sbm_group_match* calls are collective: every rank makes the same call with
 * the same n_frames, threshold, class list and capacities. */
enum { SBM_GROUP_TEMPLATES = 0, SBM_GROUP_TEMPLATES_CONTIGUOUS = 1, SBM_GROUP_FRAMES = 2 };
#define SBM_GROUP_ID_BYTES 128
int sbm_group_unique_id(void *id128);
int sbm_group_create(sbm_ctx *ctx, int n_ranks, int rank, const void *id128, int mode);
int sbm_group_destroy(sbm_ctx *ctx);
/* returns 1 when the context belongs to a group (and fills what is non-NULL), 0 when not */
int sbm_group_info(sbm_ctx *ctx, int *n_ranks, int *rank, int *mode);
/* Detector::match over the group, template sharding: same arguments and same result as sbm_match, on every rank. */
int sbm_group_match(sbm_ctx *ctx, const uint8_t *img, int w, int h, int channels, size_t stride, const uint8_t *mask,
                    float threshold, const int *class_idx, int n_classes, sbm_match_rec *out, int cap, int *n_out);
/* sbm_match_batch over the group.  Template sharding: every rank passes the same n_frames frames; out / frame_begin
 * describe n_frames lists.  Frame sharding: every rank passes its own n_frames frames (the same count on every rank); out /
 * frame_begin describe n_ranks * n_frames lists, rank-major (list r * n_frames + f = frame f of rank r), on every rank.
 * Internal block and stage capacities grow (on all ranks alike) when a frame overflows them. */
int sbm_group_match_batch(sbm_ctx *ctx, const uint8_t *const *imgs, const uint8_t *const *masks, int n_frames, int w, int h,
                          int channels, size_t stride, float threshold, const int *class_idx, int n_classes, sbm_match_rec *out,
                          int cap, int *frame_begin);
/* Device-resident and asynchronous: the frames (d_imgs + i * frame_stride) are matched on the context's frame slots, the
 * n_frames result blocks are allgathered on `stream` into d_gathered = uint8 [n_ranks][n_frames][sbm_group_block_bytes(cap)]
 * and nothing is copied to the host.  sbm_group_finalize turns a host copy of d_gathered into the final lists (as
 * sbm_group_match_batch lays them out); it fails with SBM_ERR_CAPACITY when a block's overflow flag is set. */
size_t sbm_group_block_bytes(int cap_per_frame);
int sbm_group_match_batch_device(sbm_ctx *ctx, const void *d_imgs, size_t frame_stride, int n_frames, int w, int h, int channels,
                                 float threshold, const int *class_idx, int n_classes, void *d_gathered, int cap_per_frame,
                                 void *stream);
int sbm_group_finalize(sbm_ctx *ctx, const void *h_gathered, int n_frames, int cap_per_frame, const int *class_idx, int n_classes,
                       int finalize, sbm_match_rec *out, int cap, int *frame_begin);

/* Copy an intermediate of the most recent frame to host memory `dst` (`bytes` must match exactly;
 * pass dst == NULL to query the size through *bytes_needed).  SPREAD and RESPONSE are only materialised
 * when sbm_set_debug(ctx, 1) was called before the frame. */
int sbm_set_debug(sbm_ctx *ctx, int enable);
int sbm_get_intermediate(sbm_ctx *ctx, int level, int kind, void *dst, size_t bytes, size_t *bytes_needed);

/* Raw u16 [H][W] similarity sums of one template at the coarsest level on the most recent frame. */
int sbm_similarity_map(sbm_ctx *ctx, int class_idx, int template_id, uint16_t *dst, size_t bytes);
/* Raw u16 [16][16] local sums at pyramid level `level` around centre (x,y) (level coordinates). */
int sbm_similarity_local(sbm_ctx *ctx, int class_idx, int template_id, int level, int x, int y, uint16_t *dst);

/* Training front end (Detector::addTemplate -> ColorGradientPyramid::update, line2Dup.cpp:406-444): the image is taken
 * `level` times through cv::pyrDown (size cols/2 x rows/2 each time) and then through quantizedOrientations +
 * hysteresisGradient (:218-404) on the GPU.  Any image size is accepted (no T constraints: templates are cropped
 * objects).  Outputs (each may be NULL), all [h_out][w_out] of the reached level: magnitude = dx*dx+dy*dy of the
 * strongest channel (float), angle_ori = phase in degrees (float), angle = one-hot quantised label (u8, 0 = none);
 * src_level = the pyrDown'ed source itself, u8 [h_out][w_out][channels]. */
int sbm_quantized_orientations(sbm_ctx *ctx, const uint8_t *img, int w, int h, int channels, size_t stride, int level,
                               float *magnitude, float *angle_ori, uint8_t *angle, uint8_t *src_level, int *w_out, int *h_out);

/* Test tap: cv::phase in degrees (line2Dup.cpp:327, :398) and the 16-bin label hysteresisGradient derives from it (:225,
 * :249) for n integer Sobel responses (dy[i], dx[i]), computed by the device functions the orientation kernel uses.
 * Either output may be NULL.  tests/test_gpu_phase.py runs every pair of [-1020, 1020]^2 against the OpenCV restatement. */
int sbm_phase_degrees(int device, const int *dy, const int *dx, int n, float *angle, uint8_t *label);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Pyramid level 1 of the training /
 * matching front end -- cv::pyrDown (cols/2 x rows/2) followed by quantizedOrientations + hysteresisGradient, i.e. what
 * sbm_quantized_orientations(level = 1) returns as `angle` -- for a frame that already lies in device memory, written to
 * the caller's device buffer d_angle (u8 one-hot labels, [h/2][w/2]), asynchronously on `stream` (NULL = the legacy default
 * stream) of the CURRENT device.  The call allocates nothing (no cudaMalloc / cudaFree / cudaMallocHost, no context, no
 * library that allocates): the pyrDown'ed image, the kernel's scratch, lives in a static device array of
 * SBM_STATIC_SCRATCH_BYTES inside the library, which every GPU that loads the library holds until the process exits.
 * A source whose level-1 image ((w/2) * (h/2) * channels bytes) exceeds that fails with SBM_ERR_CAPACITY and a message
 * naming both sizes: the limit is a 2048 x 2048 x 3 source.  All calls on a device share the one array (it keeps the last
 * call's image; each call overwrites every byte it later reads, so nothing has to be cleared): calls that overlap on the
 * GPU -- different streams, or two host threads -- race; serialise them on one stream or with events.  The first call on a
 * device looks the array's address up (cudaGetSymbolAddress); make that call outside stream capture.
 * weak_threshold: the detector's weak threshold (magnitude > weak^2 qualifies a pixel, line2Dup.cpp:270). */
#define SBM_STATIC_SCRATCH_BYTES (3 * 1024 * 1024)
int sbm_quantized_level1_static(const uint8_t *d_img, int w, int h, int channels, float weak_threshold, uint8_t *d_angle, void *stream);

/* Counters of the most recent frame: [0] coarse evaluations (sum over matched templates of
 * in-image features * template_positions), [1] coarse candidates, [2] local evaluations,
 * [3] records after refinement, [4] kernels launched by the most recent call, [5] layout of the coarsest level's linear
 * memories for the current frame shape: 2 = hit / any bit planes (its W is a multiple of 32; k_similarity_bits), 1 = two
 * responses per byte (W even; k_similarity_nib), 0 = one byte per response (k_similarity),
 * [6] 1 when level 0 is stored tiled for the refinement (its W is a multiple of 16; k_refine_tiled). */
int sbm_last_stats(sbm_ctx *ctx, uint64_t stats[8]);

/* Device time of the most recent match call in milliseconds (CUDA events on the launching stream;
 * needs sbm_set_timing(ctx, 1)): [0] whole pipeline, [1] front end (orientations .. linear memories),
 * [2] coarse similarity kernel (+threshold/compaction), [3] local refinement. */
int sbm_last_timing(sbm_ctx *ctx, float ms[4]);
int sbm_set_timing(sbm_ctx *ctx, int enable);
/* Average device time of ONE k_similarity launch (the roofline kernel) on the most recent frame's linear memories:
 * `iters` back-to-back launches of the pipeline's own grid between two CUDA events on the launching stream. */
int sbm_time_similarity(sbm_ctx *ctx, float threshold, int iters, float *ms_per_launch);
/* Read bandwidth of an L2-resident buffer of `bytes` bytes (GB/s; 16-byte loads that bypass L1, all SMs, CUDA events):
 * the second denominator SURVEY 8d asks for next to the HBM copy peak, since the coarse linear memories (2 MiB at
 * M1) never leave L2.  A measurement aid, not part of the match path. */
int sbm_time_l2_read(sbm_ctx *ctx, size_t bytes, int iters, float *gb_per_s);
/* Warm-cache device time per launch of every kernel of the pipeline on the most recent frame (same method):
 * ms[3l+0] pyrDown of level l (0 for l = 0), ms[3l+1] orientation, ms[3l+2] spread+response+linearize,
 * ms[3L] coarse similarity, ms[3L+1] refinement of the frame's coarse candidates at level L-2,
 * ms[4L] one launch of the whole-frame CUDA graph (0 if none is instantiated).  n_ms >= 4L+1. */
int sbm_time_kernels(sbm_ctx *ctx, float threshold, int iters, float *ms, int n_ms);

#ifdef __cplusplus
}
#endif
#endif /* SBM_CAPI_H */
