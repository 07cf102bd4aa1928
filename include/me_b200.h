/*
 * me_b200.h -- C ABI of the B200-native displacement-search library (libme_b200.so).
 *
 * This is the drop-in boundary for the hot path of ruzko/motion-estimation: the per-frame-pair
 * displacement estimate and its reduction to one (pdx, pdy, flag) per pair.  The reference has no
 * FFI or plugin interface (it is three Python scripts); the entry points below are what a ctypes
 * binding for its two call sites needs:
 *
 *     calc_ECC_transform(prevFrame, curFrame) -> (pdx, pdy, measurement_flag)
 *         estimate_vel_vid.py:306-369 (caller :522) and estimate_vel.py:286-297 (caller :361)
 *     end_process(clean_arrs) pair loop
 *         estimate_vel_vid.py:493-583, estimate_vel.py:333-378
 *
 * Conventions: plain C types only, no exceptions, every function returns 0 (ME_OK) or a negative
 * ME_E* code and leaves a message for me_last_error().  The caller owns every host buffer it
 * passes; the library owns all device memory.  One context per GPU; a context is thread-compatible
 * (use it from one thread at a time), contexts on different GPUs are independent.
 *
 * Displacements are in FULL-resolution pixels, +x content motion gives pdx > 0 (the reference's
 * sign).  The Python wrapper halves them where it emulates the video script, whose ECC result is in
 * half-resolution pixels (estimate_vel_vid.py:330,342-346,358).
 *
 * The search itself is specified in oracle/sad_oracle.c ("the spec"): full-search SAD over BxB
 * blocks and +-R displacements restricted to in-frame candidates, first-minimum-in-raster-order
 * tie-break, per-axis 3-point parabola refinement, per-axis median over textured+peaked blocks.
 */
#ifndef ME_B200_H
#define ME_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ME_OK            0
#define ME_EINVAL       -1   /* bad argument (sizes, NULL, slot out of range ...)            */
#define ME_EUNSUPPORTED -2   /* legal in the spec but not built yet (see message)             */
#define ME_ECUDA        -3   /* a CUDA runtime/driver call failed; message has the CUDA text  */
#define ME_ENOMEM       -4
#define ME_ESTATE       -5   /* call order (fetch before search ...)                          */

/* per-block flag bits (same values as oracle/sad_oracle.c) */
#define ME_FLAG_TEXTURED   1u
#define ME_FLAG_X_INTERIOR 2u
#define ME_FLAG_Y_INTERIOR 4u
#define ME_FLAG_X_PEAKED   8u
#define ME_FLAG_Y_PEAKED   16u

typedef struct me_ctx me_ctx;

/* Search parameters.  The reference has none of these (its ECC call takes motionType/criteria,
 * estimate_vel_vid.py:308-309); they are the knobs of the block search that replaces it. */
typedef struct me_params {
    int32_t block;      /* B: 8, 16 or 32                                                     */
    int32_t range;      /* R: 1 .. 128 (large ranges are cut into slabs, one CTA each)        */
    int32_t min_range;  /* texture gate: block max-min >= min_range grey levels               */
    int32_t subpixel;   /* 0 = integer displacements only                                     */
    int32_t min_valid;  /* pair flag = nvalid_x >= max(min_valid, 1)                          */
    int32_t min_curv;   /* peak gate on cm - 2 c0 + cp; values < 1 mean 1                     */
    int32_t range_y;    /* Ry: vertical range, 1 .. range; 0 = the same as range (square search).  The
                           example recordings move along x only: a small Ry cuts the work accordingly */
} me_params;

/* One record per frame pair: replaces the (pdx, pdy, measurement_flag) tuple returned at
 * estimate_vel_vid.py:369. */
typedef struct me_pair_result {
    float   pdx, pdy;       /* full-resolution pixels                                         */
    int32_t flag;           /* 1 = measurement accepted (the reference's measurement_flag)    */
    int32_t nvalid_x;       /* blocks that entered the x median                               */
    int32_t nvalid_y;
    int32_t n_textured;     /* blocks that passed the texture gate: nvalid_x / n_textured is the share of
                               usable blocks with a strict interior minimum (low when the motion left +-R) */
    float   mean_dx, mean_dy; /* frame mean of the valid blocks' refined displacements: the role of the
                               reference's calculate_dense_flow (estimate_vel_vid.py:385-391, summed at
                               :524-526), a by-product of the search                              */
} me_pair_result;

/* Static information about a context's geometry. */
typedef struct me_geometry {
    int32_t height, width;
    int32_t nbx, nby;           /* block grid (floor)                                         */
    int32_t max_frames;         /* ring slots                                                 */
    int32_t max_pairs;          /* pairs per me_search call                                   */
    int32_t tile_blocks_x;      /* blocks per CTA tile                                        */
    int32_t threads_per_cta;
    int32_t smem_bytes;
    int32_t tiles_per_pair;     /* tiles per pair; CTAs per pair = tiles * slabs_x * slabs_y  */
    int64_t cand_px_per_pair;         /* blocks * B^2 * (2R+1)^2 (nominal, what the ALU executes) */
    int64_t cand_px_per_pair_valid;   /* in-frame candidates only (algorithmic work)              */
    int32_t sm_count;
    int32_t uses_tma;
    int32_t slabs_x, slabs_y;   /* the candidate square is cut into slabs_x * slabs_y CTAs per tile   */
    int32_t window_pitch;       /* bytes per shared-memory window row (a template constant)         */
    int32_t resident_kernel;    /* 1: launches of >= 2*sm_count tiles run the one-CTA-per-SM resident
                                   kernel (producer warp + consumer warps); 0: always one CTA per tile */
    int32_t tile_blocks_y;      /* block rows per CTA tile (> 1 for small ranges: the rows share the halo) */
} me_geometry;

const char *me_last_error(void);          /* thread-local, never NULL                         */
const char *me_version(void);

/* Create a context on CUDA device `device` for frames of height x width luma bytes.
 * Allocates the device frame ring (max_frames slots), pinned staging and result buffers. */
int me_create(int device, int height, int width, const me_params *params, int max_frames, int max_pairs,
              me_ctx **out);
int me_destroy(me_ctx *ctx);
int me_get_geometry(const me_ctx *ctx, me_geometry *out);

/* Frame ingest -- replaces the decode-and-stack loop at estimate_vel_vid.py:165-195 from the point
 * where luma bytes exist.  Copies n_frames frames (row_pitch bytes between rows, frames contiguous:
 * frame f starts at luma + f*height*row_pitch) into the pinned staging ring and from there to device
 * slots first_slot.. with cudaMemcpyAsync on the copy stream.  Returns once the host buffer may be
 * reused; the device copy may still be in flight (me_search orders itself after it).  A source that
 * is already page-locked is copied directly, without the staging hop: it must then stay untouched
 * until the next me_search / me_sync. */
int me_upload(me_ctx *ctx, const uint8_t *luma, size_t row_pitch, int first_slot, int n_frames);

/* Same for interleaved BGR frames as cv.VideoCapture delivers them (estimate_vel_vid.py:173-178):
 * the BGR2GRAY conversion runs on the GPU, bit-exact with cv.cvtColor's 15-bit fixed point. */
int me_upload_bgr(me_ctx *ctx, const uint8_t *bgr, size_t row_pitch, int first_slot, int n_frames);

/* Contrast-limited adaptive histogram equalisation of ring slots, in place, on the GPU -- replaces the
 * per-frame cv.createCLAHE(clipLimit=2.0, tileGridSize=(8,8)).apply(z) of adaptive_histogram_equalization
 * (estimate_vel_vid.py:201-219), bit-exact with OpenCV's 8-bit CLAHE.  Runs on the copy stream, after the
 * uploads issued so far; later searches wait for it. */
int me_clahe(me_ctx *ctx, int first_slot, int n_frames, double clip_limit, int tiles_x, int tiles_y);

/* Non-local-means denoising of ring slots [first_slot, first_slot + n_frames), in place, on the GPU --
 * replaces cv.fastNlMeansDenoising(src, h, templateWindowSize, searchWindowSize) and
 * cv.fastNlMeansDenoisingMulti(srcImgs, k, temporalWindowSize, h, ...) as denoise_hf5 / denoising call them
 * (estimate_vel_vid.py:226-232, 246-262), bit-exact with OpenCV's 8-bit implementation.
 * temporal_window[k] (odd; 1 = the single-image function; 0 = leave frame k as it is, it only serves as a
 * neighbour) is the window of frame k of the call; every frame is denoised from the ORIGINAL neighbours, as
 * the reference does (it reads eq_arrs, writes a new array).  A window that leaves the call's frames is ME_EINVAL (cv2 raises there).  template_window 7 or
 * 19 are built; search_window <= 41.  Runs on the copy stream like me_clahe. */
int me_denoise(me_ctx *ctx, int first_slot, int n_frames, const int32_t *temporal_window, float h,
               int template_window, int search_window);

/* Copy ring slots back to host memory (row_pitch bytes between rows); waits for pending uploads and CLAHE.
 * The reference keeps every stage's frames as (N,H,W) host arrays (estimate_vel_vid.py:215-219). */
int me_download(me_ctx *ctx, uint8_t *luma, size_t row_pitch, int first_slot, int n_frames);

/* The same towards device memory of this GPU (d_dst: n_frames x height rows of dst_pitch bytes): pre-processed frames
 * go on to another GPU consumer -- e.g. the send buffer of the all_gather that reassembles a recording whose frame
 * ranges were cleaned on different GPUs -- without passing through the host.  Ordered after the uploads / CLAHE /
 * denoising issued so far; asynchronous on `stream` (NULL = the context's copy stream). */
int me_copy_frames_device(me_ctx *ctx, void *d_dst, size_t dst_pitch, int first_slot, int n_frames, void *stream);

/* Launch the search for n_pairs pairs of ring slots: pairs[2k] = previous frame, pairs[2k+1] =
 * current frame.  Asynchronous; results are read with me_fetch_*.  This is the call that replaces
 * the body of calc_ECC_transform for a whole batch. */
int me_search(me_ctx *ctx, const int32_t *pairs, int n_pairs);

/* Wait for the last me_search and copy its per-pair records to `out` (n_pairs of them). */
int me_fetch_pairs(me_ctx *ctx, me_pair_result *out, int n_pairs);

/* Searches in flight.  me_search only queues work: it may be called again before the records of the search before it were
 * fetched (pair lists and pair records live in a ring of four slots; the per-block records do not -- me_fetch_blocks always
 * sees the most recent search).  me_last_ticket names the search just launched; me_fetch_pairs_ticket waits for THAT search
 * only (on a stream of its own, not behind the searches launched after it) and copies its records.  A ticket is valid until
 * three more searches have been launched.  Keeping two searches in flight removes the idle gap between a search and the
 * launch of the next one: batch users of the reference's consecutive-pair loop (estimate_vel.py:333-378). */
int me_last_ticket(const me_ctx *ctx);
int me_fetch_pairs_ticket(me_ctx *ctx, int ticket, me_pair_result *out, int n_pairs);

/* Per-block results of pair `pair_index` of the last me_search; any pointer may be NULL.
 * Layouts (row-major over by, bx): mv int16[nb][2] (dx,dy); sad uint32[nb]; sub float[nb][2];
 * flags uint8[nb].  This is the dense field the reference only has through Farneback
 * (estimate_vel_vid.py:385-391). */
int me_fetch_blocks(me_ctx *ctx, int pair_index, int16_t *mv, uint32_t *sad, float *sub, uint8_t *flags);

/* One-shot: upload two host frames, search, fetch.  What calc_ECC_transform(prev, cur) binds. */
int me_estimate_pair(me_ctx *ctx, const uint8_t *prev, const uint8_t *cur, size_t row_pitch,
                     me_pair_result *out);

/* Search frames that already live in device memory (e.g. a torch uint8 tensor): d_frames is
 * (n_frames, height, pitch) with pitch % 16 == 0 and a 16-byte aligned base.  Uses the context's
 * result buffers; `stream` is a cudaStream_t (NULL = the context's compute stream). */
int me_search_device(me_ctx *ctx, const uint8_t *d_frames, size_t pitch, int n_frames, const int32_t *pairs,
                     int n_pairs, void *stream);

int me_sync(me_ctx *ctx);

/* Device time of the most recent search kernel, measured with CUDA events on its stream. */
int me_last_search_ms(me_ctx *ctx, float *ms);
/* Number of kernels this context has launched so far (search, ingest conversion ...). */
int64_t me_launch_count(const me_ctx *ctx);
/* Device-side stopwatch on the context's compute stream (CUDA events): everything launched between
 * start and stop.  me_timer_stop blocks until that work has finished. */
int me_timer_start(me_ctx *ctx);
int me_timer_stop(me_ctx *ctx, float *ms);

/* The same stopwatch on the copy stream, where me_upload, me_clahe and me_denoise run. */
int me_copy_timer_start(me_ctx *ctx);
int me_copy_timer_stop(me_ctx *ctx, float *ms);

/* The anchored pair walk of the video script -- the control flow of end_process, estimate_vel_vid.py:506-534 and
 * :575 -- run inside the library so that no interpreter sits between two dependent searches.  Frame k is compared
 * with the last ACCEPTED frame (the anchor), not with k-1: a pair is skipped, and the anchor kept, when
 * pdx * scale < gate and flag == 1 (:531-534); otherwise frame k becomes the anchor (:575).
 *
 * Frames first_slot .. first_slot + n_frames - 1 of the ring are classified in order; the walk starts with the
 * anchor in ring slot anchor_slot (any slot: the caller may keep an older anchor frame in a spare slot).  out[i]
 * describes frame first_slot + i: the slot it was compared with, whether it was accepted, and the pair record of
 * that comparison (full-resolution pixels).  Pairs are searched speculatively in batches (anchor, k), (anchor, k+1),
 * ... sized from the distance of the last accept; d(anchor, k) does not depend on the batch it was computed in, so
 * the result equals the one-pair-at-a-time loop.  scale: 0.5 for the reference's half-resolution unit, 1 for
 * full-resolution pixels; gate: 2 in the reference.  searches / launches (may be NULL) count the work done.
 * Returns after the last record is on the host. */
typedef struct me_scan_record {
    int32_t anchor_slot;
    int32_t accepted;
    me_pair_result pair;
} me_scan_record;
int me_anchored_scan(me_ctx *ctx, int anchor_slot, int first_slot, int n_frames, float scale, float gate,
                     me_scan_record *out, int *searches, int *launches);

/* Hand the per-pair records of the last search to another GPU consumer without a host bounce: copies n_pairs
 * records from the context's result buffer to d_dst (device memory of the same GPU, e.g. the send buffer of an
 * NCCL all_gather -- the one exchange step of the sharded path) on `stream`, ordered after the search.
 * Asynchronous.  `stream` NULL means the context's compute stream, NOT the CUDA default stream (whose handle is also 0): a
 * consumer that runs on the default stream must pass a stream of its own here and order itself after it. */
int me_copy_pair_records_device(me_ctx *ctx, void *d_dst, int n_pairs, void *stream);

/* Integer-SAD issue-rate microbenchmark: runs a register-only VABSDIFF4 loop on every SM and
 * reports VABSDIFF4 thread-instructions per clock per SM and the SM clock it saw.  Used by bench.py
 * to put a measured number under the integer-ALU roofline.  mode 0 = SAD ops only (a register 4x4
 * patch, the search kernel's operand pattern); mode 1 adds one LDS.128 and 4 ALU ops per 16. */
int me_microbench_vabsdiff4(int device, int mode, double *instr_per_clk_per_sm, double *sm_mhz_seen,
                            int *sm_count);

/* Search with no allocation of any kind -- for callers that cannot allocate at the moment they need
 * a search: a plug-in inside a framework whose pool already owns the GPU's memory, a real-time loop
 * that must not trigger cudaMalloc's device-wide synchronisation, a worker with no memory budget of
 * its own.  (No counterpart in the reference, which never leaves the CPU; added for such hosts.)
 *
 * The caller passes what it already has: device frames (n_frames, height, pitch), 16-byte aligned,
 * pitch % 16 == 0; a HOST list of n_pairs (prev, cur) frame indices; a DEVICE array d_out of n_pairs
 * records; its stream (NULL = default stream).  The kernel's working memory (per-block records,
 * per-pair tickets) comes from a fixed ME_STATIC_SCRATCH_BYTES array inside the library, held by
 * every GPU that loads it until the process exits.  A request that needs more returns ME_EINVAL
 * (1080p with 16x16 blocks: 137 KB per pair, about 30 pairs per call).
 *
 * All calls share that array and it keeps the previous call's contents; what must be zero is
 * cleared on `stream` before each launch.  Calls whose kernels could overlap on the GPU race on it:
 * keep them on one stream or order them.  The first call on a device also performs that device's
 * one-time set-up (thread-safe) and should be made outside stream capture.  Asynchronous: d_out is
 * valid once `stream` has drained -- provided `pairs` is page-locked host memory that stays untouched
 * until then; a pageable `pairs` is staged by the CUDA runtime, which makes the call wait for the
 * stream first.  Neither cudaMalloc/cudaFree nor any allocating library routine is called. */
#define ME_STATIC_SCRATCH_BYTES (4u << 20)
int me_search_static_scratch(const uint8_t *d_frames, size_t pitch, int n_frames, int height, int width,
                             const me_params *params, const int32_t *pairs, int n_pairs,
                             me_pair_result *d_out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* ME_B200_H */
