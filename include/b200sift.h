/*
 * b200sift -- C ABI of the B200-native SIFT engine (sm_100a).
 *
 * This is the drop-in boundary for the hot path of Cryptum169/another-cuda-sift. Every entry
 * point names the reference interface it replaces (paths relative to the reference repo).
 * Plain pointers and sizes only: no C++ / thrust / torch types cross this line. The C++ mirror
 * of the reference API (sift_cuda::Detector, sift_cuda::matchBruteForce) in
 * include/sift_cuda/ is a header-only wrapper over these functions.
 *
 * Conventions: every function returns 0 on success or a negative B200SIFT_E* code;
 * b200sift_last_error() gives the message of the last failure on the calling thread.
 * A handle belongs to the CUDA device that was current when it was created and is not
 * thread-safe (same as the reference, Detector.hh:24). There is no CPU fallback: without a
 * CUDA device every call fails with B200SIFT_ECUDA.
 */
#ifndef B200SIFT_H
#define B200SIFT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200SIFT_API __attribute__((visibility("default")))
#else
#define B200SIFT_API
#endif

#define B200SIFT_OK 0
#define B200SIFT_EINVAL (-1) /* bad argument / unsupported configuration */
#define B200SIFT_ECUDA (-2)  /* CUDA runtime error, see b200sift_last_error() */
#define B200SIFT_ESTATE (-3) /* call order violated (e.g. results before any frame) */

/* Mirrors `struct CudaSiftConfig` (sift_cuda/types/CudaSiftConfig.hh:3-14), same field order,
 * names (including the reference's spellings) and defaults. `upscale` must be 0 unless
 * b200sift_options::working_upscale is set: the reference documents it as broken (CudaSiftConfig.hh:12). */
typedef struct b200sift_config {
    int col_width;
    int row_width;
    int numFeatures;           /* 5000 */
    int numOctaveLayers;       /* 3; 1..6 supported (runtime parameter, as in the reference: Detector.cu:63-94) */
    double contrastThreshould; /* 0.04 */
    double edgeThreshould;     /* 10 */
    double sigma;              /* 1.6; any value whose layer blurs need <= 85 taps (GaussianUtils.cc:39-67) */
    int upscale;               /* 0 */
} b200sift_config;

typedef struct b200sift_detector b200sift_detector; /* opaque handle */

/* Process environment: when the library is loaded it sets CUDA_DEVICE_MAX_CONNECTIONS=32 unless the variable is already set
 * (CUDA reads it when the context is created; the default of 8 hardware work queues lets only ~4 of the frames in flight overlap:
 * +7.5% end to end at 1920x1200, +9-11% at 752x480 with 32). Set the variable yourself to keep another value. */

/* Engine options that have no counterpart in CudaSiftConfig. All zero (b200sift_default_options sets frames_in_flight
 * to its default) = the reference's behaviour, bit for bit. The three fixes are the TODOs the reference itself lists;
 * they change results and are therefore opt-in. */
typedef struct b200sift_options {
    int frames_in_flight;        /* internal frame slots, 3..32 (0 = default 16): throughput vs ~85 MB of HBM per slot at 1920x1200 */
    int remove_duplicates;       /* interface/Detector.cu:220 "TODO: remove duplicate": drop keypoints whose (x, y, size, angle)
                                    equal an earlier keypoint of the same octave before description (first one kept) */
    int interpolate_orientation; /* sift_func/SiftOps.cu:359 "TODO: interpolate the peak": parabolic peak interpolation of the
                                    orientation histogram instead of angle = 360 - 10 * bin */
    int working_upscale;         /* types/CudaSiftConfig.hh:11-12 "broken if set to true": with cfg.upscale = 1 the frame is
                                    doubled (bilinear, the reference's resize) and keypoints are reported in input pixels with
                                    octave -1 conventions (position, size and packed octave consistent) */
    int reserved[4];
} b200sift_options;

B200SIFT_API void b200sift_default_config(b200sift_config* cfg);
B200SIFT_API const char* b200sift_last_error(void);
B200SIFT_API const char* b200sift_version(void);

/* Detector::Detector + Detector::gpuWarmUpAndAllocate (interface/Detector.hh:26-31,
 * interface/Detector.cu:17-131): allocates every device / pinned buffer once and captures the
 * per-frame CUDA graph. Fails with EINVAL when col_width or row_width is 0 (Detector.cu:22-25). */
B200SIFT_API int b200sift_create(const b200sift_config* cfg, b200sift_detector** out);
B200SIFT_API void b200sift_default_options(b200sift_options* opt);
/* b200sift_create with options (opt == NULL: defaults). */
B200SIFT_API int b200sift_create_ex(const b200sift_config* cfg, const b200sift_options* opt, b200sift_detector** out);
B200SIFT_API int b200sift_destroy(b200sift_detector* det);

/* Geometry the reference derives in Detector.hh:27 and Detector.cu:83,109. */
B200SIFT_API int b200sift_num_octaves(const b200sift_detector* det);
/* dims[0..3] = width, height, row pitch in floats, keypoint slot capacity of octave `o`. */
B200SIFT_API int b200sift_octave_info(const b200sift_detector* det, int o, int* dims);

/* Detector::detectAndCompute(const Imagef&) (interface/Detector.cu:133-233), host image:
 * row-major float32, rows x cols, values 0..255 (HostImage.hh:36). The H2D copy is enqueued on
 * the detector's stream from `img` (pinned memory makes it asynchronous) followed by the frame
 * graph. `_host` blocks until the frame is done and returns the keypoint count in *total
 * (Detector::total_size); `_host_async` only enqueues (img must stay valid until
 * b200sift_sync). */
B200SIFT_API int b200sift_detect_host(b200sift_detector* det, const float* img, int rows, int cols, int* total);
B200SIFT_API int b200sift_detect_host_async(b200sift_detector* det, const float* img, int rows, int cols);
/* Same frame, input already resident in HBM (`pitch_bytes` between rows); copied device to
 * device into the detector's input image on the frame's own stream. `producer_stream` (a cudaStream_t, NULL = the
 * legacy default stream) is the stream whose work produced `d_img`: the engine records an event there and makes its
 * copy wait for it -- its own streams are non-blocking and would otherwise not be ordered after the producer.
 * The image may be overwritten once b200sift_sync (or any later event on b200sift_stream) has completed. */
B200SIFT_API int b200sift_detect_device_async(b200sift_detector* det, const float* d_img, size_t pitch_bytes, void* producer_stream);
/* 8-bit input: converted to float on the device, 4x less PCIe traffic. */
B200SIFT_API int b200sift_detect_host_u8_async(b200sift_detector* det, const uint8_t* img, int rows, int cols);
/* Waits for the frames enqueued so far; *total = keypoints of the last frame. */
B200SIFT_API int b200sift_sync(b200sift_detector* det, int* total);
/* The CUDA stream (cudaStream_t) of the LATEST frame, so callers can order their own work after it. The
 * engine keeps several frames in flight, each on its own stream: fetch this after every detect call. */
B200SIFT_API void* b200sift_stream(b200sift_detector* det);

/* Public result members of Detector (interface/Detector.hh:54-62), borrowed device pointers
 * owned by the handle. They refer to the LATEST frame (prev_desc: the frame before it, in place -- the
 * reference copies it, Detector.cu:136-141) and rotate over the internal frame slots (b200sift_frames_in_flight): fetch them after every
 * detect call; a pointer stays valid until b200sift_frames_in_flight() - 1 further detect calls. kpts = float3[numFeatures]
 * (x, y, layer), feats = float4[numFeatures] (packed octave, size, |contrast|, angle deg),
 * desc / prev_desc = half[numFeatures*128]. Any out pointer may be NULL. */
B200SIFT_API int b200sift_results(b200sift_detector* det, const float** kpts, const float** feats,
                     const uint16_t** desc, const uint16_t** prev_desc, int* total);
/* Detector::copyToHost(bool descriptor) (interface/Detector.cu:606-634): first `total` entries
 * into caller buffers (any may be NULL); blocks until done. */
B200SIFT_API int b200sift_copy_to_host(b200sift_detector* det, float* kpts3, float* feats4, uint16_t* desc, int with_desc);

/* Test / profiling access to intermediate state (the reference exposes the same through
 * Detector::setDataGen dumps, Detector.cu:145-229). which: 0 Gaussian stack (6 layers).
 * Dense copy (no pitch) into host memory: layers*h*w floats. */
B200SIFT_API int b200sift_debug_get_stack(b200sift_detector* det, int o, int which, float* out);
/* which: 0 cand float3, 1 candF float4, 2 coll float3, 3 collF float4 (cap entries each). */
B200SIFT_API int b200sift_debug_get_kpt_buffer(b200sift_detector* det, int o, int which, void* out, int max_elems);
/* counts[0..2] = P raw extrema, M refined, T oriented (after the last frame). */
B200SIFT_API int b200sift_debug_get_counts(b200sift_detector* det, int o, int* counts);
/* Host-generated Gaussian taps: which = -1 initial blur, 0..5 layer taps. Returns tap count. */
B200SIFT_API int b200sift_debug_get_taps(b200sift_detector* det, int which, float* out, int max_elems);
/* Stage dump of the latest frame: what Detector::setDataGen(path) makes the reference write after every
 * detectAndCompute (Detector.cu:145-229, perf/PerfData.cu:4-85), as raw little-endian arrays instead of
 * msgpack+zlib. Writes into the existing directory `dir`: input.f32, oct{o}_gauss.f32 (6 x h x w),
 * oct{o}_refined_kpts.f32x3 / refined_features.f32x4 (M refined keypoints, raster order),
 * oct{o}_oriented_kpts.f32x3 / oriented_features.f32x4 (T oriented keypoints, (keypoint, bin) order),
 * final_kpts.f32x3 / final_features.f32x4 / final_descriptors.f16x128 and manifest.json
 * (config, octave dims, counts). tools/stage_replay.py re-runs the stages on input.f32 and compares. Blocks. */
B200SIFT_API int b200sift_dump_frame(b200sift_detector* det, const char* dir);
/* Runs the frame pipeline only up to and including `last_stage` (1 pyramid, 3 extrema,
 * 4 refine, 5 orientation, 7 all) eagerly, for stage-level parity tests. Blocks. */
B200SIFT_API int b200sift_debug_run_stages(b200sift_detector* det, const float* img, int rows, int cols, int last_stage);
/* The reference's public stage methods (createInitialImage ... collectAcrossOctaves, interface/Detector.hh:36-43) as a
 * compat shim: re-runs the pipeline on the LATEST frame's input up to and including `last_stage` (numbering as above).
 * Eager, blocking; ESTATE before the first frame. */
B200SIFT_API int b200sift_rerun_stages(b200sift_detector* det, int last_stage);
/* How many frames the engine keeps in flight (internal slots; det == NULL: the default). Buffers a caller hands to the
 * *_async calls of consecutive frames (match indices, host result buffers) should rotate over at least this many. */
B200SIFT_API int b200sift_frames_in_flight(const b200sift_detector* det);
/* Number of kernel launches (graph kernel nodes) one frame issues. */
B200SIFT_API int b200sift_launches_per_frame(const b200sift_detector* det);
/* Average device time (ms) of one frame with the input resident in HBM: `iters` graph launches
 * bracketed by CUDA events on the detector's stream after `warmup` launches. */
B200SIFT_API int b200sift_time_frames(b200sift_detector* det, int warmup, int iters, float* ms_per_frame);
/* Same with all frame slots in flight (one frame per internal slot, each on its own stream): ms per frame of the
 * pipelined engine, i.e. 1 / throughput. */
B200SIFT_API int b200sift_time_frames_pipelined(b200sift_detector* det, int warmup, int iters, float* ms_per_frame);
/* Development aid (no reference counterpart): the same pipelined replay with a timing event before part A and after parts A, B, C of each
 * of n_frames frames; ms_out[4*i + k] = milliseconds since the common start (tools/dev_timeline.py prints the steady-state schedule). */
B200SIFT_API int b200sift_debug_timeline(b200sift_detector* det, int n_frames, float* ms_out);

/* Pipelined (batched-frame) use, no host synchronisation per frame. After b200sift_detect_*_async:
 *  - b200sift_match_consecutive_async: matchBruteForce(prev_descriptor, previous total, device_descriptor, this
 *    total) exactly as tool/extract_and_match_example.cc:82 calls it, with both sizes read on the device;
 *    d_out_idx = int[numFeatures] on the device, written entirely by the engine on the latest frame's stream
 *    (entries beyond the previous frame's total are set to -1): the caller must not touch it on another stream
 *    until an event recorded on b200sift_stream after this call has completed;
 *  - b200sift_copy_results_async: all numFeatures slots of kpts / feats / desc (any may be NULL) and
 *    counts2[0..1] = {this frame's total, previous frame's total} to (pinned) host memory.
 * Both are enqueued on the latest frame's stream and are ordered before the detect call that reuses the
 * buffers they read (several frames are in flight); host frames of later b200sift_detect_host_async calls are
 * copied on a separate copy stream meanwhile. */
B200SIFT_API int b200sift_match_consecutive_async(b200sift_detector* det, int* d_out_idx);
B200SIFT_API int b200sift_copy_results_async(b200sift_detector* det, float* kpts3, float* feats4, uint16_t* desc, int* counts2);
/* The loop of tool/extract_and_match_example.cc:57-101 over a batch of n host frames in ONE call, pipelined, with a
 * single synchronisation at the end: for every frame b200sift_detect_host_async (or the 8-bit variant when is_u8),
 * then, where match_idx && match_idx[i], b200sift_match_consecutive_async followed by a copy of the numFeatures
 * indices to match_idx[i] (the match of frame i-1 -> frame i), then b200sift_copy_results_async into kpts3[i] /
 * feats4[i] / desc[i] / counts2[i] (any of the four arrays, or single entries, may be NULL). Host buffers should be
 * pinned for the copies to overlap. Blocks until the whole batch is done. */
B200SIFT_API int b200sift_detect_batch_host(b200sift_detector* det, const void* const* frames, int n, int rows, int cols, int is_u8,
                                            float* const* kpts3, float* const* feats4, uint16_t* const* desc, int* const* counts2,
                                            int* const* match_idx);

/* Per-kernel-class device time of one frame (ms, averaged over `iters` eager frames with a CUDA event
 * after every launch on the detector's stream): ms_per_class[B200SIFT_K_COUNT]; launches_per_class (may
 * be NULL) = launches (event intervals) charged to each class per frame. bench.py's roofline figures
 * come from here. */
#define B200SIFT_K_COPY 0       /* prev-descriptor copy, count memset, count read-back */
#define B200SIFT_K_BLUR 1       /* fused V+H Gaussian blur, octave 0 (the full-resolution layers) */
#define B200SIFT_K_BLUR_SMALL 2 /* the same kernel on octaves >= 1 */
#define B200SIFT_K_DOWNSAMPLE 3
#define B200SIFT_K_DETECT 4     /* DoG + 3x3x3 extrema bitmask */
#define B200SIFT_K_COMPACT 5    /* row scan / scatter / refined + oriented compaction */
#define B200SIFT_K_REFINE 6
#define B200SIFT_K_ORIENT 7
#define B200SIFT_K_DESCRIBE 8
#define B200SIFT_K_COUNT 9
B200SIFT_API int b200sift_profile_frame(b200sift_detector* det, int warmup, int iters, float* ms_per_class,
                           int* launches_per_class);

/* Cumulative in-graph times (ms): the frame graph truncated after the pyramid, extrema, refinement, orientation
 * and complete, each replayed `iters` times on the last frame's input: ms5[0..4]. */
B200SIFT_API int b200sift_time_stages(b200sift_detector* det, int iters, float* ms5);

/* sift_cuda::matchBruteForce (sift_func/Match.cuh:9-14, Match.cu:8-33): for each of the nq query
 * descriptors (half[nq*128]) the index of its nearest train descriptor (half[nt*128]) under the
 * reference's half-precision L2 recipe, or -1 if the 0.8 ratio test fails. Device pointers;
 * out_idx = int[nq] on the device. `stream` is a cudaStream_t (NULL = legacy default stream).
 * Asynchronous. Scratch memory is ONE cached buffer per device, grown on demand: calls on different streams are
 * serialised through an event (a call waits for the previous call's kernels, whatever stream they ran on), and a call
 * that has to grow the buffer synchronises the device first -- so the first call of a size class must not be made
 * during stream capture. One host thread per device may call concurrently with threads of other devices. */
B200SIFT_API int b200sift_match_device(const uint16_t* d_query, int nq, const uint16_t* d_train, int nt,
                          int* d_out_idx, void* stream);
/* Host-buffer convenience (copies in, matches, copies out, blocks). */
B200SIFT_API int b200sift_match_host(const uint16_t* query, int nq, const uint16_t* train, int nt, int* out_idx);
/* No counterpart in the reference. b200sift_match_device for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): same matcher, same
 * arguments, bit-identical result, but its scratch (survivor records + squared norms) lives in a file-scope
 * `static __device__` array of the library (csrc/noalloc.cu, ~8.1 MiB, held by every GPU that loads the library until the
 * process exits) instead of the cached cudaMalloc'd buffer. The call path allocates nothing: no cudaMalloc / cudaFree /
 * allocating library routine. Limits and rules:
 *   - nq <= B200SIFT_NOALLOC_MAX_ROWS and nt <= B200SIFT_NOALLOC_MAX_ROWS, else B200SIFT_EINVAL with a message naming the limit;
 *   - every call on a device shares the one array, which keeps the previous call's contents (nothing in it has to start
 *     at zero: a call writes every norm and survivor record before reading it); calls that OVERLAP on the GPU (different
 *     streams without ordering) race on it and return garbage -- order them on one stream or with events;
 *   - asynchronous on `stream` like b200sift_match_device; inputs and output are the caller's device buffers. */
#define B200SIFT_NOALLOC_MAX_ROWS 8192
B200SIFT_API int b200sift_match_device_noalloc(const uint16_t* d_query, int nq, const uint16_t* d_train, int nt,
                                               int* d_out_idx, void* stream);
/* B200SIFT_NOALLOC_MAX_ROWS as the loaded library was built with. */
B200SIFT_API int b200sift_match_noalloc_max_rows(void);
/* Kernel launches issued by one b200sift_match_device call for this problem size. */
B200SIFT_API int b200sift_match_launches(int nq, int nt);

/* ---- Query-row sharded matching over the GPUs of one box (BASELINE.json "100k x 100k ... query-row
 * sharded across 8xB200 with NCCL gather"). The sharded semantic is sift_cuda::matchBruteForce (sift_func/Match.cu:8-33):
 * the train set is replicated, rank r owns the query rows [r*per, min(nq, (r+1)*per)), per = ceil(nq / world).
 * One process (or thread) per GPU; NCCL is loaded at run time (libnccl.so.2), only by these calls.
 *   b200sift_nccl_unique_id   ncclGetUniqueId on one rank; the caller ships the 128 bytes to the other ranks
 *                             (torch.distributed / MPI / a file -- rendezvous is the host's business)
 *   b200sift_comm_create      ncclCommInitRank on the CURRENT device; collective over all ranks
 *   b200sift_shard_rows       the row block of `rank` ([*begin, *end)); returns per
 *   b200sift_match_sharded    enqueues on `stream`, with no host synchronisation in between:
 *                               ncclBroadcast of d_train (nt x 128 halves, in place) from rank train_root
 *                                 (train_root < 0: every rank already holds the train set),
 *                               the matcher on this rank's rows (d_query_block = its first row),
 *                               ncclAllGather of `per` int32 per rank into d_out_all (world * per ints on every rank;
 *                                 the first nq_total are matchBruteForce's result, the padding is -1).
 *                             Collective: every rank must call it with the same nq_total / nt / train_root. */
typedef struct b200sift_comm b200sift_comm;
B200SIFT_API int b200sift_nccl_unique_id(uint8_t* id128);
B200SIFT_API int b200sift_comm_create(const uint8_t* id128, int rank, int world, b200sift_comm** out);
B200SIFT_API int b200sift_comm_destroy(b200sift_comm* comm);
B200SIFT_API int b200sift_comm_info(const b200sift_comm* comm, int* rank, int* world, int* nccl_version);
B200SIFT_API int b200sift_shard_rows(int n_rows, int world, int rank, int* begin, int* end);
B200SIFT_API int b200sift_match_sharded(b200sift_comm* comm, const uint16_t* d_query_block, int nq_total, uint16_t* d_train, int nt,
                                        int train_root, int* d_out_all, void* stream);
/* The same call with a CUDA event between its stages; blocks, ms5 = {broadcast, norms (+ |b|^2 chunks), tcgen05 filter,
 * exact re-rank, all-gather} in milliseconds on this rank (bench.py's per-stage breakdown). */
B200SIFT_API int b200sift_match_sharded_profile(b200sift_comm* comm, const uint16_t* d_query_block, int nq_total, uint16_t* d_train,
                                                int nt, int train_root, int* d_out_all, void* stream, float* ms5);

/* Matcher implementation: 0 (default) = tcgen05 tensor-core filter + exact re-rank (csrc/match_tc.cu),
 * 1 = all-CUDA-core exact kernel (csrc/match.cu), kept as the on-device cross-check, 2 = as 0 but train sets
 * larger than 8192 rows go through the first, unpipelined tcgen05 filter (cross-check of the streaming one), 3 = as 0 with
 * the single-pass variant of the streaming filter (exact, measured slower: A/B only).
 * All are bit-exact with the reference; process-wide setting. */
B200SIFT_API int b200sift_match_set_path(int path);
/* Test access to the tensor-core stage: d_dist[nq*nt] (device) receives the fp32 squared distances
 * |a|^2 + |b|^2 - 2ab^T the filter ranks by; d_out_idx[nq] the final result. Blocks. */
B200SIFT_API int b200sift_debug_match_dist(const uint16_t* d_query, int nq, const uint16_t* d_train, int nt,
                              float* d_dist, int* d_out_idx);

#ifdef __cplusplus
}
#endif
#endif /* B200SIFT_H */
