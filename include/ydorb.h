/*
 * ydorb.h -- C ABI of libydorb.so: the B200 (sm_100a) implementation of YDORBSLAM's ORB front-end hot path.
 *
 * The reference has no FFI layer; its boundary is the C++ class surface of
 *   YDORBSLAM::OrbExtractor   (src/orbExtractor.hpp:31-73)
 *   YDORBSLAM::OrbMatcher     (src/orbMatcher.hpp:24-66)
 *   YDORBSLAM::Frame hooks    (src/frame.hpp:83,96,166: computeStereoMatches, getKeyPointsInArea,
 *                              assignKeyPointsToGrid)
 * Each entry point below names the reference member it replaces.  INTEGRATION.md shows the C++ shim a
 * maintainer adds on the reference side (include/ydorb.hpp is that shim without OpenCV types; tests/cpp/shim_check.cpp
 * drives it and tests/test_cpp_shim.py checks its outputs against the oracle).
 *
 * Conventions: plain pointers and sizes, no exceptions, int status return (0 = ok, <0 = YDORB_E_*),
 * caller-owned output buffers.  "host" entry points take host memory and synchronise before returning;
 * "_device" entry points take device memory, enqueue on the handle's stream and do NOT synchronise.
 * There is no CPU fallback: every entry point fails with YDORB_E_CUDA if no sm_100 device is usable.
 * Handles may be used from different host threads, one thread at a time per handle (the reference runs the
 * left and right extractor concurrently on two instances, src/frame.cpp:84-87).
 */
#ifndef YDORB_H
#define YDORB_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YDORB_OK 0
#define YDORB_E_INVALID (-1)  /* bad argument */
#define YDORB_E_CUDA (-2)     /* CUDA runtime error / no device */
#define YDORB_E_CAPACITY (-3) /* an internal or caller buffer is too small */
#define YDORB_E_STATE (-4)    /* call order (e.g. pyramid requested before any extract) */

#define YDORB_MAX_LEVELS 16

/* cv::KeyPoint (28 bytes): pt.x, pt.y, size, angle [deg], response, octave, class_id */
typedef struct {
  float x, y, size, angle, response;
  int32_t octave, class_id;
} YdKeyPoint;

/* OrbExtractor ctor arguments, src/orbExtractor.hpp:33-34.  NOTE the reference ctor ignores minThFAST and uses
 * iniThFAST for the fallback (src/orbExtractor.cpp:318); set reference_min_th_quirk=1 to reproduce that. */
typedef struct {
  int32_t n_features;
  float scale_factor;
  int32_t n_levels;
  int32_t ini_th_fast;
  int32_t min_th_fast;
  int32_t reference_min_th_quirk;
} YdExtractorParams;

typedef struct YdExtractor YdExtractor;

const char* ydorb_last_error(void); /* thread-local message of the last failing call */
int ydorb_device_count(void);
/* number of CUDA kernels this library has launched in this process (evidence for benchmarks) */
long long ydorb_kernel_launches(void);
/* Page-locked host memory for the arrays handed to the host-pointer entry points (frames, key points, descriptors).
 * Every entry point accepts ordinary pageable memory (a cv::Mat's own buffer); buffers from ydorb_host_alloc are copied
 * by DMA without the driver's staging pass, which is what bounds large batches (64 x 1080p: 3.5 k -> 11.3 k frames/s).
 * A cv::Mat header can wrap such a buffer: cv::Mat(rows, cols, CV_8UC1, ptr, step). */
int ydorb_host_alloc(size_t bytes, void** out);
void ydorb_host_free(void* p);

/* Diagnostics (no reference counterpart): measured issue rates of the integer pipes, thread-operations per second:
 * out[0] POPC, [1] LOP3, [2] VIMNMX3.S16x2 (DPX), [3] IMAD, [4] LOP3+IMAD alternating.  Denominators of the integer
 * roofline bench.py reports for the FAST and Hamming kernels. */
int ydorb_diag_int_peaks(int device, double* ops_per_s, int n);
/* Diagnostics: the device's sinf/cosf of computeDescriptors (src/orbExtractor.cpp:426 resolves to libm sincosf) for the
 * `count` consecutive floats whose bit patterns start at first_bits; out[2*i] = sin, out[2*i+1] = cos.  Lets a test
 * compare the kernel's trigonometry with the host libm over the whole input domain. */
int ydorb_diag_sincosf(int device, uint32_t first_bits, uint32_t count, float* out_sin_cos);

/* ---- OrbExtractor -------------------------------------------------------------------------------------- */
/* OrbExtractor::OrbExtractor (src/orbExtractor.cpp:315-354).  max_* size the device buffers. */
int ydorb_extractor_create(const YdExtractorParams* params, int max_width, int max_height, int max_batch, int device,
                           YdExtractor** out);
void ydorb_extractor_destroy(YdExtractor* ex);
/* getLevelsNum / getScaleFactor(s) / getInvScaleFactors / getScaleFactorSquares / getInvScaleFactorSquares
 * (src/orbExtractor.hpp:42-48); each output array has n_levels entries, any may be NULL. */
int ydorb_extractor_levels(const YdExtractor* ex);
int ydorb_extractor_scale_tables(const YdExtractor* ex, float* sf, float* inv_sf, float* sf2, float* inv_sf2);
int ydorb_extractor_quotas(const YdExtractor* ex, int32_t* per_level);
/* CUDA stream (cudaStream_t) the handle enqueues on; lets callers time with events on that stream. */
void* ydorb_extractor_stream(const YdExtractor* ex);

/* OrbExtractor::extractAndCompute (src/orbExtractor.cpp:355-399): one CV_8UC1 image in; keypoints (level-major)
 * and an N x 32 descriptor matrix out.  Fresh-extractor semantics (pyramid staleness not
 * reproduced).  *n_out = keypoints written (<= capacity; YDORB_E_CAPACITY if more were produced). */
int ydorb_extract(YdExtractor* ex, const uint8_t* image, int width, int height, int stride, YdKeyPoint* keypoints,
                  uint8_t* descriptors, int capacity, int32_t* n_out);
/* batch of same-sized frames: frame b at images + b*frame_stride; outputs at keypoints + b*capacity,
 * descriptors + b*capacity*32; n_out[b]. */
int ydorb_extract_batch(YdExtractor* ex, const uint8_t* images, int width, int height, int stride, size_t frame_stride,
                        int batch, YdKeyPoint* keypoints, uint8_t* descriptors, int capacity, int32_t* n_out);
/* same, all pointers in device memory, asynchronous on the handle's stream */
int ydorb_extract_batch_device(YdExtractor* ex, const uint8_t* d_images, int width, int height, int stride,
                               size_t frame_stride, int batch, YdKeyPoint* d_keypoints, uint8_t* d_descriptors,
                               int capacity, int32_t* d_n_out);
int ydorb_extractor_sync(YdExtractor* ex);
/* enqueue on a caller-owned cudaStream_t instead of the handle's own stream (NULL restores it) */
int ydorb_extractor_set_stream(YdExtractor* ex, void* cuda_stream);
/* per-stage CUDA-event timing of the last extract call: ms[0..3] = pyramid, FAST, quadtree, describe */
int ydorb_extractor_set_profiling(YdExtractor* ex, int enable);
int ydorb_extractor_stage_ms(YdExtractor* ex, float* ms4);

/* m_v_imagePyramid (src/orbExtractor.hpp:50): padded level of the last extract call, device memory.
 * *d_level points at the level's (0,0) pixel; 19 px of REFLECT_101 border are readable around it. */
int ydorb_pyramid_level(const YdExtractor* ex, int frame, int level, const uint8_t** d_level, int* width, int* height,
                        int* stride);
/* copy a level to host; padded!=0 copies (w+38)x(h+38) including the border */
int ydorb_pyramid_download(YdExtractor* ex, int frame, int level, int padded, uint8_t* dst, int dst_stride);

/* Stage outputs of the last extract call, for parity tests (computeKeyPointsPyramid, src/orbExtractor.cpp:545-604):
 * FAST candidates of one level in reference order (cells row-major, row-major inside a cell), coordinates relative
 * to the 16 px border as distributeQuadTree sees them.  xy[2*i], xy[2*i+1]; resp[i]. */
int ydorb_debug_candidates(YdExtractor* ex, int frame, int level, int32_t* xy, float* resp, int capacity, int32_t* n_out);

/* ---- OrbMatcher ---------------------------------------------------------------------------------------- */
typedef struct YdMatcher YdMatcher;

/* per-query result of a top-2 scan (rule of src/orbMatcher.cpp:319-334: strict <, else-if strict <) */
typedef struct {
  int32_t best_idx;    /* -1 if no candidate */
  int32_t best_dist;   /* 256 if none */
  int32_t second_dist; /* 256 if none */
  int32_t second_idx;  /* -1 if none */
} YdTop2;

int ydorb_matcher_create(int device, YdMatcher** out);
void ydorb_matcher_destroy(YdMatcher* m);
void* ydorb_matcher_stream(const YdMatcher* m);
int ydorb_matcher_sync(YdMatcher* m);
int ydorb_matcher_set_stream(YdMatcher* m, void* cuda_stream);

/* OrbMatcher::computeDescriptorsDistance (src/orbMatcher.cpp:11-23), n pairs: dist[i] = |a_i xor b_i| */
int ydorb_hamming_pairs(YdMatcher* m, const uint8_t* a, const uint8_t* b, int n, int32_t* dist);
/* dense top-2 of every query against a descriptor database shard (inner loops of searchByBow*,
 * src/orbMatcher.cpp:319-334, without the greedy exclusion); idx_base is added to the indices (shard offset) */
int ydorb_hamming_top2(YdMatcher* m, const uint8_t* queries, int nq, const uint8_t* db, int ndb, int idx_base,
                       YdTop2* out);
int ydorb_hamming_top2_device(YdMatcher* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, int ndb, int idx_base,
                              YdTop2* d_out);
/* The same scan as ydorb_hamming_top2_device for callers that may not allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The per-query fold state lives in a file-scope device array of the library (~256 KB, held by
 * every GPU that loads the library until exit) instead of the handle's grow-only cudaMalloc'ed buffer: the call
 * reaches no cudaMalloc / cudaFree.  nq > YDORB_NOALLOC_MAX_QUERIES fails with YDORB_E_CAPACITY (split the queries;
 * ndb is not limited by it).  Same results as ydorb_hamming_top2_device, asynchronous on the matcher's stream.  All
 * calls and handles on a device share the array: calls that overlap on the GPU race -- keep them on one stream. */
#define YDORB_NOALLOC_MAX_QUERIES 16384
int ydorb_hamming_top2_device_noalloc(YdMatcher* m, const uint8_t* d_queries, int nq, const uint8_t* d_db, int ndb,
                                      int idx_base, YdTop2* d_out);
/* batched dense top-2 between frames of one descriptor block d_desc[frame][capacity][32] with per-frame counts
 * d_n[frame]: pair p matches the descriptors of frame d_q_frame[p] against those of frame d_db_frame[p] (e.g. frame
 * t against frame t-1).  d_out[p*capacity + i]; d_accept (may be NULL) gets the ratio test of src/orbMatcher.cpp:336. */
int ydorb_hamming_top2_frames_device(YdMatcher* m, const uint8_t* d_desc, const int32_t* d_n, int capacity,
                                     const int32_t* d_q_frame, const int32_t* d_db_frame, int npairs, YdTop2* d_out,
                                     uint8_t* d_accept, float ratio, int th);
/* same on the descriptors left on the device by the last ydorb_extract_batch (host API) of `ex`; host in/out.
 * out[p*capacity + i] with capacity = sum of the per-level quotas of `ex` (ydorb_extractor_quotas).  Only the first
 * n_out[q_frame[p]] rows of a pair are results; the rows after them come back as {-1, -1, -1, -1} with accept = 0. */
int ydorb_match_frames(YdMatcher* m, const YdExtractor* ex, const int32_t* q_frame, const int32_t* db_frame, int npairs,
                       float ratio, int th, YdTop2* out, uint8_t* accept);
/* merge per-shard results (shards[s*nq + i]) into the result of one scan over the whole database */
int ydorb_top2_merge_device(YdMatcher* m, const YdTop2* d_shards, int nshards, int nq, YdTop2* d_out);
/* accept[i] = best<=th && (float)best < ratio*(float)second   (src/orbMatcher.cpp:336) */
int ydorb_ratio_accept_device(YdMatcher* m, const YdTop2* d_top2, int nq, float ratio, int th, uint8_t* d_accept);

/* ---- multi-GPU: sharded Hamming top-2 (one process per GPU; SURVEY 8e, BASELINE config 4) ------------------------
 * The map-descriptor database is split into contiguous row shards; every rank scans its shard and the per-shard
 * records are merged into the result of ONE sequential scan over the whole database (rule of
 * src/orbMatcher.cpp:327-336: strict <, else-if strict <, lowest row wins ties, then the ratio test).  The result is
 * identical on every rank.  The library creates its own NCCL communicator from an ncclUniqueId the caller ships from
 * rank 0 to the other ranks by any host channel (a file, MPI, torch.distributed ...).  Exchange of the 16 B/query
 * records: peer-memory stores inside the scan kernel when the GPUs can map each other's memory (CUDA IPC over
 * NVLink), otherwise an NCCL all-gather; YDORB_SHARD_EXCHANGE=nccl forces the latter.
 * Every rank must make the same sequence of sharded calls with the same nq (they are collective). */
#define YDORB_COMM_ID_BYTES 128
typedef struct YdComm YdComm;
typedef struct YdShard YdShard;
/* rank 0: obtain the id (ncclGetUniqueId) */
int ydorb_comm_unique_id(uint8_t* id, int bytes);
/* all ranks, collectively; the matcher's device and stream are the communicator's.  id may be NULL when world_size == 1 */
int ydorb_comm_create(YdMatcher* m, const uint8_t* id, int rank, int world_size, int max_queries, YdComm** out);
void ydorb_comm_destroy(YdComm* c);
int ydorb_comm_rank(const YdComm* c);
int ydorb_comm_size(const YdComm* c);
int ydorb_comm_exchange_mode(const YdComm* c); /* 0 single GPU, 1 NCCL all-gather, 2 peer-memory push */
/* synchronises the matcher's stream and reports a sharded call whose peers did not arrive (YDORB_E_STATE) */
int ydorb_comm_check(YdComm* c);
/* this rank's rows [idx_base, idx_base + n_rows) of the database, uploaded once and kept resident */
int ydorb_shard_create(YdMatcher* m, const uint8_t* rows, int n_rows, int idx_base, YdShard** out);
void ydorb_shard_destroy(YdShard* s);
/* host queries in, merged top-2 (global row indices) and ratio-test flags out (accept may be NULL) */
int ydorb_hamming_top2_sharded(YdComm* c, const YdShard* shard, const uint8_t* queries, int nq, float ratio, int th,
                               YdTop2* out, uint8_t* accept);
/* everything resident; asynchronous on the matcher's stream (call ydorb_comm_check to synchronise and check) */
int ydorb_hamming_top2_sharded_device(YdComm* c, const uint8_t* d_queries, int nq, const uint8_t* d_db_shard, int n_rows,
                                      int idx_base, float ratio, int th, YdTop2* d_out, uint8_t* d_accept);

/* ---- Frame hooks + projection search -------------------------------------------------------------------- */
typedef struct YdFrame YdFrame;

typedef struct {
  float min_x, min_y, max_x, max_y; /* image bounds, src/frame.cpp:223-248 */
  int32_t n_levels;
  float scale_factors[YDORB_MAX_LEVELS];
} YdFrameGeom;

/* one local map point as Frame::isInCameraFrustum left it (src/frame.cpp:295-326) */
typedef struct {
  float proj_x, proj_y, proj_right_x, view_cos;
  int32_t level;  /* m_int_trackScaleLevel */
  int32_t active; /* m_b_isTrackInView && !isBad() */
} YdMapPointProj;

/* Frame feature state on the device + Frame::assignKeyPointsToGrid (src/frame.cpp:249-264).
 * right_x may be NULL (all -1). Host pointers. */
int ydorb_frame_create(YdMatcher* m, const YdFrameGeom* geom, const YdKeyPoint* keypoints, const uint8_t* descriptors,
                       const float* right_x, int n, YdFrame** out);
void ydorb_frame_destroy(YdFrame* f);
/* Frame::getKeyPointsInArea (src/frame.cpp:337-361), reference predicate and order; *n_out may exceed capacity */
int ydorb_frame_keypoints_in_area(YdFrame* f, float x, float y, float r, int min_level, int max_level, int32_t* out,
                                  int capacity, int32_t* n_out);
/* OrbMatcher::searchByProjectionInFrameAndMapPoint (src/orbMatcher.cpp:24-64).  kp_occupied[i]!=0: slot i holds a
 * map point with observations>0 before the call; mp_obs_positive[j]: map point j has observations>0.
 * match_out[i] = map point index assigned to keypoint i by this call or -1; *n_matches = return value of the
 * reference function. Host pointers. */
int ydorb_search_by_projection(YdFrame* f, const YdMapPointProj* map_points, const uint8_t* mp_descriptors,
                               const uint8_t* mp_obs_positive, int n_map_points, const uint8_t* kp_occupied, float th,
                               float ratio, int32_t* match_out, int32_t* n_matches);
/* Same search with the map point arrays, kp_occupied and match_out resident in device memory (a local map kept on the
 * GPU between frames); mp_descriptors 16-byte aligned.  *n_matches is a host int; returns after the matcher's stream
 * has drained, match_out is complete by then. */
int ydorb_search_by_projection_device(YdFrame* f, const YdMapPointProj* map_points, const uint8_t* mp_descriptors,
                                      const uint8_t* mp_obs_positive, int n_map_points, const uint8_t* kp_occupied,
                                      float th, float ratio, int32_t* match_out, int32_t* n_matches);

/* ---- Projection searches after M1 ------------------------------------------------------------------------
 * OrbMatcher::searchByProjectionInLastAndCurrentFrame (src/orbMatcher.cpp:65-155), ...InKeyFrameAndCurrentFrame
 * (:156-239), searchByProjectionInSim (:240-302), both directions of searchBySim3 (:566-681), and the candidate
 * search of fuseByProjection (:682-745) / fuseBySim3 (:746-807) share one loop: project, getKeyPointsInArea,
 * best-1 over the window, threshold, greedy claim, rotation histogram (:120-153, computeThreeMaxima :827-854).
 * The projection (the reference's own cv::Mat arithmetic, e.g. :84-93) stays in the caller's shim, which fills one
 * YdWindowQuery per map point; the library runs everything from the getKeyPointsInArea call on. */
#define YDORB_Q_ACTIVE 1 /* all gates before `!vIndices.empty()` passed (z>=0, isInImage, distance range, normal) */
#define YDORB_Q_BLOCKS 2 /* a claim by this query blocks later queries (observations>0 in :106-107; always in :194, :284) */
typedef struct {
  float x, y;                   /* projection in the searched frame */
  float r;                      /* radius handed to getKeyPointsInArea */
  int32_t min_level, max_level; /* level arguments of getKeyPointsInArea, -1 = defaulted */
  int32_t oct_lo, oct_hi;       /* caller-side gate oct_lo <= octave <= oct_hi (:285-286, :616, :716-717, :782) */
  float right_x;                /* projected right-image x (stereo gate :108-109, chi2 :715) */
  float right_r;                /* stereo gate radius; < 0: no stereo gate */
  float angle;                  /* angle of the source keypoint (rotation histogram, :121, :206) */
  int32_t flags;                /* YDORB_Q_* */
} YdWindowQuery;

typedef struct {
  int32_t max_dist;          /* m_int_highThd / m_int_lowThd / _orbDist */
  int32_t strict;            /* 1: best < max_dist (:117); 0: best <= max_dist */
  int32_t exclusive;         /* 1: keypoints held before the call (kp_occupied) or claimed by an earlier blocking query are skipped */
  int32_t check_orientation; /* m_b_isToCheckOrientation */
  int32_t chi2_gate;         /* 1: fuseByProjection's 5.99 / 7.81 reprojection gates (:713-719) */
  float inv_sf2[YDORB_MAX_LEVELS]; /* m_v_invScaleFactorSquares, read by the chi2 gate only */
} YdWindowParams;

/* best_of[j] (may be NULL) = keypoint chosen by query j or -1.  match_out[i] (may be NULL) = last accepted query that
 * claimed keypoint i; -1 = untouched; -2 = cleared by the orientation check.  *n_matches = the reference function's
 * return value.  Host pointers. */
int ydorb_search_windows(YdFrame* f, const YdWindowQuery* queries, const uint8_t* q_descriptors, int nq,
                         const YdWindowParams* prm, const uint8_t* kp_occupied, int32_t* best_of, int32_t* match_out,
                         int32_t* n_matches);

/* ---- EXTRA: searchForInitialization ------------------------------------------------------------------------
 * `north_star` names it; the reference has no such function (absent from src/orbMatcher.hpp:24-66).  This is
 * ORB-SLAM2's published ORBmatcher::SearchForInitialization on this reference's own pieces: getKeyPointsInArea
 * (src/frame.cpp:337-361), the descriptor distance (src/orbMatcher.cpp:11-23), the rotation histogram of :120-131 and
 * computeThreeMaxima (:827-854).  NO reference parity claim; parity is against the sequential restatement in oracle/.
 *   f2            frame searched (its grid and descriptors are resident)
 *   kps1 / desc1  keypoints and 32-byte descriptors of the first frame (host); only octave-0 keypoints are matched
 *   prev_matched  [n1][2] floats, in: window centres (vbPrevMatched), out: position of the match where one was made
 *   window_size   radius handed to getKeyPointsInArea; nn_ratio / th_low: best < nn_ratio * second, best <= th_low
 *   area_mode     0: the reference's window predicate; 1: ORB-SLAM2's |dx| < r && |dy| < r, octave == 0
 *   matches12     [n1] matched keypoint of f2 or -1; *n_matches = the function's return value */
int ydorb_match_windowed_top2(YdFrame* f2, const YdKeyPoint* kps1, const uint8_t* desc1, int n1, float* prev_matched,
                              int window_size, float nn_ratio, int th_low, int check_orientation, int area_mode,
                              int32_t* matches12, int32_t* n_matches);

/* ---- BoW matchers ---------------------------------------------------------------------------------------
 * DBoW3::FeatureVector (std::map<NodeId, std::vector<unsigned>>, thirdParty/DBow3/src/FeatureVector.h:19-21) in CSR
 * form: node_ids ascending (= std::map order), node_start[n_nodes+1], items in push_back order.  Every feature index
 * appears in at most one node (DBoW3 assigns each feature to one node at `levelsup`). */
typedef struct {
  const int32_t* node_ids;
  const int32_t* node_start;
  const int32_t* items;
  int32_t n_nodes;
} YdFeatureVector;

/* mode 0: OrbMatcher::searchByBowInKeyFrameAndFrame (src/orbMatcher.cpp:303-379): side 1 = key frame (valid1[i] =
 *         matched map point present and not bad), side 2 = frame; match_out[n2]: frame index -> key-frame index or -1.
 * mode 1: OrbMatcher::searchByBowInTwoKeyFrames (:380-462): valid2 required too; match_out[n1]: first -> second or -1.
 * Host pointers. */
int ydorb_search_by_bow(YdMatcher* m, const YdKeyPoint* kps1, const uint8_t* desc1, const uint8_t* valid1, int n1,
                        const YdFeatureVector* fv1, const YdKeyPoint* kps2, const uint8_t* desc2, const uint8_t* valid2,
                        int n2, const YdFeatureVector* fv2, int mode, float ratio, int check_orientation,
                        int32_t* match_out, int32_t* n_matches);

/* OrbMatcher::searchForTriangulation (src/orbMatcher.cpp:463-565) with isEpipolarLineDistCorrect (:808-819).
 * has_mp*[i]: getMapPoint(i) != null.  F: _fMatrix_first2second, row-major 3x3.  epipole: (xCord, yCord) of :469-470.
 * sf / sf2: second key frame's m_v_scaleFactors / m_v_scaleFactorSquares (n_levels entries).
 * match_out[n1]: first index -> second index or -1 (the reference emits them as pairs in ascending first index). */
int ydorb_search_for_triangulation(YdMatcher* m, const YdKeyPoint* kps1, const uint8_t* desc1, const float* right_x1,
                                   const uint8_t* has_mp1, int n1, const YdFeatureVector* fv1, const YdKeyPoint* kps2,
                                   const uint8_t* desc2, const float* right_x2, const uint8_t* has_mp2, int n2,
                                   const YdFeatureVector* fv2, const float* F, float epipole_x, float epipole_y,
                                   const float* sf, const float* sf2, int n_levels, int stereo_only,
                                   int check_orientation, int32_t* match_out, int32_t* n_matches);

/* OrbMatcher::computeThreeMaxima (src/orbMatcher.cpp:827-854) on the bin sizes; host helper, no device work. */
void ydorb_three_maxima(const int32_t* sizes, int len, int32_t* idx1, int32_t* idx2, int32_t* idx3);

/* Frame::computeStereoMatches (src/frame.cpp:362-472) on the pyramids held by two extractor handles
 * (frame index = position in their last batch).  mode 0 = reference (leftIdx lag etc.),
 * 1 = corrected (descriptor row and output slot follow the keypoint).  max_d = bf/baseline (+inf on the
 * reference's first frame).  Host pointers; right_x/depth have n_left entries. */
int ydorb_stereo_match(YdMatcher* m, const YdExtractor* left, int left_frame, const YdExtractor* right, int right_frame,
                       const YdKeyPoint* kp_left, const uint8_t* desc_left, int n_left, const YdKeyPoint* kp_right,
                       const uint8_t* desc_right, int n_right, float bf, float max_d, int mode, float* right_x,
                       float* depth);

/* ---- Map-point side of the front-end (SURVEY 8f "next" rows) --------------------------------------------------- */
/* MapPoint::computeDistinctiveDescriptors (src/mapPoint.cpp:169-218), batched over map points.  Point p owns the
 * descriptor rows [start[p], start[p+1]) of desc (what :183-187 collects from the observing key frames).  best[p] =
 * relative row whose sorted distance vector is least at index int(0.5*n) (first minimum wins, :203-210); -1 if the
 * point has no rows.  Host pointers; the _device variant takes device pointers and runs on the matcher's stream. */
int ydorb_distinctive_descriptors(YdMatcher* m, const uint8_t* desc, const int32_t* start, int n_points, int32_t* best);
int ydorb_distinctive_descriptors_device(YdMatcher* m, const uint8_t* d_desc, const int32_t* d_start, int n_points,
                                         int32_t* d_best);

/* Camera state Frame::isInCameraFrustum reads (src/frame.cpp:295-326; members set at :273-290 and :60-130) */
typedef struct {
  float R[9];             /* m_cvMat_R_c2w, row-major */
  float t[3];             /* m_cvMat_t_c2w */
  float origin[3];        /* m_cvMat_origin */
  float fx, fy, cx, cy;
  float min_x, max_x, min_y, max_y;
  float bf;               /* m_flt_baseLineTimesFx */
  float log_scale_factor; /* m_flt_logScaleFactor */
  int32_t n_levels;       /* m_int_scaleLevelsNum */
} YdCamera;

/* what the frustum test reads of one map point: getPosInWorld(), getNormal(), m_flt_maxDistance, m_flt_minDistance */
typedef struct {
  float pos[3];
  float normal[3];
  float max_distance;
  float min_distance;
} YdMapPoint3d;

/* Frame::isInCameraFrustum + MapPoint::predictScaleLevel (src/mapPoint.cpp:265-277) for n map points: out[i] is the
 * tracking state the reference stores in the map point (m_flt_trackProjX/Y/RightX, m_int_trackScaleLevel,
 * m_flt_trackViewCos, m_b_isTrackInView), i.e. exactly the input of ydorb_search_by_projection.  Callers filter bad /
 * already-seen points first (src/tracking.cpp:586-587) or clear `active` afterwards. */
int ydorb_frustum(YdMatcher* m, const YdCamera* cam, const YdMapPoint3d* map_points, int n, float view_cos_limit,
                  YdMapPointProj* out);
int ydorb_frustum_device(YdMatcher* m, const YdCamera* cam, const YdMapPoint3d* d_map_points, int n, float view_cos_limit,
                         YdMapPointProj* d_out);

/* ---- DBoW3 vocabulary transform ------------------------------------------------------------------
 * Frame::computeBoW (src/frame.cpp:265-272) calls DBoW3::Vocabulary::transform(features, BowVector&, FeatureVector&,
 * levelsup = 4) (thirdParty/DBow3/src/Vocabulary.cpp:752-816), which descends the k-ary tree once per descriptor
 * (:836-873: strict `d < best_d`, first child wins ties, node id captured at level L - levelsup). */
typedef struct YdVocabulary YdVocabulary;

/* The tree as DBoW3 holds it in m_nodes (Vocabulary.h: Node {id, weight, children, parent, descriptor, word_id}).
 * Node 0 is the root.  children of node i = children[child_start[i] .. child_start[i+1]) in the order of
 * Node::children (the order the descent visits them); a node without children is a leaf (a word). */
typedef struct {
  int32_t n_nodes;
  int32_t k, L;                /* m_k, m_L */
  int32_t weighting;           /* DBoW3::WeightingType: 0 TF_IDF, 1 TF, 2 IDF, 3 BINARY (BowVector.h:38-44) */
  int32_t scoring;             /* DBoW3::ScoringType: 0 L1_NORM, 1 L2_NORM, 2 CHI_SQUARE, 3 KL, 4 BHATTACHARYYA,
                                  5 DOT_PRODUCT (BowVector.h:47-55) */
  const int32_t* child_start;  /* [n_nodes + 1] */
  const int32_t* children;     /* [n_nodes - 1] */
  const uint8_t* descriptors;  /* [n_nodes][32]; the root's row is never read */
  const double* weights;       /* [n_nodes] Node::weight */
  const int32_t* word_ids;     /* [n_nodes] Node::word_id (meaningful for leaves) */
} YdVocabularyDesc;

/* Result of one transform of n descriptors.  Caller-owned arrays; capacity of the bow_* and fv_node_* arrays is n
 * (fv_node_start: n + 1).  Per-feature arrays may be NULL when not wanted. */
typedef struct {
  int32_t* word_id;      /* [n] WordId of each descriptor (Vocabulary.cpp:870) */
  int32_t* node_id;      /* [n] NodeId at level L - levelsup (0 = root when L - levelsup <= 0, or when the leaf is
                            reached above that level, where the reference leaves it uninitialised) */
  double* weight;        /* [n] Node::weight of the word (:871) */
  int32_t* bow_ids;      /* BowVector (std::map<WordId, WordValue>) in ascending key order */
  double* bow_values;
  int32_t n_bow;
  int32_t* fv_node_ids;  /* FeatureVector in the CSR form of YdFeatureVector: directly usable by ydorb_search_by_bow */
  int32_t* fv_node_start;
  int32_t* fv_items;     /* [n] */
  int32_t n_fv_nodes;
} YdBowResult;

#define YDORB_VOC_MAX_FEATURES 16384 /* descriptors per frame one transform accepts */

int ydorb_vocabulary_create(YdMatcher* m, const YdVocabularyDesc* desc, YdVocabulary** out);
/* Reads the binary stream Vocabulary::toStream writes (Vocabulary.cpp:1180-1256; parsed as Vocabulary::fromStream
 * does, :1335-1409) -- what Vocabulary::load(filename) tries first (:1080-1108) and what the reference's orbvoc.dbow3
 * is.  Both forms: plain, and the default one cut into QuickLZ (level 1) chunks.  The .yml / .txt forms are refused
 * with YDORB_E_INVALID. */
int ydorb_vocabulary_load(YdMatcher* m, const char* path, YdVocabulary** out);
/* Host only (no device needed): the plain form of a binary vocabulary stream given as a file image -- a compressed
 * stream comes back inflated (header flag cleared, chunk framing removed), a plain one unchanged.  *out_len receives
 * the size; with out == NULL nothing is copied (size query), else capacity must hold it. */
int ydorb_vocabulary_stream_inflate(const uint8_t* stream, size_t n, uint8_t* out, size_t capacity, size_t* out_len);
void ydorb_vocabulary_destroy(YdVocabulary* v);
/* n_nodes, k, L, weighting, scoring, number of words (leaves) */
int ydorb_vocabulary_info(const YdVocabulary* v, int32_t* info6);
/* Vocabulary::transform(features, v, fv, levelsup) for one frame; host pointers, n <= YDORB_VOC_MAX_FEATURES. */
int ydorb_vocabulary_transform(YdVocabulary* v, const uint8_t* descriptors, int n, int levels_up, YdBowResult* out);
/* Batched, device-resident: frame b owns rows [b*capacity, b*capacity + d_n[b]) of d_desc (the layout
 * ydorb_extract_batch_device writes) and the same slice of every output array (bow/fv arrays have per-frame stride
 * `capacity`, d_fv_node_start capacity + 1); d_counts[b] = {n_bow, n_fv_nodes}.  Per-feature outputs may be NULL.
 * Runs on the matcher's stream without synchronising. */
int ydorb_vocabulary_transform_batch_device(YdVocabulary* v, const uint8_t* d_desc, const int32_t* d_n, int capacity,
                                            int batch, int levels_up, int32_t* d_word_id, int32_t* d_node_id,
                                            double* d_weight, int32_t* d_bow_ids, double* d_bow_values,
                                            int32_t* d_fv_node_ids, int32_t* d_fv_node_start, int32_t* d_fv_items,
                                            int32_t* d_counts);

/* ---- Input pre-processing around extraction -----------------------------------------------------
 * The steps Tracking::grabImageStereo / grabImageRGBD and the Frame constructors run right before and after
 * extractAndCompute.  They use a YdMatcher handle only as the owner of a device and a stream. */
#define YDORB_RGB2GRAY 0  /* cv::COLOR_RGB2GRAY  (src/tracking.cpp:75, :99)  */
#define YDORB_BGR2GRAY 1  /* cv::COLOR_BGR2GRAY  (:83, :105)                 */
#define YDORB_RGBA2GRAY 2 /* cv::COLOR_RGBA2GRAY (:78, :101)                 */
#define YDORB_BGRA2GRAY 3 /* cv::COLOR_BGRA2GRAY (:86, :107)                 */

/* cv::cvtColor(src, dst, code) on 8-bit images: gray = (R*9798 + G*19235 + B*3735 + 2^14) >> 15 (OpenCV's 8-bit
 * fixed-point path; pinned against cv2 4.13.0).  Host pointers, strides in bytes. */
int ydorb_cvt_gray(YdMatcher* m, const uint8_t* src, int width, int height, int src_stride, int code, uint8_t* dst,
                   int dst_stride);
/* batched, device pointers, asynchronous on the matcher's stream; dst can be handed to ydorb_extract_batch_device */
int ydorb_cvt_gray_device(YdMatcher* m, const uint8_t* d_src, int width, int height, int src_stride, size_t src_frame_stride,
                          int code, uint8_t* d_dst, int dst_stride, size_t dst_frame_stride, int batch);

/* cv::undistortPoints(pts, pts, K, dist, cv::Mat(), P) as Frame::undistortKeyPoints and Frame::computeImageBounds call it
 * (src/frame.cpp:193-213, :224-248): 5 fixed-point iterations in double, result rounded to float.  K, P: row-major
 * 3x3 CV_32F; dist: k1 k2 p1 p2 [k3 [k4 k5 k6 [s1 s2 s3 s4]]] (n_dist in {4, 5, 8, 12}).  xy / xy_out: n (x, y) pairs
 * (may alias).  The caller keeps the reference's `if (dist[0] != 0)` guard.  Host pointers. */
int ydorb_undistort_points(YdMatcher* m, const float* xy, int n, const float* K, const float* dist, int n_dist,
                           const float* P, float* xy_out);
/* Frame::undistortKeyPoints in place on the layout ydorb_extract_batch_device writes (pt.x, pt.y of the first d_n[b]
 * key points of every frame); a no-op when dist[0] == 0, like the reference (:194).  Device pointers, asynchronous. */
int ydorb_undistort_keypoints_device(YdMatcher* m, YdKeyPoint* d_keypoints, const int32_t* d_n, int capacity, int batch,
                                     const float* K, const float* dist, int n_dist, const float* P);

#define YDORB_DEPTH_U16 0 /* CV_16U depth map (TUM RGB-D) */
#define YDORB_DEPTH_F32 1 /* CV_32F depth map */
/* Frame::computeStereoFromRGBD (src/frame.cpp:214-223) fused with the depthImage.convertTo(CV_32F, depth_map_factor) of
 * Tracking::grabImageRGBD (src/tracking.cpp:109-111): instead of converting the whole depth map, only the n pixels
 * under the (already undistorted) key points are read: d = (float)raw * factor (the factor is skipped for a CV_32F map
 * when |factor - 1| <= 1e-5, as at :109); d > 0 -> depth[i] = d, right_x[i] = pt.x - bf / d; else both -1.  The pixel
 * is ((int)pt.y, (int)pt.x) as `at<float>(float, float)` truncates; a key point undistorted to outside the map (where
 * the reference reads out of bounds) gets -1.  Host pointers; depth_stride in bytes. */
int ydorb_stereo_from_rgbd(YdMatcher* m, const YdKeyPoint* keypoints, int n, const void* depth_map, int depth_type,
                           int width, int height, int depth_stride, float depth_map_factor, float bf, float* right_x,
                           float* depth);
int ydorb_stereo_from_rgbd_device(YdMatcher* m, const YdKeyPoint* d_keypoints, const int32_t* d_n, int capacity, int batch,
                                  const void* d_depth_maps, int depth_type, int width, int height, int depth_stride,
                                  size_t depth_frame_stride, float depth_map_factor, float bf, float* d_right_x,
                                  float* d_depth);

#ifdef __cplusplus
}
#endif
#endif /* YDORB_H */
