/* orb_b200.h — C ABI of the B200-native ORB front-end (liborb_b200.so).
 *
 * Drop-in boundary for the hot path of cheukwaylee/ORB_SLAM2_detailed_comments. Every entry point
 * names the reference interface it replaces (paths relative to the reference root):
 *
 *   orb_extractor_create      <- ORBextractor::ORBextractor(int,float,int,int,int)   include/ORBextractor.h:104, src/ORBextractor.cc:776-925
 *   orb_extractor_get_*       <- GetLevels/GetScaleFactor(s)/GetInverseScaleFactors/GetScaleSigmaSquares/
 *                                GetInverseScaleSigmaSquares                          include/ORBextractor.h:134-182
 *   orb_extract_batch         <- ORBextractor::operator()(image, mask, keypoints, descriptors)
 *                                                                                     include/ORBextractor.h:122-126, src/ORBextractor.cc:1961-2084
 *                                (batched: the stereo Frame ctor runs two of them concurrently, src/Frame.cc:158-168)
 *   orb_extractor_pyramid_*   <- public member mvImagePyramid                         include/ORBextractor.h:185 (read by src/Frame.cc:910,1040)
 *   orb_extract_stereo_batch  <- the two operator() calls of the stereo Frame ctor + Frame::ComputeStereoMatches
 *                                                                                     src/Frame.cc:158-168, 880-1176
 *   orb_hamming_knn           <- ORBmatcher::DescriptorDistance in the best/second-best scan
 *                                                                                     include/ORBmatcher.h:62, src/ORBmatcher.cc:2073-2093, 304-342
 *   orb_hamming_knn_noalloc   <- (new, no reference counterpart) orb_hamming_knn without any allocation: fixed device scratch
 *   orb_hamming_knn_sharded   <- (new) the 8-GPU database scan of config C4 behind one call: shard scan + ncclAllGather + merge
 *   orb_hamming_knn_peer      <- (new) the same scan with the exchange INSIDE the kernel: peer stores over NVLink instead of NCCL
 *   orb_hamming_knn_merge     <- (new) top-2 merge of per-shard results, order-equivalent to one sequential scan
 *   orb_descriptor_distance_batch <- DescriptorDistance over N row pairs (e.g. src/Frame.cc:1009, src/MapPoint.cc:484)
 *   orb_search_for_initialization <- ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)   src/ORBmatcher.cc:570-712
 *   orb_search_by_bow         <- ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)                  src/ORBmatcher.cc:248-410
 *   orb_fuse_search           <- the candidate search of ORBmatcher::Fuse (both forms) and SearchBySim3      src/ORBmatcher.cc:1100-1700
 *   orb_search_for_triangulation <- ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) src/ORBmatcher.cc:879-1090
 *   orb_vocabulary_*, orb_bow_transform <- Frame::ComputeBoW = DBoW2 TemplatedVocabulary::transform   src/Frame.cc:750-765,
 *                                Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1136-1195, 1240-1280
 *   orb_undistort_keypoints, orb_image_bounds <- Frame::UndistortKeyPoints / ComputeImageBounds (cv::undistortPoints)
 *                                                                                     src/Frame.cc:770-806, 832-868
 *   orb_distinctive_descriptors <- MapPoint::ComputeDistinctiveDescriptors, batched over map points   src/MapPoint.cc:424-517
 *   orb_search_by_projection  <- the guided matchers' core: Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea
 *                                (src/Frame.cc:432-460, 729-745, 633-717), the candidate scan with sequential claims of
 *                                ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th) (src/ORBmatcher.cc:69-184) and
 *                                SearchByProjection(Frame&, Frame&, th, bMono) (src/ORBmatcher.cc:1702-1866), and
 *                                ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:2026-2068)
 *
 * Plain C: pointers and sizes only, no C++/torch types, no exceptions cross the boundary. All
 * buffers are caller-owned. Functions are thread-safe per handle (one handle = one CUDA stream +
 * private workspaces, like one ORBextractor instance, which is not re-entrant either).
 * There is no CPU fallback: without a CUDA device every compute call returns ORB_E_CUDA.
 */
#ifndef ORB_B200_H_
#define ORB_B200_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORB_B200_ABI_VERSION 1

enum orb_status {
  ORB_OK = 0,
  ORB_E_BADARG = -1,   /* null pointer, non-positive size, unsupported parameter */
  ORB_E_CAPACITY = -2, /* an internal or caller buffer was too small (raw keypoints / kps cap); outputs invalid */
  ORB_E_CUDA = -3,     /* CUDA runtime error or no device; see orb_last_error() */
  ORB_E_SHAPE = -4     /* image too small for the pyramid, or aspect < 0.5 (the reference divides by nIni == 0, src/ORBextractor.cc:1059-1063) */
};

enum orb_mem { ORB_MEM_HOST = 0, ORB_MEM_DEVICE = 1 };

enum orb_flags {
  ORB_FLAG_UPSTREAM_PYRAMID = 1, /* raulmur/ORB_SLAM2's ComputePyramid, i.e. the reference without the line its annotator added
                                    (`mvImagePyramid[level] = temp;`, src/ORBextractor.cc:2166): mvImagePyramid[l] is the w_l x h_l image, each
                                    level is resized from it, keypoints are in its coordinates. Non-default; same kernels with a
                                    stored border of 0 instead of 19 px. orb_extractor_pyramid_dims / _read return w_l x h_l then. */
  ORB_FLAG_LEGACY_BLUR = 2       /* OpenCV <= 3.4.0 Gaussian taps [18,34,49,55,49,34,18] instead of [18,34,48,56,48,34,18] */
};

typedef struct orb_params {
  int32_t nfeatures;   /* ORBextractor.nFeatures  (src/Tracking.cc:160) */
  float scale_factor;  /* ORBextractor.scaleFactor */
  int32_t nlevels;     /* ORBextractor.nLevels    (1..16) */
  int32_t ini_th_fast; /* ORBextractor.iniThFAST */
  int32_t min_th_fast; /* ORBextractor.minThFAST */
  int32_t flags;       /* orb_flags */
} orb_params;

/* Same memory layout as cv::KeyPoint (28 bytes): a std::vector<cv::KeyPoint> can be filled by memcpy. */
typedef struct orb_keypoint {
  float x, y;     /* pt, in level-0 padded coordinates of this fork (src/ORBextractor.cc:2066-2076) */
  float size;     /* (int)(31 * mvScaleFactor[octave]) */
  float angle;    /* IC_Angle, degrees [0,360) */
  float response; /* FAST score */
  int32_t octave;
  int32_t class_id; /* -1 */
} orb_keypoint;

typedef struct orb_knn2 {
  int32_t d1, d2; /* best / second-best Hamming distance; == sentinel where unset */
  int64_t i1, i2; /* global database row of best / second best; -1 where unset */
} orb_knn2;

typedef struct orb_extractor orb_extractor;

const char* orb_last_error(void); /* thread-local message of the last failing call */
int orb_abi_version(void);
int orb_device_count(void);

/* ---- extractor ------------------------------------------------------------------------------- */
int orb_extractor_create(const orb_params* p, int device, orb_extractor** out);
void orb_extractor_destroy(orb_extractor* e);

int orb_extractor_get_levels(const orb_extractor* e);
float orb_extractor_get_scale_factor(const orb_extractor* e);
/* what: 0 mvScaleFactor, 1 mvInvScaleFactor, 2 mvLevelSigma2, 3 mvInvLevelSigma2; out[nlevels] */
int orb_extractor_get_table(const orb_extractor* e, int what, float* out);
int orb_extractor_get_features_per_level(const orb_extractor* e, int32_t* out); /* mnFeaturesPerLevel */
/* Keypoint capacity per image for buffer sizing: nfeatures + 16*nlevels. A level's list ends at max(quota + 3, 4*nIni) at most (a split can
 * overshoot the quota by 3; DistributeOctTree starts from nIni = round(width / height) roots, src/ORBextractor.cc:1059-1063), so 16 per level covers
 * nIni <= 4, i.e. aspect ratios up to 4.5. A `cap` below what a frame produces is reported as ORB_E_CAPACITY (flag 8), never written past. */
int orb_extractor_max_keypoints(const orb_extractor* e);

/* operator() over a batch of equally sized CV_8UC1 images.
 *  images      : `batch` pointers (host array) to image rows; all host (ORB_MEM_HOST) or all device memory
 *  step        : row pitch in bytes (same for all images)
 *  kps, desc   : [batch][cap] orb_keypoint and [batch][cap][32] bytes, in `out_mem` memory; cap >= orb_extractor_max_keypoints
 *  counts      : [batch] keypoints per image (always HOST memory); level_counts [batch][nlevels] optional (HOST)
 * Order of keypoints == the reference's: levels ascending, quadtree list order within a level.
 * Blocks until the results are complete (the reference call is synchronous too). */
int orb_extract_batch(orb_extractor* e, const uint8_t* const* images, int in_mem, int step, int width, int height,
                      int batch, orb_keypoint* kps, uint8_t* desc, int out_mem, int cap, int32_t* counts,
                      int32_t* level_counts);

/* operator() for both images of `pairs` rectified stereo pairs PLUS Frame::ComputeStereoMatches (src/Frame.cc:880-1176) on the
 * device, so neither pyramid has to be copied back for it (the reference reads both extractors' mvImagePyramid, src/Frame.cc:910,
 * 1040). images = 2*pairs pointers, left and right interleaved (L0, R0, L1, R1, ...); kps / desc / counts / level_counts as in
 * orb_extract_batch for the 2*pairs images. bf = mbf (baseline * fx), b = mb (baseline); both extractors of the reference share one
 * parameter set (src/Tracking.cc:171-176), so one handle serves both images.
 *  uright, depth : [pairs][cap] float32 in `out_mem` memory: mvuRight / mvDepth of the LEFT keypoints (-1 = no match)
 *  matched       : [pairs] matches kept after the median filter (HOST, optional)
 * Requires cap <= 65535. Rows outside the level-0 image hold no candidates and an empty match list skips the median filter
 * (the reference is undefined there). */
int orb_extract_stereo_batch(orb_extractor* e, const uint8_t* const* images, int in_mem, int step, int width, int height,
                             int pairs, orb_keypoint* kps, uint8_t* desc, int out_mem, int cap, int32_t* counts,
                             int32_t* level_counts, float bf, float b, float* uright, float* depth, int32_t* matched);

/* Device-resident fast path used by the benchmark: `images` is ONE device buffer [batch][height][step];
 * outputs stay in device memory (kps/desc as above, counts_dev [batch] int32 device). Asynchronous on the
 * extractor's stream unless `sync` != 0. With sync == 0 the capacity flag (`cap` smaller than orb_extractor_max_keypoints(); the internal
 * per-cell / per-level buffers are sized by geometric bounds and cannot overflow) is NOT read by the call: an asynchronous caller polls
 * orb_extractor_check() at a point where it synchronises anyway. */
int orb_extract_batch_device(orb_extractor* e, const uint8_t* images, int step, int width, int height, int batch,
                             orb_keypoint* kps, uint8_t* desc, int cap, int32_t* counts_dev, int sync);

/* Waits for the handle's queued work and returns ORB_E_CAPACITY (message: which limit) if any call since the last check overflowed a
 * capacity, ORB_E_CUDA on a device error, else ORB_OK; clears the flags. The synchronous entry points do this themselves. */
int orb_extractor_check(orb_extractor* e);
/* mvImagePyramid of image `index` of the last internal pass (index -1 = the last image of the last call, which is what the
 * reference's member holds): padded level (cols x rows), copied to `dst` (host, pitch dst_step).
 * blurred != 0 returns the Gaussian-blurred working copy instead (test hook). */
int orb_extractor_pyramid_dims(const orb_extractor* e, int level, int* cols, int* rows);
int orb_extractor_pyramid_read(orb_extractor* e, int index, int level, int blurred, uint8_t* dst, int dst_step);
/* All levels of one resident image in one call: nlevels asynchronous 2-D copies and ONE synchronisation (orb_extractor_pyramid_read costs one
 * synchronisation per level). level_dst[l] / level_steps[l] as dst / dst_step above. */
int orb_extractor_pyramid_read_all(orb_extractor* e, int index, int blurred, uint8_t* const* level_dst, const int32_t* level_steps);
/* Registers host destinations (pinned memory recommended: orb_host_alloc) for the padded levels of the LAST image of every following
 * orb_extract_batch / orb_extract_stereo_batch call with this image size — what the drop-in's public mvImagePyramid has to hold after each
 * operator() (include/ORBextractor.h:185, read by Frame::ComputeStereoMatches, src/Frame.cc:910, 1040). The copy is issued as soon as the
 * pyramid kernels of the call's last pass are done and overlaps FAST / quadtree / descriptors; it is complete when the call returns.
 * Needs the level sizes, i.e. one image of this size must have been processed (else ORB_E_BADARG); a call with another image size drops the
 * registration. level_dst == NULL switches it off. The buffers stay the caller's and must outlive the registration. */
int orb_extractor_set_pyramid_download(orb_extractor* e, uint8_t* const* level_dst, const int32_t* level_steps);
/* Page-locked host memory for staging buffers that live as long as the caller's object (cudaMallocHost / cudaFreeHost without a CUDA
 * dependency in the caller's build). */
int orb_host_alloc(size_t bytes, void** out);
void orb_host_free(void* p);
/* Test hook: vToDistributeKeys of (image index, level) of the last batch: x, y (cell-shifted, border-relative) and response. */
int orb_extractor_raw_keypoints(orb_extractor* e, int index, int level, orb_keypoint* out, int cap);
/* CUDA stream of the handle (cudaStream_t) for callers that enqueue dependent work, and the number of kernels
 * launched by this handle so far. */
void* orb_extractor_stream(orb_extractor* e);
int64_t orb_extractor_launch_count(const orb_extractor* e);
/* Per-stage device timing: CUDA events recorded on the handle's stream around each stage of every internal pass and
 * resolved lazily by stage_times (no synchronisation inside the pass). Off by default. Stages: 0 pyramid, 1 FAST cells, 2 blur, 3 quadtree,
 * 4 orientation+descriptor+output. set_profiling resets the accumulators; stage_times returns the stage count,
 * fills ms[ORB_NUM_STAGES] (sum over passes) and *passes (internal passes = kernel launches per stage). */
#define ORB_NUM_STAGES 5
int orb_extractor_set_profiling(orb_extractor* e, int on);
int orb_extractor_stage_times(orb_extractor* e, double* ms, int64_t* passes);

/* ---- matcher ------------------------------------------------------------------------------- */
/* Threads and streams: the matcher / vocabulary / frame-geometry entry points keep their intermediates in one scratch per device (per
 * vocabulary handle for orb_bow_transform). They may be called from any thread and on any stream: host-side set-up is serialised by a mutex,
 * and every call makes its stream wait for the previous user of the scratch (an event recorded at the end of that call) before its own
 * kernels run, so device-memory calls that return before their kernels have finished cannot overwrite each other's intermediates. The price:
 * calls on one device execute one after another on the GPU, whatever their streams. orb_hamming_knn_noalloc is outside this chain (see there). */
/* Brute-force k=2 over 256-bit descriptors with the reference update rule (strict <: earliest index wins,
 * a duplicate of the minimum becomes `second`). sentinel = initial best/second (256; INT_MAX at src/ORBmatcher.cc:611).
 * query [Q][32], db [D][32] in q_mem / db_mem memory; out [Q] in out_mem memory. index_base is added to i1/i2
 * (database shard offset). stream: cudaStream_t or NULL (default stream); the call synchronises before returning
 * when any buffer is host memory. */
int orb_hamming_knn(const uint8_t* query, int q_mem, int Q, const uint8_t* db, int db_mem, int64_t D,
                    int64_t index_base, int sentinel, orb_knn2* out, int out_mem, void* stream);
/* Allocation-free variant of orb_hamming_knn. It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid allocation). It calls no
 * cudaMalloc / cudaFree / allocating library routine: the per-chunk partial keys live in a file-scope device array of
 * ORB_KNN_NOALLOC_SCRATCH_BYTES (4 MiB) that every GPU loading the library holds until exit. Same results as orb_hamming_knn.
 * Requirements: query [Q][32] and db [D][32] in DEVICE memory, 16-byte aligned; out [Q] in DEVICE memory; the caller's CURRENT device is
 * the one the buffers live on. Limits: Q <= ORB_KNN_NOALLOC_MAX_Q and ceil(D / 2^22) <= SCRATCH_BYTES / (16 * Q) (>= 64, i.e. D up to
 * 2.6e8 rows at Q = 4096); beyond them the call returns ORB_E_CAPACITY (ORB_E_BADARG for misaligned buffers) and launches nothing.
 * Asynchronous on `stream` (cudaStream_t or NULL). The first call on each device resolves the array's device address
 * (cudaGetSymbolAddress): make one warm-up call before capturing a graph. All calls on a device share the array; it keeps the last call's
 * contents and nothing in it has to start at zero (each call writes every word it reads). Calls that OVERLAP on the GPU (different
 * streams / threads) race on it: keep them on one stream. */
#define ORB_KNN_NOALLOC_SCRATCH_BYTES (4 << 20)
#define ORB_KNN_NOALLOC_MAX_Q 4096
int orb_hamming_knn_noalloc(const uint8_t* query, int Q, const uint8_t* db, int64_t D, int64_t index_base, int sentinel,
                            orb_knn2* out, void* stream);
/* Merge `nshards` per-shard results (shard-major [nshards][Q], device memory, shards in ascending index order)
 * into out[Q] (device): equals one sequential scan over the concatenated database. */
int orb_hamming_knn_merge(const orb_knn2* shards, int nshards, int Q, int sentinel, orb_knn2* out, void* stream);
/* Database sharded over the ranks of an NCCL communicator (contiguous shards, ascending rank = ascending rows, queries
 * replicated): scan of this rank's shard with GLOBAL indices (index_base = first row of the shard), ONE ncclAllGather of the Q x 24-byte
 * records, device merge. Every rank's out[Q] equals one sequential scan of src/ORBmatcher.cc:304-335 over the concatenated database (earliest
 * global index wins ties, across shards too). Memory-space flags, alignment staging and synchronisation as orb_hamming_knn (device buffers on
 * the caller's current device: asynchronous on `stream`). Collective: every rank of the communicator must call it, with the same Q. nccl_comm: an ncclComm_t of the NCCL
 * library loaded in this process (libnccl.so.2 is dlopen'ed: an already loaded copy - torch.distributed's included - is the one used), e.g.
 * from orb_nccl_comm_create or from the caller's own NCCL set-up. */
int orb_hamming_knn_sharded(const uint8_t* query, int q_mem, int Q, const uint8_t* db_shard, int db_mem, int64_t D_shard, int64_t index_base,
                            int sentinel, orb_knn2* out, int out_mem, void* nccl_comm, void* stream);
/* Communicator plumbing for hosts without an NCCL set-up of their own: rank 0 makes the 128-byte id (ncclGetUniqueId) and hands it to the
 * other ranks by any means (file, socket, MPI); every rank then creates its communicator on its CURRENT device (ncclCommInitRank). */
int orb_nccl_unique_id(uint8_t* id128);
int orb_nccl_comm_create(const uint8_t* id128, int nranks, int rank, void** comm);
void orb_nccl_comm_destroy(void* comm);
/* The sharded scan with the exchange fused into the reduction kernel (no NCCL, no collective launch): every rank owns a mailbox in device memory
 * that the other ranks of the node open with CUDA IPC; the kernel that reduces a rank's per-chunk keys stores its two best keys per query straight
 * into every peer's mailbox over NVLink, publishes an epoch flag, waits for the peers' flags in its own mailbox and merges. Same results as
 * orb_hamming_knn_sharded (the two lexicographically smallest (distance, global row) over all shards).
 *   orb_knn_peer_create   on the caller's CURRENT device; one process per GPU, 1 <= nranks <= ORB_KNN_PEER_MAX_RANKS, every call of the group has Q <= max_q;
 *   orb_knn_peer_export   64 bytes (cudaIpcMemHandle_t) to hand to the other ranks by any means (all-gather over torch.distributed / MPI / files);
 *   orb_knn_peer_connect  handles = nranks x 64 bytes in rank order (this rank's own entry is ignored); needs peer access between the GPUs (NVLink or PCIe P2P);
 *   orb_hamming_knn_peer  collective: every rank calls it with the same Q, in the same order of calls. Buffers as orb_hamming_knn_sharded.
 *                         A peer that never makes the matching call turns into ORB_E_CAPACITY (flag 32) after a bounded wait, reported by the next
 *                         synchronising call (orb_knn_peer_check, or this call itself when a host buffer makes it synchronous).
 *   orb_knn_peer_check    synchronises `stream` and returns ORB_E_CAPACITY if a wait timed out since the last check. */
#define ORB_KNN_PEER_MAX_RANKS 16
#define ORB_KNN_PEER_HANDLE_BYTES 64
typedef struct orb_knn_peer orb_knn_peer;
int orb_knn_peer_create(int rank, int nranks, int max_q, orb_knn_peer** out);
int orb_knn_peer_export(orb_knn_peer* p, uint8_t* handle64);
int orb_knn_peer_connect(orb_knn_peer* p, const uint8_t* handles);
int orb_hamming_knn_peer(orb_knn_peer* p, const uint8_t* query, int q_mem, int Q, const uint8_t* db_shard, int db_mem, int64_t D_shard,
                         int64_t index_base, int sentinel, orb_knn2* out, int out_mem, void* stream);
int orb_knn_peer_check(orb_knn_peer* p, void* stream);
void orb_knn_peer_destroy(orb_knn_peer* p);
/* DescriptorDistance for n row pairs: out[i] = popcount(a[i] ^ b[i]); all device memory. */
int orb_descriptor_distance_batch(const uint8_t* a, const uint8_t* b, int64_t n, int32_t* out, void* stream);
/* Same with HOST buffers (copies inside the call, default stream, synchronous). */
int orb_descriptor_distance_host(const uint8_t* a, const uint8_t* b, int64_t n, int32_t* out);
int64_t orb_matcher_launch_count(void);

/* ---- guided matching ----------------------------------------------------------------------
 * One projected map point = one query. The host computes the projection (pose, camera model, PredictScale stay in the
 * reference's Tracking / MapPoint code); everything from GetFeaturesInArea on runs on the device, for `frames`
 * independent (query set, target frame) pairs per call. */
enum orb_query_flags {
  ORB_QUERY_SKIP = 1,   /* not searched: !mbTrackInView, isBad(), outlier, invzc < 0, projection outside the image bounds */
  ORB_QUERY_CLAIMS = 2  /* pMP->Observations() > 0: once matched, the keypoint is skipped by all later queries (src/ORBmatcher.cc:127-130, 1798-1801) */
};

typedef struct orb_proj_query {   /* 32 bytes */
  float u, v;           /* x, y of GetFeaturesInArea: mTrackProjX/Y (src/ORBmatcher.cc:101) or the projection (:1755-1756) */
  float radius;         /* r of GetFeaturesInArea = r*mvScaleFactors[level] (:102) / th*mvScaleFactors[octave] (:1769); also the stereo tolerance (:134-141, :1804-1811) */
  float ur;             /* predicted right-image column: mTrackProjXR (:137) / u - mbf*invzc (:1806); read only where uright[i] > 0 */
  float angle;          /* angle of the query's own keypoint (rotation histogram, :1828-1839); unused without check_orientation */
  int16_t min_level;    /* GetFeaturesInArea(minLevel, maxLevel) (src/Frame.cc:673-703): levels are checked iff */
  int16_t max_level;    /*   min_level > 0 || max_level >= 0; max_level < 0 = no upper bound */
  int32_t flags;        /* orb_query_flags */
  int32_t reserved;
} orb_proj_query;

typedef struct orb_search_params {
  float min_x, min_y, max_x, max_y; /* Frame::mnMinX, mnMinY, mnMaxX, mnMaxY (src/Frame.cc:849-868, 186-202): grid of 64 x 48 cells over this box */
  int32_t th_dist;           /* accept iff bestDist <= th_dist: TH_HIGH (:170, :1824), or TH_LOW / ORBdist of the other variants */
  float nn_ratio;            /* > 0: reject when bestLevel == bestLevel2 && bestDist > nn_ratio * bestDist2 (:174); <= 0: best only (:1813-1821) */
  int32_t check_orientation; /* != 0: rotation histogram of HISTO_LENGTH = 30 bins, keep the three maxima (:1828-1863, 2026-2068) */
  int32_t reserved;
} orb_search_params;

/* For every frame f < frames, with N = n_kps[f] target keypoints and M = n_queries[f] queries processed in index order
 * (the reference's loop order; the claims make the result order-dependent):
 *   candidates of a query = GetFeaturesInArea(u, v, radius, min_level, max_level) in the reference's scan order (cell
 *   column ix, cell row iy, insertion order), minus keypoints with claimed[i] != 0 or claimed by an earlier query, minus
 *   keypoints with uright[i] > 0 && |ur - uright[i]| > radius; best / second best by DescriptorDistance with strict <;
 *   accept rule per orb_search_params; on accept assigned[f][best] = query index (later accepts overwrite, as
 *   F.mvpMapPoints[bestIdx] = pMP does) and nmatches[f]++; with check_orientation the accepts outside the three
 *   dominant rotation bins are undone (assigned = -1, nmatches--).
 * kps_un [frames][cap] (pt, octave, angle are read), desc [frames][cap][32], uright [frames][cap] or NULL (monocular: all -1),
 * claimed [frames][cap] bytes or NULL, queries [frames][qcap], qdesc [frames][qcap][32], n_kps / n_queries / nmatches [frames],
 * assigned [frames][cap] (-1 = untouched), query_match [frames][qcap] or NULL (keypoint accepted for the query and still
 * standing after the orientation filter, else -1). `mem` = ORB_MEM_HOST or ORB_MEM_DEVICE for ALL array arguments
 * (prm is always host). n_kps[f] <= cap < 2^18; descriptor arrays 16-byte aligned in device memory. stream as in orb_hamming_knn; host memory makes the call synchronous. */
int orb_search_by_projection(int mem, const orb_keypoint* kps_un, const uint8_t* desc, const float* uright, const uint8_t* claimed,
                             const int32_t* n_kps, int cap, const orb_proj_query* queries, const uint8_t* qdesc,
                             const int32_t* n_queries, int qcap, int frames, const orb_search_params* prm, int32_t* assigned,
                             int32_t* query_match, int32_t* nmatches, void* stream);

/* ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:1100-1280), Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (:1282-1430)
 * and the two projection passes of SearchBySim3 (:1433-1700): the candidate search of every projected map point on its own (these
 * loops have no claims between map points). For query j of frame f: candidates = GetFeaturesInArea(u, v, radius) with
 * min_level <= octave <= max_level (the reference's "kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel", :1188);
 * with inv_level_sigma2 != NULL (host table of 16 floats = pKF->mvInvLevelSigma2, zero-padded) a candidate is dropped when its
 * reprojection error fails the chi-square test of the first Fuse form (:1193-1218: stereo keypoints, uright[i] >= 0, use
 * ex^2 + ey^2 + er^2 against 7.8, monocular ones ex^2 + ey^2 against 5.99; float32 products, double compare); best =
 * smallest DescriptorDistance, first in scan order on ties. best_dist [frames][qcap] (optional) = bestDist (256 = no candidate),
 * best_idx [frames][qcap] = bestIdx if bestDist <= prm->th_dist (TH_LOW for Fuse, TH_HIGH for SearchBySim3) else -1. The map
 * bookkeeping (Replace / AddObservation, the mutual-consistency check of SearchBySim3) stays with the caller. Arrays as for
 * orb_search_by_projection; prm->nn_ratio and check_orientation are ignored. With inv_level_sigma2 the call synchronises the stream
 * before it returns (the table is staged through a per-device buffer). */
int orb_fuse_search(int mem, const orb_keypoint* kps_un, const uint8_t* desc, const float* uright, const int32_t* n_kps, int cap,
                    const orb_proj_query* queries, const uint8_t* qdesc, const int32_t* n_queries, int qcap, int frames,
                    const orb_search_params* prm, const float* inv_level_sigma2, int32_t* best_idx, int32_t* best_dist, void* stream);

/* Register-resident xor+popc+add microbenchmark: returns measured popc.b32 per second on `device` (roofline denominator). */
double orb_measure_popc_peak(int device);

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (src/ORBmatcher.cc:570-712) for `frames`
 * independent frame pairs. Queries = the keypoints of F1 in index order: u, v = vbPrevMatched[i1], radius = windowSize,
 * min_level = max_level = 0 (GetFeaturesInArea(…, level1, level1), :602), angle = mvKeysUn[i1].angle, ORB_QUERY_SKIP for
 * level1 > 0 (:596-597); kps_un / desc = F2. A keypoint of F2 matched at distance d stays available to a later, strictly closer
 * query, which evicts the earlier match (:624, :647-651); accept iff bestDist <= th_dist && bestDist < (float)bestDist2 * nn_ratio
 * with bestDist2 = INT_MAX when there is no second candidate (:611-612, :641-644); rotation histogram over all accepts, removal
 * of the standing matches outside the three maxima (:680-703). matches12 [frames][qcap] = vnMatches12, matches21 [frames][cap]
 * (optional) = inverse map of the standing matches, nmatches [frames]. prm as for orb_search_by_projection (bounds, th_dist = TH_LOW,
 * nn_ratio, check_orientation). cap <= 32768. The caller updates vbPrevMatched from matches12 (:707-709). */
int orb_search_for_initialization(int mem, const orb_keypoint* kps_un, const uint8_t* desc, const int32_t* n_kps, int cap,
                                  const orb_proj_query* queries, const uint8_t* qdesc, const int32_t* n_queries, int qcap, int frames,
                                  const orb_search_params* prm, int32_t* matches12, int32_t* matches21, int32_t* nmatches, void* stream);

/* ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:248-410) for `frames` independent
 * (keyframe, frame) pairs: keyframe features with a valid map point are matched against the frame's features of the same
 * vocabulary node (both FeatureVectors in the CSR layout of orb_bow_transform), best / second best with strict <, accept iff
 * best <= th_dist && (float)best < nn_ratio * (float)second (:339-342), a matched frame feature is skipped by later keyframe
 * features (:300-301), rotation histogram + ComputeThreeMaxima when check_orientation (:349-365, :393-408).
 * kf_desc [frames][qcap][32], kf_kps_un [frames][qcap] (angle read), kf_valid [frames][qcap] (vpMapPointsKF[i] && !isBad()),
 * kf_fv_* / f_fv_* as written by orb_bow_transform with cap = qcap / cap; f_desc [frames][cap][32], f_kps [frames][cap] (mvKeys,
 * angle read). assigned [frames][cap] = keyframe feature index matched to frame feature i (vpMapPointMatches[i] =
 * vpMapPointsKF[assigned]) or -1; nmatches [frames]. All arrays in `mem` memory; cap, qcap < 2^18.
 * The KeyFrame-KeyFrame form SearchByBoW(pKF1, pKF2, vpMatches12) (src/ORBmatcher.cc:722-880) is the same walk with
 * f_blocked [frames][cap] (optional, may be NULL) = "!pMP2 || pMP2->isBad()" (:771-776), th_dist = TH_LOW - 1 ("bestDist1 <
 * TH_LOW", :792) and the per-keyframe-feature result kf_match [frames][qcap] (optional, may be NULL; vpMatches12[idx1] =
 * vpMapPoints2[kf_match], -1 = none). */
int orb_search_by_bow(int mem, const uint8_t* kf_desc, const orb_keypoint* kf_kps_un, const uint8_t* kf_valid, const int32_t* kf_fv_node,
                      const int32_t* kf_fv_start, const int32_t* kf_fv_feature, const int32_t* kf_fv_count, int qcap, const uint8_t* f_desc,
                      const orb_keypoint* f_kps, const uint8_t* f_blocked, const int32_t* f_fv_node, const int32_t* f_fv_start,
                      const int32_t* f_fv_feature, const int32_t* f_fv_count, int cap, int frames, int th_dist, float nn_ratio,
                      int check_orientation, int32_t* assigned, int32_t* kf_match, int32_t* nmatches, void* stream);

/* ---- ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:879-1090) ------------
 * LocalMapping's matcher between two keyframes: features of KF1 without a map point against the features of KF2 in the same
 * vocabulary node that have no map point and are not matched yet; a candidate needs DescriptorDistance <= th_dist (TH_LOW = 50)
 * and <= the running best ("dist > bestDist -> continue": a later candidate of equal distance replaces the earlier one), must lie
 * away from the epipole when both keypoints are monocular (:1005-1013) and satisfy CheckDistEpipolarLine (:206-245); the winner
 * claims its KF2 keypoint (vbMatched2); then the rotation histogram / ComputeThreeMaxima filter. Same FeatureVector CSR layout as
 * orb_search_by_bow, for `pairs` independent keyframe pairs.
 *   kf1_free    [pairs][qcap] u8: 1 = feature takes part (!pKF1->GetMapPoint(idx1), and bStereo1 if bOnlyStereo)
 *   kf2_blocked [pairs][cap]  u8: 1 = candidate excluded (pKF2->GetMapPoint(idx2), or !bStereo2 if bOnlyStereo)
 *   kf?_uright  mvuRight of both keyframes (< 0 = monocular keypoint)
 *   epi         [pairs]: F12 row-major, the epipole (ex, ey) of camera 1 in image 2 (:885-897), pKF2->mvLevelSigma2, mvScaleFactors
 *   matches12   [pairs][qcap]: vMatches12 (index in KF2 or -1); vMatchedPairs = its non-negative entries in index order.
 * Float arithmetic is float32 without contraction; the chi-square compare is done in double as in the reference. */
typedef struct orb_epipolar {
  float f12[9];
  float ex, ey;
  float level_sigma2[16];
  float scale_factor[16];
  float reserved;
} orb_epipolar;
int orb_search_for_triangulation(int mem, const uint8_t* kf1_desc, const orb_keypoint* kf1_kps_un, const uint8_t* kf1_free, const float* kf1_uright,
                                 const int32_t* kf1_fv_node, const int32_t* kf1_fv_start, const int32_t* kf1_fv_feature, const int32_t* kf1_fv_count,
                                 int qcap, const uint8_t* kf2_desc, const orb_keypoint* kf2_kps_un, const uint8_t* kf2_blocked, const float* kf2_uright,
                                 const int32_t* kf2_fv_node, const int32_t* kf2_fv_start, const int32_t* kf2_fv_feature, const int32_t* kf2_fv_count,
                                 int cap, int pairs, const orb_epipolar* epi, int th_dist, int check_orientation, int32_t* matches12,
                                 int32_t* nmatches, void* stream);

/* ---- Frame::ComputeBoW (src/Frame.cc:750-765) --------------------------------------------------------
 * = DBoW2 TemplatedVocabulary<FORB::TDescriptor, FORB>::transform(features, BowVector&, FeatureVector&, levelsup)
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1136-1195, 1240-1280) for the ORB vocabulary's configuration
 * (TF_IDF weighting, L1_NORM scoring: ORBvoc.txt header "10 6 0 0").
 * orb_vocabulary_create takes the node table exactly as loadFromTextFile reads it (TemplatedVocabulary.h:1330-1420):
 * node 0 is the root; for node i in 1..n_nodes-1 (file line order) parent[i], is_leaf[i], desc[i][32], weight[i]
 * (entry 0 of every array is ignored). Children keep file order, word ids are assigned to leaves in file order. */
typedef struct orb_vocabulary orb_vocabulary;
int orb_vocabulary_create(int k, int L, int n_nodes, const int32_t* parent, const uint8_t* is_leaf, const uint8_t* desc,
                          const double* weight, int device, orb_vocabulary** out);
void orb_vocabulary_destroy(orb_vocabulary* v);
int orb_vocabulary_words(const orb_vocabulary* v);   /* number of leaves = vocabulary size */
/* transform() for `frames` descriptor sets desc [frames][cap][32], n [frames] rows used. Per feature: word [frames][cap] (word id) and
 * node [frames][cap] (ancestor at level L - levelsup; 0 = root when levelsup >= L), also for stopped words (weight 0).
 * BowVector per frame: bow_count[f] entries (bow_word ascending, bow_value L1-normalised doubles) in [frames][cap] arrays.
 * FeatureVector per frame as CSR: fv_count[f] nodes (fv_node ascending), features of node j =
 * fv_feature[f][fv_start[f][j] .. fv_start[f][j+1]) ascending; fv_start is [frames][cap + 1]. All arrays in `mem` memory;
 * cap <= 16384. Results equal the reference's std::map contents bit for bit (doubles included). */
int orb_bow_transform(orb_vocabulary* v, int mem, const uint8_t* desc, const int32_t* n, int cap, int frames, int levelsup, int32_t* word,
                      int32_t* node, int32_t* bow_word, double* bow_value, int32_t* bow_count, int32_t* fv_node, int32_t* fv_start,
                      int32_t* fv_feature, int32_t* fv_count, void* stream);

/* ---- Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cc:770-806, 832-868) ---------------------
 * cv::undistortPoints(pts, pts, mK, mDistCoef, cv::Mat(), mK): 5 fixed iterations in double precision (OpenCV >= 4 default
 * criteria), bit-exact against cv2 4.13. The camera holds mK and mDistCoef as the reference does (CV_32F). */
typedef struct orb_camera {
  float fx, fy, cx, cy;  /* mK */
  float dist[8];         /* mDistCoef: k1 k2 p1 p2 [k3 [k4 k5 k6]] */
  int32_t n_dist;        /* 4, 5 or 8 */
} orb_camera;
/* kps_un[f][i] = kps[f][i] with pt replaced by the undistorted point, i < n[f] (kps_un may alias kps). dist[0] == 0 copies the
 * keypoints unchanged (src/Frame.cc:772-776). Arrays in `mem` memory. */
int orb_undistort_keypoints(int mem, const orb_camera* cam, const orb_keypoint* kps, const int32_t* n, int cap, int frames,
                            orb_keypoint* kps_un, void* stream);
/* bounds[4] = mnMinX, mnMinY, mnMaxX, mnMaxY of a cols x rows image (the four undistorted corners, src/Frame.cc:836-858;
 * 0, 0, cols, rows without distortion). Host output; runs the same device code on the four corners. */
int orb_image_bounds(const orb_camera* cam, int cols, int rows, float* bounds);

/* ---- MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:424-517) ----------------------------
 * For every map point p < points with the descriptors desc[offsets[p] .. offsets[p+1]) of its observations (N_p rows, in
 * the order the reference's std::map<KeyFrame*, size_t> walk pushes them): all-pairs DescriptorDistance, per row the
 * median = element (int)(0.5 * (N_p - 1)) of the sorted row (the zero self-distance included), best = the first row with
 * the smallest median (strict <). best_idx[p] = that row (relative to offsets[p]; -1 for N_p == 0), best_median[p]
 * (optional) its median. All arrays in `mem` memory; offsets has points + 1 entries. */
int orb_distinctive_descriptors(int mem, const uint8_t* desc, const int64_t* offsets, int points, int32_t* best_idx,
                                int32_t* best_median, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ORB_B200_H_ */
