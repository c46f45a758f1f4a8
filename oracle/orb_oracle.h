/* TEST INFRASTRUCTURE ONLY — CPU oracle for the ORB front-end hot path.
 *
 * A CPU restatement of /root/reference src/ORBextractor.cc (ORBextractor ctor :776-925,
 * operator() :1961-2084, ComputePyramid :2093-2168, ComputeKeyPointsOctTree :1424-1601,
 * DistributeOctTree :1050-1417, DivideNode :963-1035, IC_Angle :104-161,
 * computeOrbDescriptor :177-254) and src/ORBmatcher.cc (DescriptorDistance :2073-2093 and the
 * best/second-best idiom :304-342), on top of the OpenCV primitives restated in cv_primitives.*.
 *
 * The reference ships no tests or golden vectors and cannot be built against OpenCV in the build
 * container (no OpenCV C++ headers; include/ORBextractor.h:26), so parity is pinned by
 * (0) the reference's own src/ORBextractor.cc and Thirdparty/DBoW2 compiled unmodified against
 * oracle/cvshim (oracle/_ref, tests/test_reference_binary.py: this oracle == that binary bit for bit),
 * (1) the primitives checked bit-exact against the real OpenCV (cv2 4.13) and (2) a whole-pipeline
 * composition of real cv2 calls (tests/golden/make_golden.py) whose outputs are committed under
 * tests/golden/ and compared with this oracle.
 *
 * Two documented canonicalisations (DESIGN.md "Determinism contract"):
 *  - quadtree "biggest first" tie-break: (size, creation sequence) instead of (size, heap pointer),
 *    src/ORBextractor.cc:1295-1308;
 *  - cos/sin of the keypoint angle: a fixed double-precision polynomial rounded once to float
 *    instead of the platform libm cosf/sinf (src/ORBextractor.cc:184-186); flag
 *    ORB_ORACLE_LIBM_TRIG selects libm for a bit-agreement diagnostic.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use
 * this library. The product (orb_slam2_detailed_comments_b200/) never links or loads it.
 */
#ifndef ORB_ORACLE_H_
#define ORB_ORACLE_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  ORB_ORACLE_UPSTREAM_PYRAMID = 1, /* raulmur/ORB_SLAM2 behaviour: mvImagePyramid = unpadded ROI (not this fork) */
  ORB_ORACLE_LEGACY_BLUR = 2,      /* OpenCV <= 3.4.0 taps [18,34,49,55,49,34,18] */
  ORB_ORACLE_LIBM_TRIG = 4         /* glibc cosf/sinf instead of the canonical polynomial */
};

typedef struct orb_oracle_kp { /* == cv::KeyPoint, 28 bytes */
  float x, y, size, angle, response;
  int32_t octave, class_id;
} orb_oracle_kp;

typedef struct orb_oracle_extractor orb_oracle_extractor;

orb_oracle_extractor* orb_oracle_create(int nfeatures, float scale_factor, int nlevels, int ini_th_fast,
                                        int min_th_fast, int flags);
void orb_oracle_destroy(orb_oracle_extractor* e);

/* Table accessors (what = 0 mvScaleFactor, 1 mvInvScaleFactor, 2 mvLevelSigma2, 3 mvInvLevelSigma2). */
void orb_oracle_get_table(const orb_oracle_extractor* e, int what, float* out);
void orb_oracle_get_quota(const orb_oracle_extractor* e, int32_t* out); /* mnFeaturesPerLevel */
void orb_oracle_get_umax(const orb_oracle_extractor* e, int32_t* out);  /* 16 ints */

/* operator(): returns N (>=0) or -1 if cap is too small / bad args. level_counts may be NULL. */
int orb_oracle_extract(orb_oracle_extractor* e, const uint8_t* img, int step, int width, int height,
                       orb_oracle_kp* kps, uint8_t* desc, int cap, int32_t* level_counts);

/* Stage dumps of the last orb_oracle_extract call. */
int orb_oracle_level_dims(const orb_oracle_extractor* e, int level, int* cols, int* rows);
const uint8_t* orb_oracle_level_ptr(const orb_oracle_extractor* e, int level);   /* mvImagePyramid[level], step = cols */
const uint8_t* orb_oracle_blurred_ptr(const orb_oracle_extractor* e, int level); /* NULL if the level had no keypoints */
int orb_oracle_raw_count(const orb_oracle_extractor* e, int level);
int orb_oracle_raw_keypoints(const orb_oracle_extractor* e, int level, orb_oracle_kp* out, int cap); /* vToDistributeKeys */

/* Primitives (for tests vs cv2). */
void orb_oracle_resize(const uint8_t* src, int sstep, int sw, int sh, uint8_t* dst, int dstep, int dw, int dh);
void orb_oracle_border101(const uint8_t* src, int sstep, int sw, int sh, uint8_t* dst, int dstep, int b);
void orb_oracle_blur(const uint8_t* src, int sstep, int w, int h, uint8_t* dst, int dstep, int legacy);
int orb_oracle_fast(const uint8_t* img, int step, int w, int h, int threshold, orb_oracle_kp* out, int cap);
void orb_oracle_fast_score_map(const uint8_t* img, int step, int w, int h, uint8_t* out /* w*h */);
float orb_oracle_fast_atan2(float y, float x);
void orb_oracle_sincos(float angle_rad, int libm, float* s, float* c);
int orb_oracle_distribute(const orb_oracle_kp* in, int n, int minx, int maxx, int miny, int maxy, int nfeat,
                          orb_oracle_kp* out, int cap);

/* ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:2026-2068) on L bin sizes -> out3 = {ind1, ind2, ind3}. */
void orb_oracle_three_maxima(const int32_t* sizes, int L, int32_t* out3);
/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2073-2093). */
int orb_oracle_descriptor_distance(const uint8_t* a, const uint8_t* b);
/* Brute-force k=2 with the reference's strict-< update rule (src/ORBmatcher.cc:304-335); sentinel
 * is the initial best/second (256 normally, INT_MAX at :611-612). Index base added to i1/i2.
 * out: per query {int32 d1, int32 d2, int64 i1, int64 i2}; i = -1 where unset. nthreads over queries. */
typedef struct orb_oracle_knn2 {
  int32_t d1, d2;
  int64_t i1, i2;
} orb_oracle_knn2;
void orb_oracle_hamming_knn(const uint8_t* q, int Q, const uint8_t* db, int64_t D, int64_t index_base,
                            int sentinel, orb_oracle_knn2* out, int nthreads);

/* Frame::ComputeStereoMatches (src/Frame.cc:880-1176) on the outputs and pyramids of the last orb_oracle_extract call of `left` and
 * `right`. bf = mbf, b = mb. uright / depth [nL], -1 = no match. Returns the number of matches kept. */
int orb_oracle_stereo_match(const orb_oracle_extractor* left, const orb_oracle_extractor* right, const orb_oracle_kp* kl,
                            const uint8_t* dl, int nL, const orb_oracle_kp* kr, const uint8_t* dr, int nR, float bf, float b,
                            float* uright, float* depth);

/* Guided matching core, one (query set, target frame) pair: Frame::AssignFeaturesToGrid / PosInGrid /
 * GetFeaturesInArea (src/Frame.cc:432-460, 729-745, 633-717), the candidate loops of ORBmatcher::SearchByProjection
 * (src/ORBmatcher.cc:69-184 and 1702-1866) and ComputeThreeMaxima (:2026-2068). Layouts as include/orb_b200.h
 * (orb_proj_query = 32 bytes, orb_search_params = 32 bytes). Returns nmatches. */
typedef struct orb_oracle_query {
  float u, v, radius, ur, angle;
  int16_t min_level, max_level;
  int32_t flags, reserved;
} orb_oracle_query;
typedef struct orb_oracle_search_params {
  float min_x, min_y, max_x, max_y;
  int32_t th_dist;
  float nn_ratio;
  int32_t check_orientation, reserved;
} orb_oracle_search_params;
int orb_oracle_search_by_projection(const orb_oracle_kp* kps_un, const uint8_t* desc, const float* uright, const uint8_t* claimed,
                                    int n_kps, const orb_oracle_query* queries, const uint8_t* qdesc, int n_queries,
                                    const orb_oracle_search_params* prm, int32_t* assigned, int32_t* query_match);
/* Candidate search of ORBmatcher::Fuse (src/ORBmatcher.cc:1100-1280; inv_level_sigma2 == NULL: the Scw form and SearchBySim3's passes),
 * one keyframe: per query best_dist (256 = none) and best_idx (-1 unless best_dist <= prm->th_dist). Level window = [min_level, max_level]. */
void orb_oracle_fuse_search(const orb_oracle_kp* kps_un, const uint8_t* desc, const float* uright, int n_kps, const orb_oracle_query* queries,
                            const uint8_t* qdesc, int n_queries, const orb_oracle_search_params* prm, const float* inv_level_sigma2,
                            int32_t* best_idx, int32_t* best_dist);
/* ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:570-712), one frame pair; queries as in include/orb_b200.h. matches12 [n_queries],
 * matches21 [n_kps] (inverse of the standing matches). Returns nmatches. */
int orb_oracle_search_for_initialization(const orb_oracle_kp* kps_un, const uint8_t* desc, int n_kps, const orb_oracle_query* queries,
                                         const uint8_t* qdesc, int n_queries, const orb_oracle_search_params* prm, int32_t* matches12,
                                         int32_t* matches21);
/* Frame::GetFeaturesInArea alone (scan order included): returns the number of indices written to out (cap >= n_kps). */
int orb_oracle_features_in_area(const orb_oracle_kp* kps_un, int n_kps, const orb_oracle_search_params* prm, float x, float y, float r,
                                int min_level, int max_level, int32_t* out);

/* ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:248-410), one pair. FeatureVectors as CSR
 * (node ascending, features of node j = feature[start[j] .. start[j+1])). assigned [n_f] = keyframe feature index or -1. Returns nmatches. */
int orb_oracle_search_by_bow(const uint8_t* kf_desc, const orb_oracle_kp* kf_kps_un, const uint8_t* kf_valid, const int32_t* kf_node,
                             const int32_t* kf_start, const int32_t* kf_feature, int kf_count, const uint8_t* f_desc,
                             const orb_oracle_kp* f_kps, int n_f, const int32_t* f_node, const int32_t* f_start, const int32_t* f_feature,
                             int f_count, int th_dist, float nn_ratio, int check_orientation, int32_t* assigned,
                             const uint8_t* f_blocked /* optional: KeyFrame-KeyFrame form, :771-776 */, int32_t* kf_match /* optional [n_kf] */,
                             int n_kf);

/* ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:879-1090) + CheckDistEpipolarLine (:206-245), one keyframe pair.
 * kf1_has_mp / kf2_has_mp: pKF->GetMapPoint(idx) != NULL; uright: mvuRight (< 0 = monocular); f12: F12 row-major; (ex, ey): the
 * epipole computed by the caller as in :885-897; level_sigma2 / scale_factor: pKF2's tables. matches12 [n1] = vMatches12. Returns nmatches. */
int orb_oracle_search_for_triangulation(const uint8_t* desc1, const orb_oracle_kp* kps1_un, const uint8_t* kf1_has_mp, const float* uright1, int n1,
                                        const int32_t* node1, const int32_t* start1, const int32_t* feature1, int count1,
                                        const uint8_t* desc2, const orb_oracle_kp* kps2_un, const uint8_t* kf2_has_mp, const float* uright2, int n2,
                                        const int32_t* node2, const int32_t* start2, const int32_t* feature2, int count2,
                                        const float* f12, float ex, float ey, const float* level_sigma2, const float* scale_factor,
                                        int only_stereo, int th_low, int check_orientation, int32_t* matches12);

/* Frame::UndistortKeyPoints (src/Frame.cc:770-806): cv::undistortPoints(pts, pts, K, dist, Mat(), K), OpenCV >= 4 default criteria
 * (5 iterations), on n points xy [n][2] (float, in place). k: n_dist (4, 5 or 8) float coefficients. dist[0] == 0 leaves xy unchanged. */
void orb_oracle_undistort_points(float fx, float fy, float cx, float cy, const float* dist, int n_dist, float* xy, int n);

/* Frame::ComputeBoW (src/Frame.cc:750-765) = DBoW2 TemplatedVocabulary::transform(features, BowVector, FeatureVector, levelsup)
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1136-1195, 1240-1280; BowVector.cpp; FeatureVector.cpp) with TF_IDF + L1_NORM.
 * Vocabulary node table as loadFromTextFile reads it (node 0 = root, entries 0 ignored). One descriptor set per call. Outputs as
 * include/orb_b200.h: word/node [n]; bow_word/bow_value [<= n] with *bow_count; fv_node [<= n], fv_start [<= n + 1], fv_feature [<= n]
 * with *fv_count. */
typedef struct orb_oracle_vocabulary orb_oracle_vocabulary;
orb_oracle_vocabulary* orb_oracle_vocabulary_create(int k, int L, int n_nodes, const int32_t* parent, const uint8_t* is_leaf,
                                                    const uint8_t* desc, const double* weight);
void orb_oracle_vocabulary_destroy(orb_oracle_vocabulary* v);
void orb_oracle_bow_transform(const orb_oracle_vocabulary* v, const uint8_t* desc, int n, int levelsup, int32_t* word, int32_t* node,
                              int32_t* bow_word, double* bow_value, int32_t* bow_count, int32_t* fv_node, int32_t* fv_start,
                              int32_t* fv_feature, int32_t* fv_count);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:424-517) for `points` map points; desc rows
 * [offsets[p], offsets[p+1]) are the observations of point p. best_idx -1 for a point without observations. */
void orb_oracle_distinctive_descriptors(const uint8_t* desc, const int64_t* offsets, int points, int32_t* best_idx,
                                        int32_t* best_median, int nthreads);

/* Multi-threaded extraction over frames, one extractor per thread (bench cpu baseline).
 * Returns total keypoints. frames: n contiguous w*h images. */
int64_t orb_oracle_extract_many(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th, int flags,
                                const uint8_t* frames, int width, int height, int n, int nthreads);

#ifdef __cplusplus
}
#endif
#endif
