/*
 * lgbcu.h -- thin extern "C" CUDA layer of the B200-native histogram tree learner.
 *
 * This is the ONLY surface the host C++ (lightgbmv1_b200/csrc/host) calls; host code never includes a CUDA
 * header.  One function per hot-path step so that each step can be parity-tested alone against the
 * reference's CPU implementation (file:line citations are relative to the reference tree,
 * dreaming-panda/LightGBMv1).
 *
 * Conventions (same as the reference's C API, include/LightGBM/c_api.h:1300-1312, src/c_api.cpp:35-49):
 *   - every function returns 0 on success, -1 on failure; lgbcu_last_error() returns a thread-local message;
 *   - plain pointers and sizes only; all pointers are HOST pointers unless the name ends in _dev;
 *   - there is no CPU fallback: if no CUDA device is usable every call fails with a message.
 *
 * Data model held in HBM by one LgbCuContext (= one TreeLearner + ScoreUpdater + Objective of the reference):
 *   bins      u8  [num_data][row_stride]  row-major; element = the reference's single-feature-group encoding
 *                                         (include/LightGBM/feature_group.h:32-50,153-165): 0 = most-frequent
 *                                         bin, else bin + 1 - [most_freq_bin == 0]
 *   grad,hess f32 [num_data]              (score_t, include/LightGBM/meta.h)
 *   score     f64 [num_data]              (src/boosting/score_updater.hpp:21-125)
 *   indices   i32 [2][num_data]           leaf-ordered row indices, ping-pong (src/treelearner/data_partition.hpp)
 *   pool      i64 [num_leaves][total_bin][2]  per-leaf histograms, 64-bit fixed point, layout
 *                                         identical to the reference's hist_t buffer (include/LightGBM/bin.h:32-42,
 *                                         src/treelearner/feature_histogram.hpp:1150-1176)
 */
#ifndef LGBCU_H_
#define LGBCU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define LGBCU_API __declspec(dllexport)
#else
#define LGBCU_API __attribute__((visibility("default")))
#endif

typedef struct LgbCuContext LgbCuContext;

/* missing_type values: include/LightGBM/bin.h (enum MissingType { None, Zero, NaN }) */
enum { LGBCU_MISSING_NONE = 0, LGBCU_MISSING_ZERO = 1, LGBCU_MISSING_NAN = 2 };
/* objective kinds implemented on device (src/objective/) */
enum { LGBCU_OBJ_NONE = 0, LGBCU_OBJ_BINARY = 1, LGBCU_OBJ_L2 = 2, LGBCU_OBJ_LAMBDARANK = 3, LGBCU_OBJ_MULTICLASS = 4,
       LGBCU_OBJ_L1 = 5, LGBCU_OBJ_QUANTILE = 6, LGBCU_OBJ_MULTICLASSOVA = 7,
       LGBCU_OBJ_HUBER = 8 /* p0 = alpha */, LGBCU_OBJ_FAIR = 9 /* p0 = fair_c */, LGBCU_OBJ_POISSON = 10 /* p0 = poisson_max_delta_step */,
       LGBCU_OBJ_GAMMA = 11, LGBCU_OBJ_TWEEDIE = 12 /* p0 = tweedie_variance_power */,
       LGBCU_OBJ_XENTROPY = 13 /* cross_entropy: labels in [0, 1] */,
       LGBCU_OBJ_MAPE = 14 /* L1 with the per-row weight 1 / max(1, |label|); no user weights */,
       LGBCU_OBJ_XENTLAMBDA = 15 /* cross_entropy_lambda */,
       LGBCU_OBJ_XENDCG = 16 /* rank_xendcg: p0 = objective_seed; needs lgbcu_set_query */ };
/* histogram kernel selection (LGBCU_HIST_AUTO picks by leaf size) */
enum { LGBCU_HIST_AUTO = 0, LGBCU_HIST_GLOBAL_ATOMIC = 1, LGBCU_HIST_WARP_PRIVATE = 2 };

#define LGBCU_BIN_NUMERICAL 0
#define LGBCU_BIN_CATEGORICAL 1

/* Per inner feature meta: what FeatureMetainfo + BinMapper give the CPU path
 * (src/treelearner/feature_histogram.hpp:25-38,1110-1148; include/LightGBM/bin.h:85-215). */
typedef struct LgbCuFeatureMeta {
  int32_t num_bin;        /* BinMapper::num_bin() */
  int32_t most_freq_bin;  /* BinMapper::GetMostFreqBin() */
  int32_t default_bin;    /* BinMapper::GetDefaultBin() */
  int32_t missing_type;   /* LGBCU_MISSING_* */
  int32_t real_index;     /* Dataset::RealFeatureIndex(inner) -- used for the tie-break of SplitInfo::operator> */
  int32_t bin_type;       /* LGBCU_BIN_NUMERICAL (0) or LGBCU_BIN_CATEGORICAL (BinType, include/LightGBM/bin.h:24-27) */
  double  penalty;        /* feature_contri (FeatureMetainfo::penalty) */
} LgbCuFeatureMeta;

/* Tree-growing parameters read by the path (include/LightGBM/config.h; used in
 * src/treelearner/serial_tree_learner.cpp:274-297 and feature_histogram.hpp:853-1049). */
typedef struct LgbCuTreeParams {
  int32_t num_leaves;
  int32_t max_depth;              /* <=0: no limit */
  int32_t min_data_in_leaf;
  int32_t reserved;
  double  min_sum_hessian_in_leaf;
  double  lambda_l1;
  double  lambda_l2;
  double  min_gain_to_split;
  double  max_delta_step;         /* <=0: off */
  double  path_smooth;            /* <=kEpsilon: off */
} LgbCuTreeParams;

/* POD twin of SplitInfo (src/treelearner/split_info.hpp:22-53), numerical splits only. */
typedef struct LgbCuSplitInfo {
  int32_t feature;        /* inner feature index, -1 if none */
  int32_t real_feature;   /* real feature index, -1 if none */
  int32_t threshold;      /* threshold in bin */
  int32_t default_left;
  int32_t left_count;
  int32_t right_count;
  double  gain;           /* kMinScore (-inf) if no split */
  double  left_output, right_output;
  double  left_sum_gradient, left_sum_hessian;
  double  right_sum_gradient, right_sum_hessian;
} LgbCuSplitInfo;

/* One applied split, in application order; the host replays them through Tree::Split
 * (src/io/tree.cpp:54-68, include/LightGBM/tree.h:459-501). */
typedef struct LgbCuSplitRecord {
  int32_t leaf;           /* leaf that was split (left child keeps this id) */
  int32_t right_leaf;     /* id of the new right child */
  int32_t feature;        /* inner feature */
  int32_t real_feature;
  int32_t threshold;      /* bin */
  int32_t default_left;
  int32_t left_count, right_count;      /* actual partition sizes (SplitInner update_cnt, serial_tree_learner.cpp:567-571) */
  float   gain;                          /* (float)(gain + min_gain_to_split), serial_tree_learner.cpp:583 */
  int32_t reserved;
  double  left_output, right_output;
  double  left_sum_hessian, right_sum_hessian;
  double  left_sum_gradient, right_sum_gradient;
} LgbCuSplitRecord;

LGBCU_API const char* lgbcu_last_error(void);
/* number of CUDA devices visible (0 if none / no driver) */
LGBCU_API int lgbcu_device_count(int* out);
/* Page-lock (cudaHostRegister) / release a host buffer that will be uploaded with lgbcu_set_dataset, so that the
 * copy runs at full PCIe rate instead of through the driver's pageable staging.  Returns -1 (buffer stays pageable,
 * everything still works) when there is no device or the registration fails.  No reference counterpart: the
 * reference's CPU learner never copies its bin matrix anywhere. */
LGBCU_API int lgbcu_pin_host_buffer(void* ptr, size_t bytes);
LGBCU_API int lgbcu_unpin_host_buffer(void* ptr);
/* number of kernels launched by this library in this process (bench.py's gpu_launches) */
LGBCU_API int64_t lgbcu_launch_count(void);

LGBCU_API int lgbcu_create(int device_id, LgbCuContext** out);
LGBCU_API int lgbcu_destroy(LgbCuContext* ctx);
LGBCU_API int lgbcu_synchronize(LgbCuContext* ctx);

/* A2: make the binned matrix resident.  `bins` is a row-major byte matrix [num_data][FS] (host), FS = one byte column
 * per feature, and two -- low byte, then high byte -- for a feature with more than 256 stored values
 * (1 + num_bin - [most_freq_bin == 0] > 256, i.e. max_bin > 255: the reference's 16-bit DenseBin, src/io/bin.cpp:
 * 635-645); FS = num_features when every feature fits a byte.  It is copied into a row-major device matrix whose row
 * stride is FS rounded up to a multiple of 16 bytes.
 * Replaces FeatureGroup/DenseBin storage (src/io/dense_bin.hpp:55-81) + TreeLearner::Init's view of it
 * (src/treelearner/serial_tree_learner.cpp:28-69). */
LGBCU_API int lgbcu_set_dataset(LgbCuContext* ctx, int32_t num_data, int32_t num_features,
                                const uint8_t* bins, const LgbCuFeatureMeta* meta);
/* Same, but the caller hands over column vectors (one u8[num_data] per inner feature), the layout the
 * reference's DenseBin uses; the transpose to row-major happens on device. */
LGBCU_API int lgbcu_set_dataset_columns(LgbCuContext* ctx, int32_t num_data, int32_t num_features,
                                        const uint8_t* const* columns, const LgbCuFeatureMeta* meta);
/* F1: value -> bin on the device (the row push loop of LGBM_DatasetCreateFromMat, src/c_api.cpp:1082-1097 ->
 * BinMapper::ValueToBin, include/LightGBM/bin.h:457-495).  The host finds the bin mappers on its sample
 * (BinMapper::FindBin); this call uploads the raw matrix once (float32 data_type 0 / float64 data_type 1, row- or
 * column-major) and bins every (row, inner feature) with the reference's binary search over the feature's upper
 * bounds (upper_bounds[bound_offset .. bound_offset + num_bin)) -- bit-identical to the host map.  The result, a
 * row-major [nrow][FS] byte matrix of stored values (FS as for lgbcu_set_dataset), stays on the device: lgbcu_set_dataset_binned makes it a
 * learner's resident matrix without a host round trip, lgbcu_binned_copy_to_host reads it back. */
typedef struct LgbCuBinnedMatrix LgbCuBinnedMatrix;
typedef struct LgbCuBinFeature {
  int32_t column;         /* column of the raw matrix */
  int32_t num_bin;        /* BinMapper::num_bin() */
  int32_t most_freq_bin;
  int32_t missing_type;   /* LGBCU_MISSING_* */
  int32_t bound_offset;   /* first upper bound of this feature in upper_bounds */
  int32_t num_categories; /* -1: numerical.  >= 0: categorical, and upper_bounds[bound_offset ..) holds the feature's
                           * num_categories category values in ascending order followed by their bins (as doubles);
                           * NaN, negative and unlisted values map to bin 0 (BinMapper::ValueToBin, bin.h:457-494) */
  int32_t pad;
  int32_t bundle;         /* 0: a feature of its own.  b > 0: member of exclusive-feature bundle b (the reference's
                           * multi-feature FeatureGroup over ONE DenseBin, include/LightGBM/feature_group.h:32-50): in a
                           * row where several members are off their most-frequent bin, only the member with the
                           * highest `column` keeps its value (the reference's pushes overwrite one another in column
                           * order, dataset.h:382-392 -> dense_bin.hpp:68-80); the others read as most-frequent */
} LgbCuBinFeature;
LGBCU_API int lgbcu_bin_matrix(int device, const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                               int32_t num_features, const LgbCuBinFeature* feats, const double* upper_bounds,
                               int32_t num_bounds, LgbCuBinnedMatrix** out);
LGBCU_API int lgbcu_binned_copy_to_host(LgbCuBinnedMatrix* m, uint8_t* out);
LGBCU_API int lgbcu_binned_device(const LgbCuBinnedMatrix* m, int* device);
LGBCU_API int lgbcu_binned_free(LgbCuBinnedMatrix* m);
LGBCU_API int lgbcu_set_dataset_binned(LgbCuContext* ctx, const LgbCuBinnedMatrix* m, const LgbCuFeatureMeta* meta);

LGBCU_API int lgbcu_set_tree_params(LgbCuContext* ctx, const LgbCuTreeParams* params);
LGBCU_API int lgbcu_set_hist_mode(LgbCuContext* ctx, int mode);

/* Metadata (include/LightGBM/dataset.h Metadata: label_t = float labels/weights, int32 query boundaries) */
LGBCU_API int lgbcu_set_label(LgbCuContext* ctx, const float* label, int32_t n);
LGBCU_API int lgbcu_set_weight(LgbCuContext* ctx, const float* weight, int32_t n); /* weight==NULL clears */
LGBCU_API int lgbcu_set_query(LgbCuContext* ctx, const int32_t* boundaries, int32_t num_queries);

/* A9: objective on device.
 * binary (src/objective/binary_objective.hpp:102-133): p0 = sigmoid, p1 = label_weight[neg], p2 = label_weight[pos]
 * l2     (src/objective/regression_objective.hpp:125-140): no params
 * lambdarank (src/objective/rank_objective.hpp:48-65,139-229): p0 = sigmoid, p1 = truncation_level, p2 = norm(0/1);
 *        label_gain via lgbcu_set_label_gain
 * multiclass (src/objective/multiclass_objective.hpp:86-130, softmax): p0 = num_class, p1 = factor = K / (K - 1);
 *        fills all K gradient / hessian planes from all K score planes in one pass */
LGBCU_API int lgbcu_set_objective(LgbCuContext* ctx, int kind, double p0, double p1, double p2);
/* Multiclass: the learner keeps num_class planes of scores, gradients and hessians ([class][row], the reference's
 * ScoreUpdater / gradients_ layout; validation sets likewise).  lgbcu_set_class picks the plane that every other call
 * -- lgbcu_begin_tree / lgbcu_train_tree, the lgbcu_add_score_* family, lgbcu_copy_score, lgbcu_set_gradients, the
 * validation-set calls -- works on, so one tree per class and iteration is grown exactly as a single-class tree is
 * (GBDT::TrainOneIter's loop over cur_tree_id, src/boosting/gbdt.cpp:353-412).  Set the number of classes before any
 * validation set is added. */
/* l1 (regression_objective.hpp:214-231) and quantile (:491-516, p0 = alpha) have no second-order step worth keeping: after a
 * tree is grown (and before its scores are added) lgbcu_renew_leaf_outputs recomputes every leaf's value as the
 * (weighted) percentile of label - score over the rows the tree grew on -- SerialTreeLearner::RenewTreeOutput,
 * serial_tree_learner.cpp:661-699, with PercentileFun / WeightedPercentileFun (regression_objective.hpp:18-88).
 * alpha_is_float: the quantile objective holds alpha as a float and forms the position in float arithmetic.
 * leaf_nonempty[l] = the leaf holds rows on this rank (data-parallel runs average over the ranks that do). */
LGBCU_API int lgbcu_renew_leaf_outputs(LgbCuContext* ctx, int32_t num_leaves, double alpha, int32_t alpha_is_float,
                                       double* leaf_values, int32_t* leaf_nonempty);
LGBCU_API int lgbcu_set_num_class(LgbCuContext* ctx, int32_t num_class);
LGBCU_API int lgbcu_set_class(LgbCuContext* ctx, int32_t class_id);
/* multiclassova (multiclass_objective.hpp:186-276; lgbcu_set_objective p0 = num_class, p1 = sigmoid): one binary logloss
 * per class whose positive rows are (int)label == class; neg_pos = that class's (label weight of the negatives, of the
 * positives) pairs, [num_class][2] (is_unbalance / scale_pos_weight, binary_objective.hpp:76-96) */
LGBCU_API int lgbcu_set_ova_label_weights(LgbCuContext* ctx, const double* neg_pos, int32_t num_class);
LGBCU_API int lgbcu_set_label_gain(LgbCuContext* ctx, const double* label_gain, int32_t n);
LGBCU_API int lgbcu_get_gradients(LgbCuContext* ctx);                      /* score -> grad,hess (device) */
LGBCU_API int lgbcu_set_gradients(LgbCuContext* ctx, const float* grad, const float* hess, int32_t n);
LGBCU_API int lgbcu_copy_gradients(LgbCuContext* ctx, float* grad, float* hess, int32_t n);

/* A10: score (src/boosting/score_updater.hpp) */
LGBCU_API int lgbcu_set_score(LgbCuContext* ctx, const double* score, int32_t n);      /* NULL -> zeros */
LGBCU_API int lgbcu_add_score_const(LgbCuContext* ctx, double val);                    /* AddScore(double), :54-61 */
LGBCU_API int lgbcu_copy_score(LgbCuContext* ctx, double* score, int32_t n);
/* score[idx] += leaf_value[leaf(idx)] over the current partition
 * (SerialTreeLearner::AddPredictionToScore, src/treelearner/serial_tree_learner.h:96-111) */
LGBCU_API int lgbcu_add_score_by_partition(LgbCuContext* ctx, const double* leaf_value, int32_t num_leaves);

/* Tree::AddPredictionToScore on the resident binned matrix (src/io/tree.cpp:109-199; decision:
 * include/LightGBM/tree.h:285-300).  rows == NULL -> every row; otherwise the listed rows (out-of-bag rows in
 * GBDT::UpdateScore, src/boosting/gbdt.cpp:465-468).  rows_tag != 0 lets the library keep the uploaded row list:
 * a call with the same non-zero tag and count reuses it.  decision_type as in the model file
 * (bit 1 = default_left, bits 2-3 = missing type). */
LGBCU_API int lgbcu_add_score_by_tree(LgbCuContext* ctx, int32_t num_leaves, const int32_t* split_feature_inner,
                                      const int32_t* threshold_bin, const int8_t* decision_type,
                                      const int32_t* left_child, const int32_t* right_child, const double* leaf_value,
                                      const int32_t* rows, int32_t num_rows, int64_t rows_tag);

/* F3: bagging on the device (GBDT::Bagging + BaggingHelper, src/boosting/gbdt.cpp:162-181, 209-243): one LCG per
 * 1024-row block (include/LightGBM/utils/random.h), seeded seed + block by lgbcu_bagging_reset (ResetBaggingConfig,
 * gbdt.cpp:739-797); lgbcu_bagging draws the next bag -- a row is in the bag when its block's NextFloat() <
 * fraction -- and keeps it on the device: in-bag rows ascending, then out-of-bag rows ascending.  Pass
 * LGBCU_DEVICE_BAG as lgbcu_begin_tree's bag_indices / LGBCU_DEVICE_OOB as lgbcu_add_score_by_tree's rows to use it.
 * lgbcu_copy_bag reads the whole list back (tests). */
#define LGBCU_DEVICE_BAG ((const int32_t*)(intptr_t)-1)
#define LGBCU_DEVICE_OOB ((const int32_t*)(intptr_t)-1)
LGBCU_API int lgbcu_bagging_reset(LgbCuContext* ctx, int32_t seed);
LGBCU_API int lgbcu_bagging(LgbCuContext* ctx, double fraction, int32_t* bag_count);
/* the same draw with the fraction chosen by the row's label: > 0 -> pos_fraction, else neg_fraction (BalancedBaggingHelper,
 * src/boosting/gbdt.cpp:183-207); needs the labels on the device (lgbcu_set_label) */
LGBCU_API int lgbcu_bagging_balanced(LgbCuContext* ctx, double pos_fraction, double neg_fraction, int32_t* bag_count);
LGBCU_API int lgbcu_copy_bag(LgbCuContext* ctx, int32_t* out, int32_t n);

/* GOSS (replaces the device-visible half of GOSS::BaggingHelper / GOSS::Bagging, src/boosting/goss.hpp:103-177).  The
 * per-row decision depends on how many rows of the same thread chunk were sampled before it, so the walk stays on the
 * host; the device hands out |gradient * hessian| per row (float, formed as the reference forms it) ...            */
LGBCU_API int lgbcu_goss_scores(LgbCuContext* ctx, float* out_host, int32_t n);
/* ... and applies the outcome: bit r of word r/32 of `inbag_bits` = row r is in the bag, of `amplify_bits` = row r is a
 * sampled small-gradient row whose gradient and hessian are multiplied by chunk_multiply[r / chunk_rows].  The bag
 * (in-bag rows ascending, then the others) is built on the device like lgbcu_bagging's; both bit arrays hold
 * ceil(n / 1024) * 32 words.  Needs lgbcu_bagging_reset first.                                                       */
LGBCU_API int lgbcu_goss_apply(LgbCuContext* ctx, const uint32_t* inbag_bits, const uint32_t* amplify_bits, int32_t chunk_rows,
                               const float* chunk_multiply, int32_t nchunks, int32_t* bag_count);

/* A8: BeforeTrain (src/treelearner/serial_tree_learner.cpp:244-272): reset partition / leaf table / pool,
 * root sums.  bag_indices==NULL -> all rows; feature_used==NULL -> all features (ColSampler by-tree mask). */
LGBCU_API int lgbcu_begin_tree(LgbCuContext* ctx, const int32_t* bag_indices, int32_t bag_count,
                               const int8_t* feature_used);

/* A3+A4+A5: ConstructHistograms for `leaf` into its pool slot (Dataset::ConstructHistograms,
 * src/io/dataset.cpp:1282-1383; DenseBin::ConstructHistogramInner, src/io/dense_bin.hpp:97-140). */
LGBCU_API int lgbcu_hist_build(LgbCuContext* ctx, int32_t leaf);
/* FixHistogram (src/io/dataset.cpp:1410-1428) applied; output = hist_t[2*total_bin] in the reference layout */
LGBCU_API int lgbcu_hist_read(LgbCuContext* ctx, int32_t leaf, double* out, int64_t out_len);
LGBCU_API int lgbcu_num_total_bin(LgbCuContext* ctx, int32_t* out);
/* A5: larger = parent - smaller in the parent's slot (FeatureHistogram::Subtract, feature_histogram.hpp:79-83) */
LGBCU_API int lgbcu_hist_subtract(LgbCuContext* ctx, int32_t parent_leaf_slot_owner, int32_t smaller_leaf);

/* A6: FindBestThreshold over all used features of `leaf` + per-leaf argmax
 * (feature_histogram.hpp:85-95,853-1049; serial_tree_learner.cpp:358-425,701-736). */
LGBCU_API int lgbcu_find_best_split(LgbCuContext* ctx, int32_t leaf, LgbCuSplitInfo* out);

/* A7: DataPartition::Split (src/treelearner/data_partition.hpp:101-120 -> dense_bin.hpp:173-307). */
LGBCU_API int lgbcu_partition(LgbCuContext* ctx, int32_t leaf, int32_t inner_feature, int32_t threshold,
                              int32_t default_left, int32_t right_leaf, int32_t* left_count);
LGBCU_API int lgbcu_leaf_indices(LgbCuContext* ctx, int32_t leaf, int32_t* out, int32_t capacity, int32_t* count);
/* set LeafSplits sums of a leaf by hand (step-wise tests; SplitInner does it in train_tree) */
LGBCU_API int lgbcu_set_leaf_sums(LgbCuContext* ctx, int32_t leaf, double sum_g, double sum_h, double output);
LGBCU_API int lgbcu_get_leaf_sums(LgbCuContext* ctx, int32_t leaf, double* sum_g, double* sum_h, int32_t* count);

/* A8: the whole SerialTreeLearner::Train loop (serial_tree_learner.cpp:152-202) with device-side control:
 * no host round trip per split.  Call lgbcu_begin_tree first.  records has room for num_leaves-1 entries. */
LGBCU_API int lgbcu_train_tree(LgbCuContext* ctx, LgbCuSplitRecord* records, int32_t* num_splits);

/* A11: data-parallel (src/treelearner/data_parallel_tree_learner.cpp) -- communicator plumbing.
 * unique_id is NCCL's 128-byte id; rank 0 creates it, the caller broadcasts it (torch.distributed / MPI / file). */
LGBCU_API int lgbcu_comm_unique_id(uint8_t* id128);
/* The communicator belongs to the rendezvous, not to the context (Network::Init is once per process in the reference,
 * src/network/network.cpp:45-58): the first context naming (id, rank, world, device) creates it, later ones attach to
 * the cached communicator; lgbcu_comm_destroy only detaches; lgbcu_comm_release_all destroys every cached
 * communicator (call it when no context uses one any more: LGBM_NetworkFree). */
LGBCU_API int lgbcu_comm_init(LgbCuContext* ctx, const uint8_t* id128, int32_t rank, int32_t world_size);
LGBCU_API int lgbcu_comm_destroy(LgbCuContext* ctx);
LGBCU_API int lgbcu_comm_release_all(void);
/* small host-side syncs of the data-parallel path (Network::GlobalSyncUpBySum / ByMean, include/LightGBM/network.h:
 * 160-260): in-place sum over ranks of n <= 64 doubles.  No-op on one rank. */
LGBCU_API int lgbcu_comm_allreduce_sum_f64(LgbCuContext* ctx, double* inout, int32_t n);

/* No counterpart in the reference (its DataPartition::Split allocates its own buffers on the host).
 * A stable two-way partition of `n` row indices by one feature column that allocates NOTHING: for callers that own
 * their input/output buffers but may not obtain device memory when they need the partition (a host running under
 * someone else's fixed GPU memory arena, or a latency path that cannot afford cudaMalloc's device-wide sync).
 * The kernel's working memory is a fixed array inside this library: every GPU that loads the library holds about
 * 0.53 MB for it until exit, and one call handles at most LGBCU_NOALLOC_MAX_ROWS = 4,194,304 rows (more is
 * refused with an error).  All pointers except `meta` are DEVICE pointers owned by the caller:
 *   bins_dev        row-major u8 matrix, `row_stride` bytes per row; `column` = byte of the split feature
 *   meta            (host) bin-mapper facts of that feature, as for lgbcu_set_dataset
 *   indices_in_dev  n row indices; indices_out_dev (a buffer that does not overlap it, or the call is refused) receives
 *                   the rows going left, in their original order, followed by the rows going right, in their order
 *   left_count_dev  one int32 on the device: number of rows that went left
 *   stream          cudaStream_t (NULL = default stream); the call only enqueues kernels and returns
 * Same decision rule as lgbcu_partition / DenseBin::SplitInner.  Calls share the built-in array: calls that can
 * overlap on one GPU race -- keep them on one stream.  The first call on a device resolves the array's address
 * (and caches it for the life of the process: a cudaDeviceReset() in between is not supported); make it before any
 * stream capture begins. */
#define LGBCU_NOALLOC_MAX_ROWS (1 << 22)
LGBCU_API int lgbcu_partition_rows_noalloc(const uint8_t* bins_dev, int64_t row_stride, int32_t column,
                                           const LgbCuFeatureMeta* meta, int32_t threshold, int32_t default_left,
                                           const int32_t* indices_in_dev, int32_t n, int32_t* indices_out_dev,
                                           int32_t* left_count_dev, void* stream);

/* ---- validation sets and evaluation metrics on the device --------------------------------
 * A validation set is a second binned matrix whose bins come from the TRAINING set's mappers (the reference refuses
 * anything else: GBDT::AddValidDataset -> Dataset::CheckAlign, src/boosting/gbdt.cpp:121-125); it owns a score vector
 * (ScoreUpdater, src/boosting/score_updater.hpp:21-125) that every new tree is added to by walking the tree on the
 * bins (Tree::AddPredictionToScore, src/io/tree.cpp:109-199). `set` = -1 names the training rows. */
LGBCU_API int lgbcu_valid_create(LgbCuContext* ctx, int32_t num_data, const uint8_t* bins_row_major, int32_t* out_set);
LGBCU_API int lgbcu_valid_set_label(LgbCuContext* ctx, int32_t set, const float* label, int32_t n);
LGBCU_API int lgbcu_valid_set_weight(LgbCuContext* ctx, int32_t set, const float* weight, int32_t n);
LGBCU_API int lgbcu_valid_set_query(LgbCuContext* ctx, int32_t set, const int32_t* boundaries, int32_t num_queries);
LGBCU_API int lgbcu_valid_set_score(LgbCuContext* ctx, int32_t set, const double* score, int32_t n);
LGBCU_API int lgbcu_valid_add_const(LgbCuContext* ctx, int32_t set, double value);
/* ScoreUpdater::MultiplyScore (score_updater.hpp:84-91) of the current class's scores: training data / a validation set */
LGBCU_API int lgbcu_multiply_score(LgbCuContext* ctx, double value);
LGBCU_API int lgbcu_valid_multiply_score(LgbCuContext* ctx, int32_t set, double value);
LGBCU_API int lgbcu_valid_add_tree(LgbCuContext* ctx, int32_t set, int32_t num_leaves, const int32_t* split_feature_inner,
                                   const int32_t* threshold_bin, const int8_t* decision_type, const int32_t* left_child,
                                   const int32_t* right_child, const double* leaf_value);
/* Categorical features (LgbCuFeatureMeta::bin_type = LGBCU_BIN_CATEGORICAL; at most 256 bins each).  A categorical
 * split sends a SET of bins left (FeatureHistogram::FindBestThresholdCategoricalInner, feature_histogram.hpp:277-513;
 * DenseBin::SplitCategorical, dense_bin.hpp:309-363).  In a split record of such a feature `threshold` is an index
 * into the tree's bin sets (8 x 32-bit words each, bit b = bin b), read with lgbcu_tree_cat_sets after lgbcu_train_tree
 * (out = NULL returns the count alone).
 * For the tree walks (lgbcu_add_score_by_tree, lgbcu_valid_add_tree) a categorical node has decision_type bit 0 set and
 * threshold_bin = the index of its set among those staged with lgbcu_set_tree_cat_sets. */
LGBCU_API int lgbcu_set_cat_params(LgbCuContext* ctx, double cat_smooth, double cat_l2, int32_t max_cat_threshold,
                                   int32_t min_data_per_group, int32_t max_cat_to_onehot);
/* Monotone constraints, basic method (src/treelearner/monotone_constraints.hpp:19-126): one of -1 / 0 / +1 per inner
 * feature (NULL: none) and Config::monotone_penalty.  Call after lgbcu_set_dataset. */
LGBCU_API int lgbcu_set_monotone(LgbCuContext* ctx, const int8_t* types, int32_t num_features, double penalty);
LGBCU_API int lgbcu_tree_cat_sets(LgbCuContext* ctx, uint32_t* out, int32_t cap_sets, int32_t* num_sets);
LGBCU_API int lgbcu_set_tree_cat_sets(LgbCuContext* ctx, const uint32_t* sets, int32_t num_sets);
LGBCU_API int lgbcu_valid_copy_score(LgbCuContext* ctx, int32_t set, double* out, int32_t n);

/* metrics: Metric::Eval of src/metric/regression_metric.hpp (l2, rmse), binary_metric.hpp (binary_logloss, auc),
 * rank_metric.hpp + dcg_calculator.cpp (ndcg@k).  lgbcu_eval_set_ndcg uploads what NDCGMetric::Init prepares on the
 * host: eval positions, 1/maxDCG per (query, position) (<= 0 marks an all-irrelevant query), optional query weights,
 * the label gains and the position discounts. */
#define LGBCU_METRIC_L2 0
#define LGBCU_METRIC_RMSE 1
#define LGBCU_METRIC_BINARY_LOGLOSS 2
#define LGBCU_METRIC_AUC 3
#define LGBCU_METRIC_NDCG 4
#define LGBCU_METRIC_MULTI_LOGLOSS 5   /* `sigmoid` > 0: softmax the K raw scores first (the objective's ConvertOutput) */
#define LGBCU_METRIC_MULTI_ERROR 6     /* `sigmoid` = +top_k (softmax first) or -top_k (scores as they are) */
#define LGBCU_METRIC_L1 7
#define LGBCU_METRIC_QUANTILE 8        /* `sigmoid` = alpha */
#define LGBCU_METRIC_HUBER 9           /* `sigmoid` = alpha */
#define LGBCU_METRIC_FAIR 10           /* `sigmoid` = fair_c */
#define LGBCU_METRIC_POISSON 11        /* poisson / gamma / tweedie see exp(score) when the objective is one of the log-link ones */
#define LGBCU_METRIC_GAMMA 12
#define LGBCU_METRIC_TWEEDIE 13        /* `sigmoid` = tweedie_variance_power */
#define LGBCU_METRIC_BINARY_ERROR 14   /* `sigmoid` as for binary_logloss */
#define LGBCU_METRIC_MAPE 16
#define LGBCU_METRIC_MAP 18            /* map@k per eval position (map_metric.hpp); uses the tables of lgbcu_eval_set_ndcg */
#define LGBCU_METRIC_GAMMA_DEVIANCE 19
#define LGBCU_METRIC_XENTLAMBDA 17     /* cross_entropy_lambda: the weight sits inside the loss, the mean is over the rows */
#define LGBCU_METRIC_XENTROPY 15       /* cross_entropy (labels in [0, 1]); `sigmoid` as for binary_logloss.  kullback_leibler is
                                        * this plus a constant of the labels, which the host adds */
LGBCU_API int lgbcu_eval_set_ndcg(LgbCuContext* ctx, int32_t set, const int32_t* eval_at, int32_t num_eval,
                                  const double* inv_max_dcg, const float* query_weight, const double* label_gain,
                                  int32_t num_label_gain, const double* discount, int32_t num_discount);
/* out[0] for the point-wise metrics and auc, out[0..num_eval) for ndcg; `sigmoid` is used by binary_logloss only */
LGBCU_API int lgbcu_eval(LgbCuContext* ctx, int32_t set, int32_t kind, double sigmoid, double* out, int32_t out_len);

/* instrumentation for bench.py: device-time (ms, CUDA events on the context's stream) spent in the histogram
 * kernel since the last reset, its launch count and the algorithmic bytes it processed
 * (n*(G+8) [+4n if indexed] + total_bin*16 per call). */
LGBCU_API int lgbcu_profile_enable(LgbCuContext* ctx, int on);
LGBCU_API int lgbcu_profile_read(LgbCuContext* ctx, double* hist_ms, int64_t* hist_calls, double* hist_bytes,
                                 int reset);

/* profile mode: milliseconds of the LAST tree spent in each phase (CUDA events around every launch of the
 * per-round sequence): [0] histogram (+ all-reduce) [2] split search + control [4] partition; [6] rounds enqueued for
 * the last tree, [7] rounds it needed; other slots 0 */
LGBCU_API int lgbcu_profile_phases(LgbCuContext* ctx, double* ms8);

#ifdef __cplusplus
}
#endif
#endif  /* LGBCU_H_ */
