/*
 * lgbm_b200.h -- the drop-in boundary: the subset of the reference's C API (include/LightGBM/c_api.h) that sits in
 * front of the training hot path, exported by lib_lightgbm_b200.so with the SAME names, argument order, types
 * and 0 / -1 + LGBM_GetLastError convention (c_api.h:1300-1312, src/c_api.cpp:35-49), so that a caller bound to
 * lib_lightgbm.so (python-package/lightgbm/basic.py, R, SWIG, tests/c_api_test/test_.py) can load this library
 * instead.  Each declaration cites the reference declaration it replaces.
 *
 * Behavioural contract: same parameter strings (aliases, defaults) for the keys the path reads; `device_type` is
 * accepted with cpu|gpu|cuda and always selects the CUDA learner (there is no CPU path); `tree_learner` is
 * serial|data.  Parameters that would change the trees but are outside the path (categorical features, monotone
 * constraints, dart/goss/rf, multiclass, ...) fail with a message instead of silently training something else.
 */
#ifndef LGBM_B200_H_
#define LGBM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LGBM_B200_API __attribute__((visibility("default")))

typedef void* DatasetHandle; /* c_api.h:23 */
typedef void* BoosterHandle; /* c_api.h:24 */

#define C_API_DTYPE_FLOAT32 (0) /* c_api.h:27-30 */
#define C_API_DTYPE_FLOAT64 (1)
#define C_API_DTYPE_INT32 (2)
#define C_API_DTYPE_INT64 (3)
#define C_API_PREDICT_NORMAL (0)    /* c_api.h:32-35 */
#define C_API_PREDICT_RAW_SCORE (1)
#define C_API_PREDICT_LEAF_INDEX (2)

/* c_api.h:47 */
LGBM_B200_API const char* LGBM_GetLastError(void);
/* c_api.h:54  log lines go to the callback instead of stdout (thread-local, like the reference's Log) */
LGBM_B200_API int LGBM_RegisterLogCallback(void (*callback)(const char*));

/* c_api.h:229-236  (src/c_api.cpp:998-1015 -> :1018-1100) */
LGBM_B200_API int LGBM_DatasetCreateFromMat(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                                            const char* parameters, const DatasetHandle reference, DatasetHandle* out);
/* c_api.h:160-170, 205-215  (src/c_api.cpp:1102-1171, 1239-1318): sparse input.  The matrix is densified on the host (absent
 * entries = 0.0) and takes the LGBM_DatasetCreateFromMat path: the same bin mappers, bundles and bins. */
LGBM_B200_API int LGBM_DatasetCreateFromCSR(const void* indptr, int indptr_type, const int32_t* indices, const void* data,
                                            int data_type, int64_t nindptr, int64_t nelem, int64_t num_col, const char* parameters,
                                            const DatasetHandle reference, DatasetHandle* out);
LGBM_B200_API int LGBM_DatasetCreateFromCSC(const void* col_ptr, int col_ptr_type, const int32_t* indices, const void* data,
                                            int data_type, int64_t ncol_ptr, int64_t nelem, int64_t num_row, const char* parameters,
                                            const DatasetHandle reference, DatasetHandle* out);
/* c_api.h:750-764  (src/c_api.cpp:1893-1943) */
LGBM_B200_API int LGBM_BoosterPredictForCSR(BoosterHandle handle, const void* indptr, int indptr_type, const int32_t* indices,
                                            const void* data, int data_type, int64_t nindptr, int64_t nelem, int64_t num_col,
                                            int predict_type, int start_iteration, int num_iteration, const char* parameter,
                                            int64_t* out_len, double* out_result);
/* c_api.h:240-255  (src/c_api.cpp:1320-1350): the listed rows (ascending) of a dataset, with its bin mappers and metadata */
LGBM_B200_API int LGBM_DatasetGetSubset(const DatasetHandle handle, const int32_t* used_row_indices, int32_t num_used_row_indices,
                                        const char* parameters, DatasetHandle* out);
/* c_api.h:344-348  (src/c_api.cpp:1422-1439) field_name: label | weight | group/query | init_score */
LGBM_B200_API int LGBM_DatasetSetField(DatasetHandle handle, const char* field_name, const void* field_data,
                                       int num_element, int type);
/* c_api.h:380-381 */
LGBM_B200_API int LGBM_DatasetGetNumData(DatasetHandle handle, int* out);
/* c_api.h:389-390 */
LGBM_B200_API int LGBM_DatasetGetNumFeature(DatasetHandle handle, int* out);
/* c_api.h:311 */
LGBM_B200_API int LGBM_DatasetFree(DatasetHandle handle);

/* c_api.h:410-412  (src/c_api.cpp:1501-1509) */
LGBM_B200_API int LGBM_BoosterCreate(const DatasetHandle train_data, const char* parameters, BoosterHandle* out);
/* c_api.h:441 */
LGBM_B200_API int LGBM_BoosterFree(BoosterHandle handle);
/* c_api.h:505-506  (src/c_api.cpp:1598-1607) */
LGBM_B200_API int LGBM_BoosterUpdateOneIter(BoosterHandle handle, int* is_finished);
/* c_api.h:530-533  (src/c_api.cpp:1609-1625) custom objective: host float gradients / hessians */
LGBM_B200_API int LGBM_BoosterUpdateOneIterCustom(BoosterHandle handle, const float* grad, const float* hess, int* is_finished);
/* c_api.h:548-549 */
LGBM_B200_API int LGBM_BoosterGetCurrentIteration(BoosterHandle handle, int* out_iteration);
/* c_api.h:557-558 */
LGBM_B200_API int LGBM_BoosterNumModelPerIteration(BoosterHandle handle, int* out_tree_per_iteration);
/* c_api.h:566-567 */
LGBM_B200_API int LGBM_BoosterNumberOfTotalModel(BoosterHandle handle, int* out_models);
/* c_api.h:496-497 */
LGBM_B200_API int LGBM_BoosterGetNumClasses(BoosterHandle handle, int* out_len);
/* c_api.h:469-470  the data set must have been created with `reference` = the training data (same bin mappers);
 * its scores live on the device and every new tree is added to them */
LGBM_B200_API int LGBM_BoosterAddValidData(BoosterHandle handle, const DatasetHandle valid_data);
/* c_api.h:575-576, 590-595, 636-639  metrics: l2, rmse, binary_logloss, auc, ndcg@k (evaluated on the device);
 * data_idx 0 = training data, i = i-th validation set */
LGBM_B200_API int LGBM_BoosterGetEvalCounts(BoosterHandle handle, int* out_len);
LGBM_B200_API int LGBM_BoosterGetEvalNames(BoosterHandle handle, const int len, int* out_len, const size_t buffer_len,
                                           size_t* out_buffer_len, char** out_strs);
LGBM_B200_API int LGBM_BoosterGetEval(BoosterHandle handle, int data_idx, int* out_len, double* out_results);
/* c_api.h:649-651  data_idx 0 = training data, i = i-th validation set */
LGBM_B200_API int LGBM_BoosterGetNumPredict(BoosterHandle handle, int data_idx, int64_t* out_len);
/* c_api.h:663-666  scores of the training / validation data after the objective's output transform */
LGBM_B200_API int LGBM_BoosterGetPredict(BoosterHandle handle, int data_idx, int64_t* out_len, double* out_result);
/* c_api.h:1007-1018  predict_type: C_API_PREDICT_NORMAL | C_API_PREDICT_RAW_SCORE (host tree walk; not on the hot path) */
LGBM_B200_API int LGBM_BoosterPredictForMat(BoosterHandle handle, const void* data, int data_type, int32_t nrow, int32_t ncol,
                                            int is_row_major, int predict_type, int start_iteration, int num_iteration,
                                            const char* parameter, int64_t* out_len, double* out_result);
/* c_api.h:1172-1178 */
LGBM_B200_API int LGBM_BoosterSaveModelToString(BoosterHandle handle, int start_iteration, int num_iteration,
                                                int feature_importance_type, int64_t buffer_len, int64_t* out_len,
                                                char* out_str);
/* c_api.h:1155-1159 */
LGBM_B200_API int LGBM_BoosterSaveModel(BoosterHandle handle, int start_iteration, int num_iteration,
                                        int feature_importance_type, const char* filename);

/* ---- what python-package/lightgbm/basic.py and engine.py reach around lgb.train on a numpy matrix ----------------- */
/* c_api.h:283-285 */
LGBM_B200_API int LGBM_DatasetSetFeatureNames(DatasetHandle handle, const char** feature_names, int num_feature_names);
/* c_api.h:299-304 */
LGBM_B200_API int LGBM_DatasetGetFeatureNames(DatasetHandle handle, const int len, int* num_feature_names,
                                              const size_t buffer_len, size_t* out_buffer_len, char** feature_names);
/* c_api.h:359-363  label | weight (float32), init_score (float64), group (int32 query boundaries) */
LGBM_B200_API int LGBM_DatasetGetField(DatasetHandle handle, const char* field_name, int* out_len, const void** out_ptr,
                                       int* out_type);
/* c_api.h:371-372  refuses changes of the Dataset-construction parameters (src/c_api.cpp:1441-1451) */
LGBM_B200_API int LGBM_DatasetUpdateParamChecking(const char* old_parameters, const char* new_parameters);
/* c_api.h:432-434, 421-423  a booster read back from model text: predicts and saves, cannot train */
LGBM_B200_API int LGBM_BoosterLoadModelFromString(const char* model_str, int* out_num_iterations, BoosterHandle* out);
LGBM_B200_API int LGBM_BoosterCreateFromModelfile(const char* filename, int* out_num_iterations, BoosterHandle* out);
/* c_api.h:487-488  learning rate and tree-growing parameters */
LGBM_B200_API int LGBM_BoosterResetParameter(BoosterHandle handle, const char* parameters);
/* c_api.h:540  (GBDT::RollbackOneIter, src/boosting/gbdt.cpp:421-444) */
LGBM_B200_API int LGBM_BoosterRollbackOneIter(BoosterHandle handle);
/* c_api.h:622-623, 609-614 */
LGBM_B200_API int LGBM_BoosterGetNumFeature(BoosterHandle handle, int* out_len);
LGBM_B200_API int LGBM_BoosterGetFeatureNames(BoosterHandle handle, const int len, int* out_len, const size_t buffer_len,
                                              size_t* out_buffer_len, char** out_strs);
/* c_api.h:1235-1238  importance_type 0 = split counts, 1 = total gains */
LGBM_B200_API int LGBM_BoosterFeatureImportance(BoosterHandle handle, int num_iteration, int importance_type,
                                                double* out_results);
/* c_api.h:707-712 */
LGBM_B200_API int LGBM_BoosterCalcNumPredict(BoosterHandle handle, int num_row, int predict_type, int start_iteration,
                                             int num_iteration, int64_t* out_len);
/* c_api.h:1266-1269, 1285-1288  the socket / external-collective linkers: with num_machines > 1 these fail with a message
 * that points at LGBM_NetworkInitNCCL (below); with num_machines <= 1 they are no-ops, as in the reference */
LGBM_B200_API int LGBM_NetworkInit(const char* machines, int local_listen_port, int listen_time_out, int num_machines);
LGBM_B200_API int LGBM_NetworkInitWithFunctions(int num_machines, int rank, void* reduce_scatter_ext_fun,
                                                void* allgather_ext_fun);

/* Data-parallel rendezvous inside one NVLink box.  Replaces LGBM_NetworkInit (c_api.h:1266-1269: machine list +
 * TCP ports) for tree_learner=data: the ranks share a 128-byte NCCL id instead of a socket mesh.  Rank 0 calls
 * LGBM_NetworkGetNCCLUniqueId and hands the bytes to the other ranks by any means (torch.distributed, MPI, a
 * file); every rank then calls LGBM_NetworkInitNCCL before LGBM_BoosterCreate.  State is thread-local, like the
 * reference's Network (src/network/network.cpp:17-27). */
LGBM_B200_API int LGBM_NetworkGetNCCLUniqueId(uint8_t* out_id128);
LGBM_B200_API int LGBM_NetworkInitNCCL(const uint8_t* id128, int rank, int num_machines);
/* c_api.h:1275 */
LGBM_B200_API int LGBM_NetworkFree(void);

/* ours: how many kernels this library has launched in this process (bench.py's gpu_launches) */
LGBM_B200_API int64_t LGBM_B200_LaunchCount(void);
/* ours: histogram-kernel instrumentation of a booster (device ms, calls, algorithmic bytes); see lgbcu.h */
LGBM_B200_API int LGBM_B200_BoosterProfile(BoosterHandle handle, int enable, int reset, double* hist_ms,
                                           int64_t* hist_calls, double* hist_bytes);
/* profile mode: per-phase milliseconds and round counts of the last tree (lgbcu_profile_phases) */
LGBM_B200_API int LGBM_B200_BoosterPhases(BoosterHandle handle, double* ms8);

/* ours: introspection used by the parity tests (the reference exposes the same facts through
 * LGBM_DatasetDumpText, src/io/dataset.cpp:994-1044).  out_bins: [num_data][num_inner_features] BIN values
 * (BinIterator::Get, not the stored encoding), row-major; real_index: real column of every inner feature. */
LGBM_B200_API int LGBM_B200_DatasetNumInnerFeatures(DatasetHandle handle, int* out);
LGBM_B200_API int LGBM_B200_DatasetDumpBins(DatasetHandle handle, int32_t* out_bins, int32_t* real_index);
/* ours: bin mapper of one inner feature: meta[0..4] = num_bin, most_freq_bin, default_bin, missing_type, real index;
 * upper_bounds receives min(num_bin, cap) doubles */
LGBM_B200_API int LGBM_B200_DatasetFeatureInfo(DatasetHandle handle, int inner_feature, int32_t* meta5,
                                               double* upper_bounds, int cap);

#ifdef __cplusplus
}
#endif
#endif /* LGBM_B200_H_ */
