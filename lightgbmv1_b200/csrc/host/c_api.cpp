// The drop-in C API (include/lgbm_b200.h): same entry points as the reference's src/c_api.cpp for the training
// path, backed by the CUDA learner.  API_BEGIN/API_END convention of src/c_api.cpp:35-49.
#include "lgbm_b200.h"

#include <omp.h>

#include <cstring>
#include <fstream>
#include <memory>
#include <mutex>

#include "common.h"
#include "config.h"
#include "dataset.h"
#include "gbdt.h"

using namespace lgb;

namespace {

thread_local char g_last_error[512] = "Everything is fine";

void SetLastError(const char* msg) { std::snprintf(g_last_error, sizeof(g_last_error), "%s", msg); }

struct Booster {
  std::unique_ptr<GBDT> gbdt;
  std::mutex mu;  // training calls are serialised (the reference holds a writer lock, src/c_api.cpp:346-349)
};

}  // namespace

#define API_BEGIN() try {
#define API_END()                                                       \
  }                                                                     \
  catch (std::exception & ex) { SetLastError(ex.what()); return -1; }   \
  catch (std::string & ex) { SetLastError(ex.c_str()); return -1; }     \
  catch (...) { SetLastError("unknown exception"); return -1; }         \
  return 0;

extern "C" {

const char* LGBM_GetLastError(void) { return g_last_error; }

int LGBM_DatasetCreateFromMat(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                              const char* parameters, const DatasetHandle reference, DatasetHandle* out) {
  API_BEGIN();
  if (!data || !out) Log::Fatal("LGBM_DatasetCreateFromMat: NULL argument");
  Config config;
  config.Set(Config::Str2Map(parameters));
  if (config.num_threads > 0) omp_set_num_threads(config.num_threads);
  if (reference == nullptr) {
    *out = Dataset::FromDenseMatrix(data, data_type, nrow, ncol, is_row_major != 0, config);
  } else {
    *out = Dataset::FromDenseMatrixAligned(data, data_type, nrow, ncol, is_row_major != 0,
                                           *reinterpret_cast<const Dataset*>(reference), config);
  }
  API_END();
}

int LGBM_DatasetSetField(DatasetHandle handle, const char* field_name, const void* field_data, int num_element, int type) {
  API_BEGIN();
  if (!handle || !field_name) Log::Fatal("LGBM_DatasetSetField: NULL argument");
  Dataset* ds = reinterpret_cast<Dataset*>(handle);
  if (type == C_API_DTYPE_FLOAT32) ds->SetFloatField(field_name, reinterpret_cast<const float*>(field_data), num_element);
  else if (type == C_API_DTYPE_FLOAT64) ds->SetDoubleField(field_name, reinterpret_cast<const double*>(field_data), num_element);
  else if (type == C_API_DTYPE_INT32) ds->SetIntField(field_name, reinterpret_cast<const int32_t*>(field_data), num_element);
  else Log::Fatal("Input data type error or field not found");
  API_END();
}

int LGBM_DatasetGetNumData(DatasetHandle handle, int* out) {
  API_BEGIN();
  *out = reinterpret_cast<Dataset*>(handle)->num_data();
  API_END();
}

int LGBM_DatasetGetNumFeature(DatasetHandle handle, int* out) {
  API_BEGIN();
  *out = reinterpret_cast<Dataset*>(handle)->num_total_features();
  API_END();
}

int LGBM_DatasetFree(DatasetHandle handle) {
  API_BEGIN();
  delete reinterpret_cast<Dataset*>(handle);
  API_END();
}

int LGBM_BoosterCreate(const DatasetHandle train_data, const char* parameters, BoosterHandle* out) {
  API_BEGIN();
  if (!train_data || !out) Log::Fatal("LGBM_BoosterCreate: NULL argument");
  Config config;
  config.Set(Config::Str2Map(parameters));
  if (config.num_threads > 0) omp_set_num_threads(config.num_threads);
  // src/c_api.cpp:132-138: without an initialised network the learner is serial whatever was asked
  if (!Net().active || Net().num_machines <= 1) config.tree_learner = "serial";
  std::unique_ptr<Booster> b(new Booster());
  b->gbdt.reset(new GBDT(reinterpret_cast<const Dataset*>(train_data), config));
  *out = b.release();
  API_END();
}

int LGBM_BoosterFree(BoosterHandle handle) {
  API_BEGIN();
  delete reinterpret_cast<Booster*>(handle);
  API_END();
}

int LGBM_BoosterUpdateOneIter(BoosterHandle handle, int* is_finished) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mu);
  *is_finished = b->gbdt->TrainOneIter(nullptr, nullptr) ? 1 : 0;
  API_END();
}

int LGBM_BoosterUpdateOneIterCustom(BoosterHandle handle, const float* grad, const float* hess, int* is_finished) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mu);
  if (!grad || !hess) Log::Fatal("LGBM_BoosterUpdateOneIterCustom: NULL gradients");
  *is_finished = b->gbdt->TrainOneIter(grad, hess) ? 1 : 0;
  API_END();
}

int LGBM_BoosterGetCurrentIteration(BoosterHandle handle, int* out_iteration) {
  API_BEGIN();
  *out_iteration = reinterpret_cast<Booster*>(handle)->gbdt->current_iteration();
  API_END();
}

int LGBM_BoosterNumModelPerIteration(BoosterHandle handle, int* out_tree_per_iteration) {
  API_BEGIN();
  (void)handle;
  *out_tree_per_iteration = 1;
  API_END();
}

int LGBM_BoosterNumberOfTotalModel(BoosterHandle handle, int* out_models) {
  API_BEGIN();
  *out_models = reinterpret_cast<Booster*>(handle)->gbdt->current_iteration();
  API_END();
}

int LGBM_BoosterGetNumClasses(BoosterHandle handle, int* out_len) {
  API_BEGIN();
  (void)handle;
  *out_len = 1;
  API_END();
}

int LGBM_BoosterGetNumPredict(BoosterHandle handle, int data_idx, int64_t* out_len) {
  API_BEGIN();
  GBDT* g = reinterpret_cast<Booster*>(handle)->gbdt.get();
  if (data_idx < 0 || data_idx > g->num_valid()) Log::Fatal("Check failed: data_idx >= 0 && data_idx <= number of validation sets");
  *out_len = data_idx == 0 ? g->num_data() : g->valid_num_data(data_idx - 1);
  API_END();
}

int LGBM_BoosterGetPredict(BoosterHandle handle, int data_idx, int64_t* out_len, double* out_result) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mu);
  if (data_idx < 0 || data_idx > b->gbdt->num_valid()) Log::Fatal("Check failed: data_idx >= 0 && data_idx <= number of validation sets");
  if (data_idx == 0) {
    b->gbdt->GetTrainScore(out_result, true);
    *out_len = b->gbdt->num_data();
  } else {
    b->gbdt->GetValidScore(data_idx - 1, out_result, true);
    *out_len = b->gbdt->valid_num_data(data_idx - 1);
  }
  API_END();
}

int LGBM_BoosterAddValidData(BoosterHandle handle, const DatasetHandle valid_data) {
  API_BEGIN();
  if (!valid_data) Log::Fatal("validation data handle is NULL");
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mu);
  b->gbdt->AddValidData(reinterpret_cast<const Dataset*>(valid_data));
  API_END();
}

int LGBM_BoosterGetEvalCounts(BoosterHandle handle, int* out_len) {
  API_BEGIN();
  *out_len = static_cast<int>(reinterpret_cast<Booster*>(handle)->gbdt->EvalNames().size());
  API_END();
}

int LGBM_BoosterGetEvalNames(BoosterHandle handle, const int len, int* out_len, const size_t buffer_len,
                             size_t* out_buffer_len, char** out_strs) {
  API_BEGIN();
  const std::vector<std::string> names = reinterpret_cast<Booster*>(handle)->gbdt->EvalNames();
  *out_buffer_len = 0;
  int idx = 0;
  for (const std::string& name : names) {  // Booster::GetEvalNames (src/c_api.cpp:764-780)
    if (idx < len && buffer_len > 0) {
      std::memcpy(out_strs[idx], name.c_str(), std::min(name.size() + 1, buffer_len));
      out_strs[idx][buffer_len - 1] = '\0';
    }
    *out_buffer_len = std::max(name.size() + 1, *out_buffer_len);
    ++idx;
  }
  *out_len = idx;
  API_END();
}

int LGBM_BoosterGetEval(BoosterHandle handle, int data_idx, int* out_len, double* out_results) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mu);
  const std::vector<double> r = b->gbdt->GetEvalAt(data_idx);
  *out_len = static_cast<int>(r.size());
  for (size_t i = 0; i < r.size(); ++i) out_results[i] = r[i];
  API_END();
}

int LGBM_BoosterPredictForMat(BoosterHandle handle, const void* data, int data_type, int32_t nrow, int32_t ncol,
                              int is_row_major, int predict_type, int start_iteration, int num_iteration,
                              const char* parameter, int64_t* out_len, double* out_result) {
  API_BEGIN();
  (void)parameter;
  if (start_iteration != 0) Log::Fatal("start_iteration != 0 is outside this library's path");
  if (predict_type != C_API_PREDICT_NORMAL && predict_type != C_API_PREDICT_RAW_SCORE)
    Log::Fatal("predict_type %d is outside this library's path (normal, raw score)", predict_type);
  Booster* b = reinterpret_cast<Booster*>(handle);
  b->gbdt->PredictMat(data, data_type, nrow, ncol, is_row_major != 0, predict_type == C_API_PREDICT_RAW_SCORE,
                      num_iteration, out_result);
  *out_len = nrow;
  API_END();
}

int LGBM_BoosterSaveModelToString(BoosterHandle handle, int start_iteration, int num_iteration, int feature_importance_type,
                                  int64_t buffer_len, int64_t* out_len, char* out_str) {
  API_BEGIN();
  (void)feature_importance_type;
  const std::string model = reinterpret_cast<Booster*>(handle)->gbdt->SaveModelToString(start_iteration, num_iteration);
  *out_len = static_cast<int64_t>(model.size()) + 1;
  if (*out_len <= buffer_len) std::memcpy(out_str, model.c_str(), *out_len);
  API_END();
}

int LGBM_BoosterSaveModel(BoosterHandle handle, int start_iteration, int num_iteration, int feature_importance_type,
                          const char* filename) {
  API_BEGIN();
  (void)feature_importance_type;
  const std::string model = reinterpret_cast<Booster*>(handle)->gbdt->SaveModelToString(start_iteration, num_iteration);
  std::ofstream f(filename, std::ios::binary);
  if (!f) Log::Fatal("Model file %s is not available for writes", filename);
  f.write(model.data(), static_cast<std::streamsize>(model.size()));
  API_END();
}

int LGBM_NetworkGetNCCLUniqueId(uint8_t* out_id128) {
  API_BEGIN();
  if (lgbcu_comm_unique_id(out_id128) != 0) Log::Fatal("%s", lgbcu_last_error());
  API_END();
}

int LGBM_NetworkInitNCCL(const uint8_t* id128, int rank, int num_machines) {
  API_BEGIN();
  if (!id128 || num_machines < 1 || rank < 0 || rank >= num_machines) Log::Fatal("LGBM_NetworkInitNCCL: bad argument");
  NetworkState& n = Net();
  std::memcpy(n.unique_id, id128, 128);
  n.rank = rank; n.num_machines = num_machines; n.active = num_machines > 1;
  API_END();
}

int LGBM_NetworkFree(void) {
  API_BEGIN();
  lgbcu_comm_release_all();  // boosters that trained over it must have been freed (as with Network::Dispose)
  Net() = NetworkState();
  API_END();
}

int64_t LGBM_B200_LaunchCount(void) { return lgbcu_launch_count(); }

int LGBM_B200_BoosterProfile(BoosterHandle handle, int enable, int reset, double* hist_ms, int64_t* hist_calls,
                             double* hist_bytes) {
  API_BEGIN();
  LgbCuContext* ctx = reinterpret_cast<Booster*>(handle)->gbdt->device();
  if (lgbcu_profile_enable(ctx, enable) != 0) Log::Fatal("%s", lgbcu_last_error());
  if (lgbcu_profile_read(ctx, hist_ms, hist_calls, hist_bytes, reset) != 0) Log::Fatal("%s", lgbcu_last_error());
  API_END();
}

int LGBM_B200_BoosterPhases(BoosterHandle handle, double* ms8) {
  API_BEGIN();
  LgbCuContext* ctx = reinterpret_cast<Booster*>(handle)->gbdt->device();
  if (lgbcu_profile_phases(ctx, ms8) != 0) Log::Fatal("%s", lgbcu_last_error());
  API_END();
}

int LGBM_B200_DatasetNumInnerFeatures(DatasetHandle handle, int* out) {
  API_BEGIN();
  *out = reinterpret_cast<Dataset*>(handle)->num_features();
  API_END();
}

int LGBM_B200_DatasetDumpBins(DatasetHandle handle, int32_t* out_bins, int32_t* real_index) {
  API_BEGIN();
  const Dataset* ds = reinterpret_cast<Dataset*>(handle);
  const int F = ds->num_features(), N = ds->num_data();
  for (int f = 0; f < F; ++f) real_index[f] = ds->real_feature_index(f);
  for (int r = 0; r < N; ++r) {
    for (int f = 0; f < F; ++f) {
      const BinMapper& m = ds->mapper(f);
      const int stored = ds->bins()[static_cast<size_t>(r) * F + f];
      const int offset = m.most_freq_bin() == 0 ? 1 : 0;
      out_bins[static_cast<size_t>(r) * F + f] = stored == 0 ? static_cast<int>(m.most_freq_bin()) : stored - 1 + offset;
    }
  }
  API_END();
}

int LGBM_B200_DatasetFeatureInfo(DatasetHandle handle, int inner_feature, int32_t* meta5, double* upper_bounds, int cap) {
  API_BEGIN();
  const Dataset* ds = reinterpret_cast<Dataset*>(handle);
  if (inner_feature < 0 || inner_feature >= ds->num_features()) Log::Fatal("inner feature %d out of range", inner_feature);
  const BinMapper& m = ds->mapper(inner_feature);
  meta5[0] = m.num_bin(); meta5[1] = static_cast<int>(m.most_freq_bin()); meta5[2] = static_cast<int>(m.default_bin());
  meta5[3] = m.missing_type(); meta5[4] = ds->real_feature_index(inner_feature);
  for (int i = 0; i < m.num_bin() && i < cap; ++i) upper_bounds[i] = m.upper_bounds()[i];
  API_END();
}

}  // extern "C"
