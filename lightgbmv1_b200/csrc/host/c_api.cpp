// The drop-in C API (include/lgbm_b200.h): same entry points as the reference's src/c_api.cpp for the training
// path, backed by the CUDA learner.  API_BEGIN/API_END convention of src/c_api.cpp:35-49.
#include "lgbm_b200.h"

#include <omp.h>

#include <cstring>
#include <cmath>
#include <fstream>
#include <sstream>
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

// A model read back from its text (LGBM_BoosterLoadModelFromString / CreateFromModelfile, GBDT::LoadModelFromString,
// src/boosting/gbdt_model_text.cpp:399-560): trees + header, good for prediction and for writing the text again.
struct LoadedModel {
  std::string text;
  std::vector<std::unique_ptr<Tree>> trees;
  std::vector<std::string> feature_names;
  int max_feature_idx = -1;
  std::string objective;   // the header's objective line, e.g. "binary sigmoid:1"
  double sigmoid = 0.0;    // > 0: predictions are 1 / (1 + exp(-sigmoid * raw))
  bool exp_out = false;    // the log-link regressions (poisson, gamma, tweedie): predictions are exp(raw)
  bool softplus_out = false;   // cross_entropy_lambda: predictions are log(1 + exp(raw))
  bool average_output = false; // boosting=rf: the header's "average_output" line
  bool square_out = false;     // "regression sqrt": predictions are sign(raw) * raw^2
  int num_tree_per_iteration = 1;   // multiclass: trees (and outputs per row) per iteration
  bool softmax = false;             // "multiclass ...": predictions are the softmax of a row's raw scores
  double ova_sigmoid = 0.0;         // "multiclassova ...": > 0, a sigmoid per class
};

struct Booster {
  std::unique_ptr<GBDT> gbdt;
  std::unique_ptr<LoadedModel> loaded;
  std::mutex mu;  // training calls are serialised (the reference holds a writer lock, src/c_api.cpp:346-349)
};

GBDT* Trainer(BoosterHandle handle) {
  Booster* b = reinterpret_cast<Booster*>(handle);
  if (!b || !b->gbdt) Log::Fatal("this booster was loaded from a model string: it can predict and be saved, not trained");
  return b->gbdt.get();
}

LoadedModel* ParseModel(const std::string& text) {
  std::unique_ptr<LoadedModel> m(new LoadedModel());
  m->text = text;
  size_t pos = 0;
  auto next_line = [&](std::string* line) {
    if (pos >= text.size()) return false;
    size_t e = text.find('\n', pos);
    if (e == std::string::npos) e = text.size();
    *line = text.substr(pos, e - pos);
    pos = e + 1;
    return true;
  };
  std::string line;
  bool in_header = true;
  while (in_header && next_line(&line)) {
    if (line.rfind("Tree=", 0) == 0) { pos -= line.size() + 1; break; }
    const size_t eq = line.find('=');
    if (eq == std::string::npos) { if (line == "average_output") m->average_output = true; continue; }
    const std::string k = line.substr(0, eq), v = line.substr(eq + 1);
    if (k == "max_feature_idx") m->max_feature_idx = std::atoi(v.c_str());
    else if (k == "objective") m->objective = v;
    else if (k == "feature_names") m->feature_names = SplitAny(v, " ");
    else if (k == "num_tree_per_iteration") m->num_tree_per_iteration = std::max(1, std::atoi(v.c_str()));
  }
  m->softmax = m->objective.rfind("multiclass", 0) == 0 && m->objective.rfind("multiclassova", 0) != 0;
  if (m->objective.rfind("multiclassova", 0) == 0) {   // a sigmoid per class
    m->ova_sigmoid = 1.0;
    const size_t p = m->objective.find("sigmoid:");
    if (p != std::string::npos) m->ova_sigmoid = std::strtod(m->objective.c_str() + p + 8, nullptr);
  }
  if (m->max_feature_idx < 0) Log::Fatal("Model file doesn't specify max_feature_idx");
  if (m->objective.rfind("binary", 0) == 0) {
    m->sigmoid = 1.0;
    const size_t p = m->objective.find("sigmoid:");
    if (p != std::string::npos) m->sigmoid = std::strtod(m->objective.c_str() + p + 8, nullptr);
  }
  if (m->objective.rfind("cross_entropy", 0) == 0 && m->objective.rfind("cross_entropy_lambda", 0) != 0) m->sigmoid = 1.0;
  if (m->objective.rfind("cross_entropy_lambda", 0) == 0) m->softplus_out = true;
  if (m->objective.rfind("regression", 0) == 0 && m->objective.find(" sqrt") != std::string::npos) m->square_out = true;
  if (m->objective.rfind("poisson", 0) == 0 || m->objective.rfind("gamma", 0) == 0 || m->objective.rfind("tweedie", 0) == 0) m->exp_out = true;
  // trees: "Tree=i" ... blank line, until "end of trees"
  std::string cur;
  bool in_tree = false;
  while (next_line(&line)) {
    if (line.rfind("end of trees", 0) == 0) break;
    if (line.rfind("Tree=", 0) == 0) { in_tree = true; cur.clear(); continue; }
    if (in_tree) {
      if (line.empty()) { m->trees.emplace_back(Tree::FromString(cur)); in_tree = false; }
      else { cur += line; cur += '\n'; }
    }
  }
  if (in_tree && !cur.empty()) m->trees.emplace_back(Tree::FromString(cur));
  return m.release();
}

}  // namespace

#define API_BEGIN() try {
#define API_END()                                                       \
  }                                                                     \
  catch (std::exception & ex) { SetLastError(ex.what()); return -1; }   \
  catch (std::string & ex) { SetLastError(ex.c_str()); return -1; }     \
  catch (...) { SetLastError("unknown exception"); return -1; }         \
  return 0;

namespace {

// Sparse input (LGBM_DatasetCreateFromCSR / CSC, src/c_api.cpp:1102-1171, 1239-1318) -> the dense row-major matrix the
// device path consumes: absent entries are 0.0.  The reference samples and pushes the same values either way (its
// per-column sample lists hold the non-zero sampled values in sample order for a dense, a CSR and a CSC source alike),
// so the dataset is the one LGBM_DatasetCreateFromMat builds from the densified matrix.  (One corner differs: in a CSR
// row the reference pushes the implied zeros of features whose zero bin is not their most frequent bin AFTER the
// stored entries, Dataset::FinishOneRow, which changes who wins a bundle conflict in such a row.)
int64_t IndptrAt(const void* p, int type, int64_t i) {
  return type == C_API_DTYPE_INT32 ? static_cast<int64_t>(static_cast<const int32_t*>(p)[i]) : static_cast<const int64_t*>(p)[i];
}

template <typename T>
void Densify(bool csr, const void* ptr, int ptr_type, const int32_t* indices, const void* data, int data_type, int64_t nptr,
             int64_t nrow, int64_t ncol, int64_t row0, int64_t rows, T* out) {
  std::fill(out, out + static_cast<size_t>(rows) * ncol, T(0));
  int bad = 0;
  auto value = [&](int64_t k) -> T {
    return data_type == C_API_DTYPE_FLOAT32 ? static_cast<T>(static_cast<const float*>(data)[k]) : static_cast<T>(static_cast<const double*>(data)[k]);
  };
  if (csr) {
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < rows; ++r) {
      const int64_t b = IndptrAt(ptr, ptr_type, row0 + r), e = IndptrAt(ptr, ptr_type, row0 + r + 1);
      for (int64_t k = b; k < e; ++k) {
        const int32_t c = indices[k];
        if (c < 0 || c >= ncol) { bad = 1; continue; }   // (no exception may leave a parallel region)
        out[static_cast<size_t>(r) * ncol + c] = value(k);
      }
    }
  } else {
#pragma omp parallel for schedule(static)
    for (int64_t c = 0; c < nptr - 1; ++c) {
      const int64_t b = IndptrAt(ptr, ptr_type, c), e = IndptrAt(ptr, ptr_type, c + 1);
      for (int64_t k = b; k < e; ++k) {
        const int64_t r = indices[k];
        if (r < 0 || r >= nrow) { bad = 1; continue; }
        if (r >= row0 && r < row0 + rows) out[static_cast<size_t>(r - row0) * ncol + c] = value(k);
      }
    }
  }
  if (bad) Log::Fatal(csr ? "Check failed: (inner_data.first) < (num_col)" : "row index out of range in CSC input");
}

Dataset* DatasetFromSparse(bool csr, const void* ptr, int ptr_type, const int32_t* indices, const void* data, int data_type,
                           int64_t nptr, int64_t nrow, int64_t ncol, const char* parameters, const DatasetHandle reference) {
  if (!ptr || !indices || !data) Log::Fatal("LGBM_DatasetCreateFromCS%c: NULL argument", csr ? 'R' : 'C');
  if (ptr_type != C_API_DTYPE_INT32 && ptr_type != C_API_DTYPE_INT64) Log::Fatal("Unknown data type in RowFunctionFromCSR");
  if (data_type != C_API_DTYPE_FLOAT32 && data_type != C_API_DTYPE_FLOAT64) Log::Fatal("Unknown data type in RowFunctionFromCSR");
  if (ncol <= 0) Log::Fatal("The number of columns should be greater than zero.");
  if (ncol >= INT32_MAX) Log::Fatal("The number of columns should be smaller than INT32_MAX.");
  if (nrow <= 0 || nrow >= INT32_MAX) Log::Fatal("Check failed: (num_data) > (0)");
  const size_t cells = static_cast<size_t>(nrow) * static_cast<size_t>(ncol);
  const size_t esz = data_type == C_API_DTYPE_FLOAT32 ? 4 : 8;
  if (cells * esz > (static_cast<size_t>(96) << 30))
    Log::Fatal("sparse input of %lld x %lld would need %.1f GB as the dense matrix this library trains on", static_cast<long long>(nrow),
               static_cast<long long>(ncol), static_cast<double>(cells * esz) / 1073741824.0);
  Config config;
  config.Set(Config::Str2Map(parameters));
  if (config.num_threads > 0) omp_set_num_threads(config.num_threads);
  std::unique_ptr<char[]> dense(new char[cells * esz]);
  if (esz == 4) Densify<float>(csr, ptr, ptr_type, indices, data, data_type, nptr, nrow, ncol, 0, nrow, reinterpret_cast<float*>(dense.get()));
  else Densify<double>(csr, ptr, ptr_type, indices, data, data_type, nptr, nrow, ncol, 0, nrow, reinterpret_cast<double*>(dense.get()));
  if (reference == nullptr)
    return Dataset::FromDenseMatrix(dense.get(), data_type, static_cast<int32_t>(nrow), static_cast<int32_t>(ncol), true, config);
  return Dataset::FromDenseMatrixAligned(dense.get(), data_type, static_cast<int32_t>(nrow), static_cast<int32_t>(ncol), true,
                                         *reinterpret_cast<const Dataset*>(reference), config);
}

}  // namespace

extern "C" {

const char* LGBM_GetLastError(void) { return g_last_error; }

int LGBM_RegisterLogCallback(void (*callback)(const char*)) {
  API_BEGIN();
  Log::callback() = callback;
  API_END();
}

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

int LGBM_DatasetCreateFromCSR(const void* indptr, int indptr_type, const int32_t* indices, const void* data, int data_type,
                              int64_t nindptr, int64_t nelem, int64_t num_col, const char* parameters,
                              const DatasetHandle reference, DatasetHandle* out) {
  API_BEGIN();
  (void)nelem;
  if (!out) Log::Fatal("LGBM_DatasetCreateFromCSR: NULL argument");
  *out = DatasetFromSparse(true, indptr, indptr_type, indices, data, data_type, nindptr, nindptr - 1, num_col, parameters, reference);
  API_END();
}

int LGBM_DatasetCreateFromCSC(const void* col_ptr, int col_ptr_type, const int32_t* indices, const void* data, int data_type,
                              int64_t ncol_ptr, int64_t nelem, int64_t num_row, const char* parameters,
                              const DatasetHandle reference, DatasetHandle* out) {
  API_BEGIN();
  (void)nelem;
  if (!out) Log::Fatal("LGBM_DatasetCreateFromCSC: NULL argument");
  *out = DatasetFromSparse(false, col_ptr, col_ptr_type, indices, data, data_type, ncol_ptr, num_row, ncol_ptr - 1, parameters, reference);
  API_END();
}

int LGBM_DatasetGetSubset(const DatasetHandle handle, const int32_t* used_row_indices, int32_t num_used_row_indices,
                          const char* parameters, DatasetHandle* out) {
  API_BEGIN();
  if (!handle || !used_row_indices || !out) Log::Fatal("LGBM_DatasetGetSubset: NULL argument");
  Config config;
  config.Set(Config::Str2Map(parameters));
  if (config.num_threads > 0) omp_set_num_threads(config.num_threads);
  *out = Dataset::Subset(*reinterpret_cast<const Dataset*>(handle), used_row_indices, num_used_row_indices);
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
  Trainer(handle);
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
  Booster* b = reinterpret_cast<Booster*>(handle);
  *out_iteration = b->loaded ? static_cast<int>(b->loaded->trees.size()) / b->loaded->num_tree_per_iteration : b->gbdt->current_iteration();
  API_END();
}

int LGBM_BoosterNumModelPerIteration(BoosterHandle handle, int* out_tree_per_iteration) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  *out_tree_per_iteration = b->loaded ? b->loaded->num_tree_per_iteration : b->gbdt->num_tree_per_iteration();
  API_END();
}

int LGBM_BoosterNumberOfTotalModel(BoosterHandle handle, int* out_models) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  *out_models = b->loaded ? static_cast<int>(b->loaded->trees.size()) : b->gbdt->num_models();
  API_END();
}

int LGBM_BoosterGetNumClasses(BoosterHandle handle, int* out_len) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  *out_len = b->loaded ? b->loaded->num_tree_per_iteration : b->gbdt->num_tree_per_iteration();
  API_END();
}

int LGBM_BoosterGetNumPredict(BoosterHandle handle, int data_idx, int64_t* out_len) {
  API_BEGIN();
  GBDT* g = reinterpret_cast<Booster*>(handle)->gbdt.get();
  if (data_idx < 0 || data_idx > g->num_valid()) Log::Fatal("Check failed: data_idx >= 0 && data_idx <= number of validation sets");
  *out_len = static_cast<int64_t>(data_idx == 0 ? g->num_data() : g->valid_num_data(data_idx - 1)) * g->num_tree_per_iteration();
  API_END();
}

int LGBM_BoosterGetPredict(BoosterHandle handle, int data_idx, int64_t* out_len, double* out_result) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mu);
  if (data_idx < 0 || data_idx > b->gbdt->num_valid()) Log::Fatal("Check failed: data_idx >= 0 && data_idx <= number of validation sets");
  if (data_idx == 0) {
    b->gbdt->GetTrainScore(out_result, true);
    *out_len = static_cast<int64_t>(b->gbdt->num_data()) * b->gbdt->num_tree_per_iteration();
  } else {
    b->gbdt->GetValidScore(data_idx - 1, out_result, true);
    *out_len = static_cast<int64_t>(b->gbdt->valid_num_data(data_idx - 1)) * b->gbdt->num_tree_per_iteration();
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
  Booster* b = reinterpret_cast<Booster*>(handle);
  *out_len = b->loaded ? 0 : static_cast<int>(b->gbdt->EvalNames().size());
  API_END();
}

int LGBM_BoosterGetEvalNames(BoosterHandle handle, const int len, int* out_len, const size_t buffer_len,
                             size_t* out_buffer_len, char** out_strs) {
  API_BEGIN();
  Booster* bb = reinterpret_cast<Booster*>(handle);
  const std::vector<std::string> names = bb->loaded ? std::vector<std::string>() : bb->gbdt->EvalNames();
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
  if (predict_type != C_API_PREDICT_NORMAL && predict_type != C_API_PREDICT_RAW_SCORE && predict_type != C_API_PREDICT_LEAF_INDEX)
    Log::Fatal("predict_type %d is outside this library's path (normal, raw score, leaf index)", predict_type);
  Booster* b = reinterpret_cast<Booster*>(handle);
  if (predict_type == C_API_PREDICT_LEAF_INDEX) {
    // GBDT::PredictLeafIndex (gbdt_prediction.cpp:77-86): the leaf of every tree, [row][tree]
    const int n_models_all = b->loaded ? static_cast<int>(b->loaded->trees.size()) : b->gbdt->num_models();
    const int K = b->loaded ? b->loaded->num_tree_per_iteration : b->gbdt->num_tree_per_iteration();
    const int want_cols = b->loaded ? b->loaded->max_feature_idx + 1 : b->gbdt->max_feature_idx() + 1;
    if (ncol != want_cols) Log::Fatal("The number of features in data (%d) is not the same as it was in training data (%d).", ncol, want_cols);
    int n_models = n_models_all;
    if (num_iteration > 0) n_models = std::min(n_models, num_iteration * K);
#pragma omp parallel
    {
      std::vector<double> row(ncol);
#pragma omp for schedule(static)
      for (int32_t r = 0; r < nrow; ++r) {
        for (int c = 0; c < ncol; ++c) {
          const size_t i = is_row_major ? static_cast<size_t>(r) * ncol + c : static_cast<size_t>(c) * nrow + r;
          row[c] = data_type == 0 ? static_cast<double>(static_cast<const float*>(data)[i]) : static_cast<const double*>(data)[i];
        }
        for (int t = 0; t < n_models; ++t) {
          const Tree& tree = b->loaded ? *b->loaded->trees[t] : b->gbdt->model(t);
          out_result[static_cast<size_t>(r) * n_models + t] = tree.PredictLeaf(row.data());
        }
      }
    }
    *out_len = static_cast<int64_t>(nrow) * n_models;
    return 0;
  }
  if (b->loaded) {
    const LoadedModel& m = *b->loaded;
    if (ncol != m.max_feature_idx + 1)
      Log::Fatal("The number of features in data (%d) is not the same as it was in training data (%d).", ncol, m.max_feature_idx + 1);
    const int K = m.num_tree_per_iteration;
    int n_iter = static_cast<int>(m.trees.size()) / K;
    if (num_iteration > 0) n_iter = std::min(n_iter, num_iteration);
    const bool raw = predict_type == C_API_PREDICT_RAW_SCORE;
#pragma omp parallel
    {
      std::vector<double> row(ncol), sc(K);
#pragma omp for schedule(static)
      for (int32_t r = 0; r < nrow; ++r) {
        for (int c = 0; c < ncol; ++c) {
          const size_t i = is_row_major ? static_cast<size_t>(r) * ncol + c : static_cast<size_t>(c) * nrow + r;
          row[c] = data_type == 0 ? static_cast<double>(static_cast<const float*>(data)[i]) : static_cast<const double*>(data)[i];
        }
        for (int k = 0; k < K; ++k) sc[k] = 0.0;
        for (int t = 0; t < n_iter; ++t) for (int k = 0; k < K; ++k) sc[k] += m.trees[static_cast<size_t>(t) * K + k]->Predict(row.data());
        if (!raw && m.average_output && n_iter > 0) for (int k = 0; k < K; ++k) sc[k] /= n_iter;
        if (!raw && m.sigmoid > 0) sc[0] = 1.0f / (1.0f + std::exp(-m.sigmoid * sc[0]));
        if (!raw && m.exp_out) sc[0] = std::exp(sc[0]);
        if (!raw && m.square_out) sc[0] = ((sc[0] > 0.0) - (sc[0] < 0.0)) * sc[0] * sc[0];
        if (!raw && m.softplus_out) sc[0] = std::log(1.0f + std::exp(sc[0]));
        if (!raw && m.ova_sigmoid > 0) for (int k = 0; k < K; ++k) sc[k] = 1.0f / (1.0f + std::exp(-m.ova_sigmoid * sc[k]));
        if (!raw && m.softmax) {   // Common::Softmax
          double wmax = sc[0];
          for (int k = 1; k < K; ++k) wmax = std::max(sc[k], wmax);
          double wsum = 0.0f;
          for (int k = 0; k < K; ++k) { sc[k] = std::exp(sc[k] - wmax); wsum += sc[k]; }
          for (int k = 0; k < K; ++k) sc[k] /= static_cast<double>(wsum);
        }
        for (int k = 0; k < K; ++k) out_result[static_cast<size_t>(r) * K + k] = sc[k];
      }
    }
    *out_len = static_cast<int64_t>(nrow) * K;
    return 0;
  }
  b->gbdt->PredictMat(data, data_type, nrow, ncol, is_row_major != 0, predict_type == C_API_PREDICT_RAW_SCORE,
                      num_iteration, out_result);
  *out_len = static_cast<int64_t>(nrow) * b->gbdt->num_tree_per_iteration();
  API_END();
}

int LGBM_BoosterSaveModelToString(BoosterHandle handle, int start_iteration, int num_iteration, int feature_importance_type,
                                  int64_t buffer_len, int64_t* out_len, char* out_str) {
  API_BEGIN();
  (void)feature_importance_type;
  Booster* b = reinterpret_cast<Booster*>(handle);
  if (b->loaded && !(start_iteration <= 0 && num_iteration <= 0)) Log::Fatal("a loaded model is written back whole (start_iteration / num_iteration not supported)");
  const std::string model = b->loaded ? b->loaded->text : b->gbdt->SaveModelToString(start_iteration, num_iteration);
  *out_len = static_cast<int64_t>(model.size()) + 1;
  if (*out_len <= buffer_len) std::memcpy(out_str, model.c_str(), *out_len);
  API_END();
}

int LGBM_BoosterSaveModel(BoosterHandle handle, int start_iteration, int num_iteration, int feature_importance_type,
                          const char* filename) {
  API_BEGIN();
  (void)feature_importance_type;
  Booster* b = reinterpret_cast<Booster*>(handle);
  const std::string model = b->loaded ? b->loaded->text : b->gbdt->SaveModelToString(start_iteration, num_iteration);
  std::ofstream f(filename, std::ios::binary);
  if (!f) Log::Fatal("Model file %s is not available for writes", filename);
  f.write(model.data(), static_cast<std::streamsize>(model.size()));
  API_END();
}

// ---- entry points the reference's python package reaches around lgb.train (python-package/lightgbm/basic.py) ------

static void CopyNames(const std::vector<std::string>& names, const int len, int* out_len, const size_t buffer_len,
                      size_t* out_buffer_len, char** out_strs) {
  *out_buffer_len = 0;
  int idx = 0;
  for (const std::string& name : names) {
    if (idx < len && buffer_len > 0) {
      std::memcpy(out_strs[idx], name.c_str(), std::min(name.size() + 1, buffer_len));
      out_strs[idx][buffer_len - 1] = '\0';
    }
    *out_buffer_len = std::max(name.size() + 1, *out_buffer_len);
    ++idx;
  }
  *out_len = idx;
}

int LGBM_DatasetSetFeatureNames(DatasetHandle handle, const char** feature_names, int num_feature_names) {
  API_BEGIN();
  if (!handle || !feature_names) Log::Fatal("LGBM_DatasetSetFeatureNames: NULL argument");
  std::vector<std::string> names;
  for (int i = 0; i < num_feature_names; ++i) names.emplace_back(feature_names[i]);
  reinterpret_cast<Dataset*>(handle)->set_feature_names(names);
  API_END();
}

int LGBM_DatasetGetFeatureNames(DatasetHandle handle, const int len, int* num_feature_names, const size_t buffer_len,
                                size_t* out_buffer_len, char** feature_names) {
  API_BEGIN();
  CopyNames(reinterpret_cast<Dataset*>(handle)->feature_names(), len, num_feature_names, buffer_len, out_buffer_len, feature_names);
  API_END();
}

int LGBM_DatasetGetField(DatasetHandle handle, const char* field_name, int* out_len, const void** out_ptr, int* out_type) {
  API_BEGIN();
  if (!handle || !field_name || !out_len || !out_ptr || !out_type) Log::Fatal("LGBM_DatasetGetField: NULL argument");
  const Metadata& md = reinterpret_cast<Dataset*>(handle)->metadata();
  const std::string k(field_name);
  // Dataset::GetFloatField / GetDoubleField / GetIntField (src/io/dataset.cpp:880-930); "group" returns the boundaries
  if (k == "label" || k == "target") { *out_ptr = md.label.data(); *out_len = static_cast<int>(md.label.size()); *out_type = C_API_DTYPE_FLOAT32; }
  else if (k == "weight" || k == "weights") { *out_ptr = md.weights.data(); *out_len = static_cast<int>(md.weights.size()); *out_type = C_API_DTYPE_FLOAT32; }
  else if (k == "init_score") { *out_ptr = md.init_score.data(); *out_len = static_cast<int>(md.init_score.size()); *out_type = C_API_DTYPE_FLOAT64; }
  else if (k == "query" || k == "group") { *out_ptr = md.query_boundaries.data(); *out_len = static_cast<int>(md.query_boundaries.size()); *out_type = C_API_DTYPE_INT32; }
  else Log::Fatal("Field not found");
  if (*out_len == 0) *out_ptr = nullptr;
  API_END();
}

int LGBM_DatasetUpdateParamChecking(const char* old_parameters, const char* new_parameters) {
  API_BEGIN();
  // Dataset-construction parameters cannot change once the Dataset exists (src/c_api.cpp:1441-1451 ->
  // Config::... "Cannot change X after constructed Dataset handle")
  Config a, b;
  auto ma = Config::Str2Map(old_parameters);
  a.Set(ma);
  auto mb = ma;
  for (auto& kv : Config::Str2Map(new_parameters)) mb[kv.first] = kv.second;
  b.Set(mb);
  if (a.max_bin != b.max_bin) Log::Fatal("Cannot change max_bin after constructed Dataset handle.");
  if (a.bin_construct_sample_cnt != b.bin_construct_sample_cnt) Log::Fatal("Cannot change bin_construct_sample_cnt after constructed Dataset handle.");
  if (a.min_data_in_bin != b.min_data_in_bin) Log::Fatal("Cannot change min_data_in_bin after constructed Dataset handle.");
  if (a.use_missing != b.use_missing) Log::Fatal("Cannot change use_missing after constructed Dataset handle.");
  if (a.zero_as_missing != b.zero_as_missing) Log::Fatal("Cannot change zero_as_missing after constructed Dataset handle.");
  if (a.enable_bundle != b.enable_bundle) Log::Fatal("Cannot change enable_bundle after constructed Dataset handle.");
  if (a.data_random_seed != b.data_random_seed) Log::Fatal("Cannot change data_random_seed after constructed Dataset handle.");
  if (a.max_bin_by_feature != b.max_bin_by_feature) Log::Fatal("Cannot change max_bin_by_feature after constructed Dataset handle.");
  if (a.categorical_feature != b.categorical_feature) Log::Fatal("Cannot change categorical_feature after constructed Dataset handle.");
  if (a.feature_pre_filter != b.feature_pre_filter) Log::Fatal("Cannot change feature_pre_filter after constructed Dataset handle.");
  if (a.feature_pre_filter && a.min_data_in_leaf < b.min_data_in_leaf)
    Log::Fatal("Reducing `min_data_in_leaf` with `feature_pre_filter=true` may cause unexpected behaviour for features that were pre-filtered by the larger `min_data_in_leaf`.");
  API_END();
}

int LGBM_BoosterLoadModelFromString(const char* model_str, int* out_num_iterations, BoosterHandle* out) {
  API_BEGIN();
  if (!model_str || !out_num_iterations || !out) Log::Fatal("LGBM_BoosterLoadModelFromString: NULL argument");
  std::unique_ptr<Booster> b(new Booster());
  b->loaded.reset(ParseModel(model_str));
  *out_num_iterations = static_cast<int>(b->loaded->trees.size());
  *out = b.release();
  API_END();
}

int LGBM_BoosterCreateFromModelfile(const char* filename, int* out_num_iterations, BoosterHandle* out) {
  API_BEGIN();
  if (!filename || !out_num_iterations || !out) Log::Fatal("LGBM_BoosterCreateFromModelfile: NULL argument");
  std::ifstream f(filename, std::ios::binary);
  if (!f) Log::Fatal("Model file %s is not available for reads", filename);
  std::stringstream ss;
  ss << f.rdbuf();
  std::unique_ptr<Booster> b(new Booster());
  b->loaded.reset(ParseModel(ss.str()));
  *out_num_iterations = static_cast<int>(b->loaded->trees.size());
  *out = b.release();
  API_END();
}

int LGBM_BoosterResetParameter(BoosterHandle handle, const char* parameters) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  GBDT* g = Trainer(handle);
  std::lock_guard<std::mutex> lock(b->mu);
  auto merged = g->config().raw;
  for (auto& kv : Config::Str2Map(parameters)) merged[kv.first] = kv.second;
  Config c;
  c.Set(merged);
  if (!Net().active || Net().num_machines <= 1) c.tree_learner = "serial";
  g->ResetParameter(c);
  API_END();
}

int LGBM_BoosterRollbackOneIter(BoosterHandle handle) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  GBDT* g = Trainer(handle);
  std::lock_guard<std::mutex> lock(b->mu);
  g->RollbackOneIter();
  API_END();
}

int LGBM_BoosterGetNumFeature(BoosterHandle handle, int* out_len) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  *out_len = b->loaded ? b->loaded->max_feature_idx + 1 : b->gbdt->train_data()->num_total_features();
  API_END();
}

int LGBM_BoosterGetFeatureNames(BoosterHandle handle, const int len, int* out_len, const size_t buffer_len,
                                size_t* out_buffer_len, char** out_strs) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  CopyNames(b->loaded ? b->loaded->feature_names : b->gbdt->train_data()->feature_names(), len, out_len, buffer_len,
            out_buffer_len, out_strs);
  API_END();
}

int LGBM_BoosterFeatureImportance(BoosterHandle handle, int num_iteration, int importance_type, double* out_results) {
  API_BEGIN();
  if (importance_type != 0 && importance_type != 1) Log::Fatal("Unknown importance type: only support split=0 and gain=1");
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::vector<double> imp;
  if (b->loaded) {
    imp.assign(b->loaded->max_feature_idx + 1, 0.0);
    int n = static_cast<int>(b->loaded->trees.size());
    if (num_iteration > 0) n = std::min(n, num_iteration);
    for (int m = 0; m < n; ++m) {
      const Tree& t = *b->loaded->trees[m];
      for (int i = 0; i < t.num_leaves() - 1; ++i)
        if (t.split_gain()[i] > 0) imp[t.split_feature()[i]] += importance_type == 0 ? 1.0 : t.split_gain()[i];
    }
  } else {
    imp = b->gbdt->FeatureImportance(num_iteration, importance_type);
  }
  for (size_t i = 0; i < imp.size(); ++i) out_results[i] = imp[i];
  API_END();
}

// LGBM_BoosterPredictForCSR (src/c_api.cpp:1893-1943): rows densified in chunks and sent through the dense prediction
int LGBM_BoosterPredictForCSR(BoosterHandle handle, const void* indptr, int indptr_type, const int32_t* indices, const void* data,
                              int data_type, int64_t nindptr, int64_t nelem, int64_t num_col, int predict_type, int start_iteration,
                              int num_iteration, const char* parameter, int64_t* out_len, double* out_result) {
  API_BEGIN();
  (void)nelem;
  if (!indptr || !indices || !data || !out_len || !out_result || num_col <= 0 || num_col >= INT32_MAX || nindptr < 1)
    Log::Fatal("LGBM_BoosterPredictForCSR: bad argument");
  const int64_t nrow = nindptr - 1;
  const int64_t chunk = std::max<int64_t>(1, std::min<int64_t>(nrow, (static_cast<int64_t>(256) << 20) / (8 * num_col)));
  std::vector<double> dense(static_cast<size_t>(chunk) * num_col);
  int64_t total = 0;
  for (int64_t r0 = 0; r0 < nrow; r0 += chunk) {
    const int64_t rows = std::min(chunk, nrow - r0);
    Densify<double>(true, indptr, indptr_type, indices, data, data_type, nindptr, nrow, num_col, r0, rows, dense.data());
    int64_t len = 0;
    if (LGBM_BoosterPredictForMat(handle, dense.data(), C_API_DTYPE_FLOAT64, static_cast<int32_t>(rows), static_cast<int32_t>(num_col), 1,
                                  predict_type, start_iteration, num_iteration, parameter, &len, out_result + total) != 0)
      throw std::runtime_error(LGBM_GetLastError());
    total += len;
  }
  *out_len = total;
  API_END();
}

int LGBM_BoosterCalcNumPredict(BoosterHandle handle, int num_row, int predict_type, int start_iteration, int num_iteration,
                               int64_t* out_len) {
  API_BEGIN();
  (void)start_iteration;
  if (predict_type != C_API_PREDICT_NORMAL && predict_type != C_API_PREDICT_RAW_SCORE && predict_type != C_API_PREDICT_LEAF_INDEX)
    Log::Fatal("predict_type %d is outside this library's path (normal, raw score, leaf index)", predict_type);
  Booster* b = reinterpret_cast<Booster*>(handle);
  if (predict_type == C_API_PREDICT_LEAF_INDEX) {
    const int K = b->loaded ? b->loaded->num_tree_per_iteration : b->gbdt->num_tree_per_iteration();
    int n_models = b->loaded ? static_cast<int>(b->loaded->trees.size()) : b->gbdt->num_models();
    if (num_iteration > 0) n_models = std::min(n_models, num_iteration * K);
    *out_len = static_cast<int64_t>(num_row) * n_models;
    return 0;
  }
  *out_len = static_cast<int64_t>(num_row) * (b->loaded ? b->loaded->num_tree_per_iteration : b->gbdt->num_tree_per_iteration());   // one value per row and class
  API_END();
}

// The reference's socket / MPI linkers are not part of this library: inside one NVLink box the ranks meet over NCCL.
int LGBM_NetworkInit(const char* machines, int local_listen_port, int listen_time_out, int num_machines) {
  API_BEGIN();
  (void)machines; (void)local_listen_port; (void)listen_time_out;
  if (num_machines > 1)
    Log::Fatal("LGBM_NetworkInit (TCP machine list) is not available in this library: data-parallel training runs over "
               "NCCL -- rank 0 calls LGBM_NetworkGetNCCLUniqueId, every rank LGBM_NetworkInitNCCL(id, rank, num_machines)");
  API_END();
}

int LGBM_NetworkInitWithFunctions(int num_machines, int rank, void* reduce_scatter_ext_fun, void* allgather_ext_fun) {
  API_BEGIN();
  (void)rank; (void)reduce_scatter_ext_fun; (void)allgather_ext_fun;
  if (num_machines > 1)
    Log::Fatal("LGBM_NetworkInitWithFunctions (external collectives on host buffers) is not available in this library: "
               "the histograms never leave the GPUs; use LGBM_NetworkInitNCCL");
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
  const int FS = ds->num_storage_columns();
  std::vector<int> scol(F), wide(F);
  for (int f = 0, s = 0; f < F; ++f) { wide[f] = ds->stored_values(f) > 256; scol[f] = s; s += wide[f] ? 2 : 1; }
  const std::vector<uint8_t>& bytes = ds->bins();
  for (int f = 0; f < F; ++f) real_index[f] = ds->real_feature_index(f);
  for (int r = 0; r < N; ++r) {
    for (int f = 0; f < F; ++f) {
      const BinMapper& m = ds->mapper(f);
      const uint8_t* p = bytes.data() + static_cast<size_t>(r) * FS + scol[f];
      const int stored = wide[f] ? (p[0] | (p[1] << 8)) : p[0];
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
  for (int i = 0; i < m.num_bin() && i < cap; ++i) upper_bounds[i] = m.BinToValue(i);   // categorical: the bin's category
  API_END();
}

}  // extern "C"
