// Binned data store (A2): what Dataset / FeatureGroup / DenseBin / Metadata are to the reference's training path
// (include/LightGBM/dataset.h, feature_group.h; src/io/dataset.cpp:315-432; src/c_api.cpp:1018-1100).
// Host side keeps the bin mappers, the feature order and a row-major u8 matrix of stored values that the CUDA
// learner uploads once; one feature per group (dense data), stored value = the reference's single-feature
// group encoding.
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "bin_mapper.h"
#include "config.h"
#include "lgbcu.h"

namespace lgb {

struct Metadata {
  std::vector<label_t> label;
  std::vector<label_t> weights;
  std::vector<data_size_t> query_boundaries;  // num_queries + 1, empty if none
  std::vector<double> init_score;
};

class Dataset {
 public:
  // LGBM_DatasetCreateFromMat with reference == nullptr: sample, find bins, bin every row
  static Dataset* FromDenseMatrix(const void* data, int data_type, int32_t nrow, int32_t ncol, bool row_major,
                                  const Config& config);
  // LGBM_DatasetGetSubset: the listed rows (ascending) of `full` with its mappers, bins and metadata
  // (Dataset::CopyFeatureMapperFrom + CopySubrow, src/io/dataset.cpp:742-829; Metadata::Init(fullset, indices))
  static Dataset* Subset(const Dataset& full, const int32_t* used_row_indices, int32_t num_used);
  // ... with a reference dataset: reuse its mappers and feature order (Dataset::CreateValid)
  static Dataset* FromDenseMatrixAligned(const void* data, int data_type, int32_t nrow, int32_t ncol, bool row_major,
                                         const Dataset& reference, const Config& config);

  ~Dataset();
  Dataset(const Dataset&) = delete;
  Dataset& operator=(const Dataset&) = delete;

  data_size_t num_data() const { return num_data_; }
  int num_features() const { return static_cast<int>(inner_to_real_.size()); }
  // stored values of an inner feature (FeatureGroup::num_total_bin_ of its single-feature group) and the byte columns of
  // the binned matrix: one per feature, two for a feature with more than 256 stored values (max_bin > 255)
  int stored_values(int inner) const { return mappers_[inner]->num_bin() + (mappers_[inner]->most_freq_bin() == 0 ? 0 : 1); }
  int num_storage_columns() const { int n = 0; for (int f = 0; f < num_features(); ++f) n += stored_values(f) > 256 ? 2 : 1; return n; }
  int num_total_features() const { return num_total_features_; }
  int real_feature_index(int inner) const { return inner_to_real_[inner]; }
  int bundle_of(int inner) const { return bundle_of_.empty() ? 0 : bundle_of_[inner]; }
  int inner_feature_index(int real) const { return real_to_inner_[real]; }
  const BinMapper& mapper(int inner) const { return *mappers_[inner]; }
  // row-major [num_data][num_features] stored values on the HOST.  When the matrix was binned on the device it is
  // fetched on first use (validation sets, dumps); training adopts the device copy directly (device_bins()).
  const std::vector<uint8_t>& bins() const;
  LgbCuBinnedMatrix* device_bins() const { return dev_bins_; }
  uint64_t num_total_bin() const;
  Metadata& metadata() { return meta_; }
  const Metadata& metadata() const { return meta_; }
  const std::vector<std::string>& feature_names() const { return feature_names_; }
  void set_feature_names(const std::vector<std::string>& names);  // Dataset::set_feature_names (dataset.h:621-652)
  std::vector<std::string> feature_infos() const;
  // value of feature `real` in bin space for one raw value (used by prediction on binned data)
  uint32_t stored_value(int inner, double v) const;

  void SetFloatField(const char* name, const float* data, int n);
  void SetDoubleField(const char* name, const double* data, int n);
  void SetIntField(const char* name, const int32_t* data, int n);

 private:
  explicit Dataset(data_size_t n) : num_data_(n) {}
  void BinAllRows(const void* data, int data_type, int32_t nrow, int32_t ncol, bool row_major, int device);

  data_size_t num_data_;
  int num_total_features_ = 0;
  std::vector<std::shared_ptr<BinMapper>> mappers_;  // per inner feature
  std::vector<int> inner_to_real_, real_to_inner_;
  // exclusive-feature bundle of every inner feature: 0 = none, b > 0 = the reference would keep it in one column with
  // the other features of bundle b (dataset.cpp BundledFeatureOrder; the override rule is applied in BinAllRows)
  std::vector<int> bundle_of_;
  mutable std::vector<uint8_t> bins_;
  mutable bool bins_pinned_ = false;  // page-locked for the upload (lgbcu_pin_host_buffer)
  LgbCuBinnedMatrix* dev_bins_ = nullptr;  // the matrix as binned on the device (lgbcu_bin_matrix), if it was
  bool label_pinned_ = false;
  std::vector<std::string> feature_names_;
  Metadata meta_;
};

}  // namespace lgb
