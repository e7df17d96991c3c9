// Value -> bin mapping (A1): BinMapper::FindBin / ValueToBin of the reference (src/io/bin.cpp:78-155,256-312,
// 325-520; include/LightGBM/bin.h:457-495), numerical and categorical features.  Must be bit-exact: the bin boundaries
// decide every stored byte of the device matrix and the real-valued thresholds written into the model.
#pragma once
#include <string>
#include <unordered_map>
#include <vector>

#include "common.h"

namespace lgb {

class BinMapper {
 public:
  // values: the sampled NON-ZERO values of one column (zeros are implied: total_sample_cnt - values.size()).
  void FindBin(std::vector<double> values, size_t total_sample_cnt, int max_bin, int min_data_in_bin,
               int min_split_data, bool pre_filter, bool use_missing, bool zero_as_missing, bool categorical = false);
  uint32_t ValueToBin(double value) const;
  // BinMapper::BinToValue (bin.h:120-127): the bin's upper bound, or the category of a categorical bin
  double BinToValue(uint32_t bin) const { return categorical_ ? static_cast<double>(bin_2_cat_[bin]) : upper_[bin]; }
  bool is_categorical() const { return categorical_; }
  const std::vector<int>& bin_2_categorical() const { return bin_2_cat_; }

  int num_bin() const { return num_bin_; }
  int missing_type() const { return missing_type_; }
  bool is_trivial() const { return trivial_; }
  uint32_t default_bin() const { return default_bin_; }
  uint32_t most_freq_bin() const { return most_freq_bin_; }
  double sparse_rate() const { return sparse_rate_; }
  double min_val() const { return min_val_; }
  double max_val() const { return max_val_; }
  const std::vector<double>& upper_bounds() const { return upper_; }
  std::string info_string() const;  // BinMapper::bin_info_string (bin.h:193-207)

 private:
  int num_bin_ = 1;
  int missing_type_ = kMissingNone;
  bool trivial_ = true;
  double sparse_rate_ = 1.0;
  std::vector<double> upper_;
  double min_val_ = 0, max_val_ = 0;
  uint32_t default_bin_ = 0, most_freq_bin_ = 0;
  // BinType::CategoricalBin: bin 0 = NaN / negative / rare categories (category -1), bins 1.. = categories by frequency
  bool categorical_ = false;
  std::vector<int> bin_2_cat_;
  std::unordered_map<int, uint32_t> cat_2_bin_;
};

}  // namespace lgb
