// Tree model: the host bookkeeping the reference keeps in Tree (include/LightGBM/tree.h:459-501,
// src/io/tree.cpp:20-68,223-262).  The device grows the tree; the host replays the split records through
// Split() so that node numbering, leaf ids and the model text are the reference's.
#pragma once
#include <string>
#include <vector>

#include "common.h"

namespace lgb {

class Tree {
 public:
  explicit Tree(int max_leaves);
  // numerical split; returns the id of the new right leaf
  int Split(int leaf, int feature_inner, int feature_real, uint32_t threshold_bin, double threshold_double,
            double left_value, double right_value, int left_cnt, int right_cnt, double left_weight,
            double right_weight, float gain, int missing_type, bool default_left);
  // categorical split (Tree::SplitCategorical, src/io/tree.cpp:70-91): `bin_set` = 8 words, bit b = bin b goes left
  // (the device's form of cat_threshold_inner_); `cat_set` = the bitset over the real category values (cat_threshold_)
  int SplitCategorical(int leaf, int feature_inner, int feature_real, const uint32_t* bin_set, const std::vector<uint32_t>& cat_set,
                       double left_value, double right_value, int left_cnt, int right_cnt, double left_weight,
                       double right_weight, float gain, int missing_type);
  int num_cat() const { return num_cat_; }
  const std::vector<uint32_t>& cat_bin_sets() const { return cat_bin_sets_; }   // [num_cat][8]
  void Shrinkage(double rate);
  void AddBias(double val);
  void AsConstantTree(double val) { num_leaves_ = 1; shrinkage_ = 1.0f; leaf_value_[0] = val; }
  std::string ToString() const;
  // the inverse of ToString (Tree::Tree(const char* str, size_t*), src/io/tree.cpp:580-700): a tree of a saved model,
  // good for Predict / ToString (bin-space thresholds are not part of the text)
  static Tree* FromString(const std::string& text);
  double Predict(const double* features) const;  // raw feature values (Tree::Predict / NumericalDecision)
  int PredictLeaf(const double* features) const;  // Tree::PredictLeafIndex (tree.h:420-432)

  int num_leaves() const { return num_leaves_; }
  double leaf_value(int i) const { return leaf_value_[i]; }
  // Tree::SetLeafOutput (tree.h:92-94): MaybeRoundToZero(output)
  void SetLeafOutput(int leaf, double v) { leaf_value_[leaf] = (v >= -kZeroThreshold && v <= kZeroThreshold) ? 0 : v; }
  const std::vector<double>& leaf_values() const { return leaf_value_; }
  const std::vector<int>& split_feature_inner() const { return split_feature_inner_; }
  const std::vector<int>& split_feature() const { return split_feature_; }
  const std::vector<uint32_t>& threshold_in_bin() const { return threshold_in_bin_; }
  const std::vector<int8_t>& decision_type() const { return decision_type_; }
  const std::vector<int>& left_child() const { return left_child_; }
  const std::vector<int>& right_child() const { return right_child_; }
  const std::vector<float>& split_gain() const { return split_gain_; }
  void ScaleLeafValues(double rate) { Shrinkage(rate); }

  static const int8_t kCategoricalMask = 1;
  static const int8_t kDefaultLeftMask = 2;  // include/LightGBM/tree.h: kCategoricalMask = 1, kDefaultLeftMask = 2

 private:
  int max_leaves_, num_leaves_;
  std::vector<int> left_child_, right_child_, split_feature_inner_, split_feature_;
  std::vector<uint32_t> threshold_in_bin_;
  std::vector<double> threshold_;
  std::vector<int8_t> decision_type_;
  std::vector<float> split_gain_;
  std::vector<int> leaf_parent_;
  std::vector<double> leaf_value_, leaf_weight_;
  std::vector<int> leaf_count_;
  std::vector<double> internal_value_, internal_weight_;
  std::vector<int> internal_count_;
  std::vector<int> leaf_depth_;
  double shrinkage_;
  int num_cat_ = 0;
  std::vector<int> cat_boundaries_{0};
  std::vector<uint32_t> cat_threshold_;
  std::vector<uint32_t> cat_bin_sets_;
};

}  // namespace lgb
