// Host half of the objectives north_star names: everything ObjectiveFunction::Init / BoostFromScore /
// ConvertOutput / ToString do in the reference (src/objective/binary_objective.hpp:56-158,
// regression_objective.hpp:93-199, rank_objective.hpp:100-140).  GetGradients itself runs on device
// (lgbcu_get_gradients); this class only derives its parameters.
#pragma once
#include <functional>
#include <memory>
#include <string>

#include "config.h"
#include "dataset.h"
#include "lgbcu.h"

namespace lgb {

class Objective {
 public:
  static Objective* Create(const Config& config);  // ObjectiveFunction::CreateObjectiveFunction
  virtual ~Objective() {}
  // `sum_over_ranks` sums n doubles over the machines of a data-parallel run in place (Network::GlobalSyncUpBySum);
  // it is a no-op on one machine
  typedef std::function<void(double*, int)> SumSync;
  virtual void Init(const Metadata& meta, data_size_t num_data, const SumSync& sum_over_ranks) = 0;
  virtual void Configure(LgbCuContext* ctx) const = 0;  // push kind + parameters to the device learner
  virtual double BoostFromScore(int class_id = 0) const = 0;
  virtual bool NeedTrain(int class_id = 0) const { return true; }   // ObjectiveFunction::ClassNeedTrain
  // ObjectiveFunction::IsRenewTreeOutput: leaf values are recomputed as a percentile of the residuals after a tree is
  // grown (lgbcu_renew_leaf_outputs); RenewAlpha / RenewAlphaIsFloat say which one and in which arithmetic
  virtual bool IsRenewTreeOutput() const { return false; }
  virtual double RenewAlpha() const { return 0.5; }
  virtual bool RenewAlphaIsFloat() const { return false; }
  virtual int NumModelPerIteration() const { return 1; }           // trees (= score planes) per boosting iteration
  // ConvertOutput works on one row's NumModelPerIteration() raw scores
  virtual data_size_t NumPositiveData() const { return 0; }  // rows of this machine with label > 0 (balanced bagging)
  virtual bool IsConstantHessian() const { return false; }
  virtual void ConvertOutput(const double* in, double* out) const { out[0] = in[0]; }
  virtual std::string ToString() const = 0;
  virtual const char* Name() const = 0;
};

}  // namespace lgb
