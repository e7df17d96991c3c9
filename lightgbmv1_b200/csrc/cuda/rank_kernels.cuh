// LambdarankNDCG::GetGradients on device (src/objective/rank_objective.hpp:48-65,139-229).
#pragma once
namespace {
int rank_prepare(LgbCuContext* c) { (void)c; return fail("lambdarank on device: not built yet"); }
int rank_launch(LgbCuContext* c) { (void)c; return 0; }
}  // namespace
