// The reference's tiny LCG (include/LightGBM/utils/random.h:20-112).  Bin-construction sampling, bagging and
// column sampling all draw from it, so matching its sequence is part of "same inputs -> same trees".
#pragma once
#include <cmath>
#include <set>
#include <vector>

namespace lgb {

class Random {
 public:
  explicit Random(int seed = 123456789) : x_(static_cast<unsigned>(seed)) {}
  int NextShort(int lo, int hi) { return Rand16() % (hi - lo) + lo; }
  int NextInt(int lo, int hi) { return Rand32() % (hi - lo) + lo; }
  float NextFloat() { return static_cast<float>(Rand16()) / 32768.0f; }

  // K ordered samples of {0..N-1} (random.h:67-99)
  std::vector<int> Sample(int N, int K) {
    std::vector<int> ret;
    ret.reserve(K > 0 ? K : 0);
    if (K > N || K <= 0) return ret;
    if (K == N) {
      for (int i = 0; i < N; ++i) ret.push_back(i);
    } else if (K > 1 && K > (N / std::log2(K))) {
      for (int i = 0; i < N; ++i) {
        const double prob = (K - ret.size()) / static_cast<double>(N - i);
        if (NextFloat() < prob) ret.push_back(i);
      }
    } else {
      std::set<int> picked;
      for (int r = N - K; r < N; ++r) {
        const int v = NextInt(0, r);
        if (!picked.insert(v).second) picked.insert(r);
      }
      ret.assign(picked.begin(), picked.end());
    }
    return ret;
  }

 private:
  int Rand16() { x_ = 214013u * x_ + 2531011u; return static_cast<int>((x_ >> 16) & 0x7FFF); }
  int Rand32() { x_ = 214013u * x_ + 2531011u; return static_cast<int>(x_ & 0x7FFFFFFF); }
  unsigned x_;
};

}  // namespace lgb
