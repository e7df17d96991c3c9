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
      // Floyd's sampling as in random.h:85-95; the reference keeps the picks in a std::set and returns them in
      // ascending order -- a bitmap over 0..N-1 gives the same set and the same order without the tree
      std::vector<unsigned long long> picked((static_cast<size_t>(N) + 63) / 64, 0ull);
      auto test_and_set = [&](int v) {
        unsigned long long& w = picked[static_cast<size_t>(v) >> 6];
        const unsigned long long bit = 1ull << (v & 63);
        const bool was = (w & bit) != 0;
        w |= bit;
        return was;
      };
      for (int r = N - K; r < N; ++r) {
        const int v = NextInt(0, r);
        if (test_and_set(v)) test_and_set(r);
      }
      for (size_t w = 0; w < picked.size(); ++w) {
        unsigned long long bits = picked[w];
        while (bits) {
          const int b = __builtin_ctzll(bits);
          ret.push_back(static_cast<int>(w * 64 + b));
          bits &= bits - 1;
        }
      }
    }
    return ret;
  }

 private:
  int Rand16() { x_ = 214013u * x_ + 2531011u; return static_cast<int>((x_ >> 16) & 0x7FFF); }
  int Rand32() { x_ = 214013u * x_ + 2531011u; return static_cast<int>(x_ & 0x7FFFFFFF); }
  unsigned x_;
};

}  // namespace lgb
