// Value -> bin on the device (replaces the row push loop of LGBM_DatasetCreateFromMat,
// src/c_api.cpp:1082-1097 -> FeatureGroup::PushData -> BinMapper::ValueToBin, include/LightGBM/bin.h:457-495).
// The host keeps what is sequential and tiny (sampling, BinMapper::FindBin); the N x F map runs here: one thread per
// (row, inner feature) does the reference's binary search over the feature's upper bounds -- same comparisons, in
// double, so the result is bit-identical -- and stores the single-feature-group encoding (0 = most frequent bin,
// else bin + 1 - [most_freq_bin == 0]; feature_group.h:153-165).  The raw matrix crosses PCIe once, in chunks staged
// through two page-locked buffers (filled by the host threads while the previous chunk is in flight).
#pragma once

struct LgbCuBinnedMatrix {
  int device = 0; int32_t nrow = 0, nfeat = 0;
  uint8_t* d_bins = nullptr;     // [nrow][nfeat] stored values, row-major, no padding; nfeat = BYTE columns: one per
                                 // feature, two (low byte first) per feature with more than 256 stored values
};

struct BinFeatDev { int column, num_bin, mfb, missing_type, bound_off, scol, num_cat, pad; };   // num_cat >= 0: categorical   // scol = (first) byte column of the output

template <typename T>
__global__ void __launch_bounds__(256) k_bin_rows(const T* __restrict__ x, long long rows, int ncol, int row_major, long long ld,
                                                  int F, int FS, const BinFeatDev* __restrict__ feats, const double* __restrict__ bounds,
                                                  uint8_t* __restrict__ out) {
  const long long total = rows * F;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / F;
    const int f = (int)(i - r * F);
    const BinFeatDev d = feats[f];
    double value = (double)(row_major ? x[r * ncol + d.column] : x[(long long)d.column * ld + r]);
    int bin;
    if (d.num_cat >= 0) {
      // BinMapper::ValueToBin of a categorical feature (include/LightGBM/bin.h:457-460, 483-494): NaN, negative and unseen
      // categories -> bin 0.  The feature's `bounds` hold its categories in ascending order, then their bins.
      bin = 0;
      if (value == value) {
        // static_cast<int>(value): values outside the int range are undefined there; they are taken as unseen here
        const int iv = (value > -2147483649.0 && value < 2147483648.0) ? (int)value : -1;
        if (iv >= 0) {
          const double* cats = bounds + d.bound_off;
          int l = 0, rr = d.num_cat - 1;
          while (l < rr) { const int m = (l + rr) >> 1; if (cats[m] < (double)iv) l = m + 1; else rr = m; }
          if (d.num_cat > 0 && cats[l] == (double)iv) bin = (int)cats[d.num_cat + l];
        }
      }
    } else if (value != value && d.missing_type == LGBCU_MISSING_NAN) {
      bin = d.num_bin - 1;
    } else {
      if (value != value) value = 0.0;
      int l = 0, rr = d.num_bin - 1;
      if (d.missing_type == LGBCU_MISSING_NAN) rr -= 1;
      const double* ub = bounds + d.bound_off;
      while (l < rr) {
        const int m = (rr + l - 1) / 2;
        if (value <= ub[m]) rr = m; else l = m + 1;
      }
      bin = l;
    }
    int stored = 0;
    if (bin != d.mfb) stored = bin + (d.mfb == 0 ? 0 : 1);   // bin + 1 - [mfb == 0]
    uint8_t* o = out + r * FS + d.scol;
    o[0] = (uint8_t)(stored & 255);
    if (d.num_bin + (d.mfb == 0 ? 0 : 1) > 256) o[1] = (uint8_t)(stored >> 8);   // a wide feature: two byte columns
  }
}

// Exclusive-feature bundles (LgbCuBinFeature::bundle): the reference keeps the members of a bundle in ONE column, so in
// a row where several members are off their most-frequent bin the pushes overwrite one another and the member pushed
// last -- the highest raw column -- survives; to every other member the row reads as most-frequent (its value lies
// outside their ranges: DenseBin::SplitInner / DenseBinIterator::Get with min_bin / max_bin, dense_bin.hpp:173-307,
// 446-460; the histogram leaves it to FixHistogram).  Here every feature has its own column(s), so the same rows are
// set to 0 after the map.  One thread per (row, bundle); `mem` lists a bundle's members by DESCENDING raw column.
struct BundleMember { int scol, wide; };
__global__ void __launch_bounds__(256) k_bundle_override(uint8_t* __restrict__ out, long long rows, int FS, int nb,
                                                         const int* __restrict__ start, const BundleMember* __restrict__ mem) {
  const long long total = rows * nb;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / nb;
    const int b = (int)(i - r * nb);
    uint8_t* o = out + r * FS;
    bool seen = false;
    for (int m = start[b]; m < start[b + 1]; ++m) {
      const BundleMember M = mem[m];
      const int v = o[M.scol] | (M.wide ? o[M.scol + 1] << 8 : 0);
      if (v == 0) continue;
      if (seen) { o[M.scol] = 0; if (M.wide) o[M.scol + 1] = 0; }
      seen = true;
    }
  }
}

// Page-locked staging for lgbcu_bin_matrix, kept between calls: allocating 2 x 64 MB of pinned memory costs ~35 ms, a
// third of binning 10.5M x 28 -- and datasets are created more than once per process (validation sets, shards,
// repeated training).  One pair per process, handed to one caller at a time; a concurrent caller allocates its own.
struct PinCache { void* p[2] = {nullptr, nullptr}; size_t bytes = 0; bool busy = false; std::mutex mu; };
static PinCache g_pin;
static bool pin_acquire(size_t bytes, void** out2) {
  std::lock_guard<std::mutex> lock(g_pin.mu);
  if (g_pin.busy) return false;
  if (g_pin.bytes < bytes) {
    for (void*& q : g_pin.p) { if (q) lgb_host_free(q); q = nullptr; }
    g_pin.bytes = 0;
    if (lgb_host_malloc(&g_pin.p[0], bytes) != cudaSuccess || lgb_host_malloc(&g_pin.p[1], bytes) != cudaSuccess) {
      cudaGetLastError();
      for (void*& q : g_pin.p) { if (q) lgb_host_free(q); q = nullptr; }
      return false;
    }
    g_pin.bytes = bytes;
  }
  g_pin.busy = true; out2[0] = g_pin.p[0]; out2[1] = g_pin.p[1];
  return true;
}
static void pin_release() { std::lock_guard<std::mutex> lock(g_pin.mu); g_pin.busy = false; }

extern "C" {

int lgbcu_binned_free(LgbCuBinnedMatrix* m) {
  if (!m) return 0;
  ScopedDevice dev_(m->device);
  if (m->d_bins) lgb_dev_free(m->d_bins);
  delete m;
  return 0;
}

int lgbcu_bin_matrix(int device, const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                     int32_t num_features, const LgbCuBinFeature* feats, const double* upper_bounds,
                     int32_t num_bounds, LgbCuBinnedMatrix** out) {
  NEED(data && feats && upper_bounds && out, "lgbcu_bin_matrix: NULL argument");
  NEED(nrow > 0 && ncol > 0 && num_features > 0, "lgbcu_bin_matrix: empty matrix");
  NEED(data_type == 0 || data_type == 1, "lgbcu_bin_matrix: data_type must be 0 (float32) or 1 (float64)");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail("lgbcu_bin_matrix: no CUDA device available"); }
  NEED(device >= 0 && device < ndev, "lgbcu_bin_matrix: device %d out of range (have %d)", device, ndev);
  USE_DEVICE(device);
  struct Guard {   // everything but the result is released on every way out
    void* dev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr}; void* pin[2] = {nullptr, nullptr}; bool pin_cached = false;
    cudaStream_t s[2] = {nullptr, nullptr}; cudaEvent_t ev[2] = {nullptr, nullptr}; LgbCuBinnedMatrix* m = nullptr;
    ~Guard() {
      for (void* p : dev) if (p) lgb_dev_free(p);
      if (pin_cached) pin_release(); else for (void* p : pin) if (p) lgb_host_free(p);
      for (auto e : ev) if (e) cudaEventDestroy(e);
      for (auto st : s) if (st) cudaStreamDestroy(st);
      if (m) lgbcu_binned_free(m);
    }
  } G;
  const size_t esz = data_type == 0 ? 4 : 8;
  const bool timing = std::getenv("LGBM_B200_TIMING") != nullptr;
  auto t_prev = std::chrono::steady_clock::now();
  auto tick = [&](const char* what) {
    if (!timing) return;
    const auto now = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[b200 timing]   bin_matrix: %-24s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_prev).count());
    t_prev = now;
  };
  std::vector<BinFeatDev> hf(num_features);
  for (int f = 0; f < num_features; ++f) {
    const int ncat = feats[f].num_categories;   // < 0: numerical
    NEED(feats[f].column >= 0 && feats[f].column < ncol && feats[f].num_bin >= 1 && feats[f].bound_offset >= 0 &&
         feats[f].bound_offset + (ncat >= 0 ? 2 * ncat : feats[f].num_bin) <= num_bounds, "lgbcu_bin_matrix: bad descriptor of feature %d", f);
    hf[f] = BinFeatDev{feats[f].column, feats[f].num_bin, feats[f].most_freq_bin, feats[f].missing_type, feats[f].bound_offset, 0, ncat, 0};
  }
  int FS = 0;   // byte columns of the result
  for (int f = 0; f < num_features; ++f) { hf[f].scol = FS; FS += (hf[f].num_bin + (hf[f].mfb == 0 ? 0 : 1) > 256) ? 2 : 1; }
  // exclusive-feature bundles: members by descending raw column
  int nb = 0;
  for (int f = 0; f < num_features; ++f) { NEED(feats[f].bundle >= 0, "lgbcu_bin_matrix: bad bundle of feature %d", f); nb = std::max(nb, (int)feats[f].bundle); }
  int* d_bstart = nullptr; BundleMember* d_bmem = nullptr;
  if (nb > 0) {
    std::vector<std::vector<int>> members(nb);
    for (int f = 0; f < num_features; ++f) if (feats[f].bundle > 0) members[feats[f].bundle - 1].push_back(f);
    std::vector<int> bstart(1, 0); std::vector<BundleMember> bmem;
    for (auto& g : members) {
      std::sort(g.begin(), g.end(), [&](int a, int b) { return feats[a].column > feats[b].column; });
      for (int f : g) bmem.push_back(BundleMember{hf[f].scol, (hf[f].num_bin + (hf[f].mfb == 0 ? 0 : 1) > 256) ? 1 : 0});
      bstart.push_back((int)bmem.size());
    }
    CU(lgb_dev_malloc(&G.dev[4], sizeof(int) * bstart.size())); d_bstart = static_cast<int*>(G.dev[4]);
    CU(lgb_dev_malloc(&G.dev[5], sizeof(BundleMember) * std::max<size_t>(1, bmem.size()))); d_bmem = static_cast<BundleMember*>(G.dev[5]);
    CU(cudaMemcpy(d_bstart, bstart.data(), sizeof(int) * bstart.size(), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_bmem, bmem.data(), sizeof(BundleMember) * bmem.size(), cudaMemcpyHostToDevice));
  }
  BinFeatDev* d_feats = nullptr; double* d_bounds = nullptr;
  CU(lgb_dev_malloc(&G.dev[0], sizeof(BinFeatDev) * num_features)); d_feats = static_cast<BinFeatDev*>(G.dev[0]);
  CU(lgb_dev_malloc(&G.dev[1], sizeof(double) * num_bounds)); d_bounds = static_cast<double*>(G.dev[1]);
  CU(cudaMemcpy(d_feats, hf.data(), sizeof(BinFeatDev) * num_features, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d_bounds, upper_bounds, sizeof(double) * num_bounds, cudaMemcpyHostToDevice));
  G.m = new LgbCuBinnedMatrix();
  G.m->device = device; G.m->nrow = nrow; G.m->nfeat = FS;
  CU(lgb_dev_malloc(reinterpret_cast<void**>(&G.m->d_bins), (size_t)nrow * FS));
  for (int i = 0; i < 2; ++i) { CU(cudaStreamCreateWithFlags(&G.s[i], cudaStreamNonBlocking)); CU(cudaEventCreateWithFlags(&G.ev[i], cudaEventDisableTiming)); }
  tick("descriptors + result alloc");
  int sm_count = 0;
  CU(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device));
  const int max_grid = sm_count * 16;
  static const int hw = [] { const long n = sysconf(_SC_NPROCESSORS_ONLN); return n > 0 ? (int)n : 1; }();
  const int nthreads = std::max(omp_get_max_threads(), std::min(8, hw));
  if (is_row_major) {
    // chunks of whole rows: host threads copy a chunk into a page-locked buffer while the previous one is in flight
    const size_t row_bytes = (size_t)ncol * esz;
    const size_t chunk_rows = std::max<size_t>(1, std::min<size_t>((size_t)nrow, ((size_t)64 << 20) / row_bytes));
    G.pin_cached = pin_acquire(chunk_rows * row_bytes, G.pin);
    for (int i = 0; i < 2; ++i) { if (!G.pin_cached) CU(lgb_host_malloc(&G.pin[i], chunk_rows * row_bytes)); CU(lgb_dev_malloc(&G.dev[2 + i], chunk_rows * row_bytes)); }
    tick("staging buffers");
    int k = 0;
    for (size_t r0 = 0; r0 < (size_t)nrow; r0 += chunk_rows, ++k) {
      const int b = k & 1;
      const size_t rows = std::min(chunk_rows, (size_t)nrow - r0);
      CU(cudaEventSynchronize(G.ev[b]));   // the copy that last used this staging buffer has finished (no-op the first time)
      const char* src = static_cast<const char*>(data) + r0 * row_bytes;
      const size_t bytes = rows * row_bytes;
      // one contiguous piece per thread (several MB: large enough for memcpy's streaming-store path)
      const size_t piece = ((bytes + (size_t)nthreads - 1) / (size_t)nthreads + 4095) & ~(size_t)4095;
#pragma omp parallel for schedule(static) num_threads(nthreads)
      for (int p = 0; p < nthreads; ++p) {
        const size_t a = (size_t)p * piece, e = std::min(bytes, a + piece);
        if (a < e) std::memcpy(static_cast<char*>(G.pin[b]) + a, src + a, e - a);
      }
      CU(cudaMemcpyAsync(G.dev[2 + b], G.pin[b], bytes, cudaMemcpyHostToDevice, G.s[b]));
      CU(cudaEventRecord(G.ev[b], G.s[b]));
      const long long work = (long long)rows * num_features;
      const int grid = (int)std::min<long long>(max_grid, (work + 255) / 256);
      uint8_t* dst = G.m->d_bins + r0 * FS;
      if (data_type == 0) k_bin_rows<float><<<grid, 256, 0, G.s[b]>>>(static_cast<const float*>(G.dev[2 + b]), (long long)rows, ncol, 1, 0, num_features, FS, d_feats, d_bounds, dst);
      else k_bin_rows<double><<<grid, 256, 0, G.s[b]>>>(static_cast<const double*>(G.dev[2 + b]), (long long)rows, ncol, 1, 0, num_features, FS, d_feats, d_bounds, dst);
      g_launches.fetch_add(1, std::memory_order_relaxed);
      if (nb > 0) {
        const int bgrid = (int)std::min<long long>(max_grid, ((long long)rows * nb + 255) / 256);
        k_bundle_override<<<bgrid, 256, 0, G.s[b]>>>(dst, (long long)rows, FS, nb, d_bstart, d_bmem);
        g_launches.fetch_add(1, std::memory_order_relaxed);
      }
      CU(cudaGetLastError());
    }
  } else {
    // column-major input: upload the used columns only, whole, then one launch
    const size_t col_bytes = (size_t)nrow * esz;
    CU(lgb_dev_malloc(&G.dev[2], col_bytes * num_features));
    std::vector<BinFeatDev> hc = hf;
    for (int f = 0; f < num_features; ++f) {
      CU(cudaMemcpyAsync(static_cast<char*>(G.dev[2]) + f * col_bytes, static_cast<const char*>(data) + (size_t)hf[f].column * col_bytes,
                         col_bytes, cudaMemcpyHostToDevice, G.s[0]));
      hc[f].column = f;
    }
    CU(cudaMemcpyAsync(d_feats, hc.data(), sizeof(BinFeatDev) * num_features, cudaMemcpyHostToDevice, G.s[0]));
    const long long work = (long long)nrow * num_features;
    const int grid = (int)std::min<long long>(max_grid, (work + 255) / 256);
    if (data_type == 0) k_bin_rows<float><<<grid, 256, 0, G.s[0]>>>(static_cast<const float*>(G.dev[2]), (long long)nrow, num_features, 0, (long long)nrow, num_features, FS, d_feats, d_bounds, G.m->d_bins);
    else k_bin_rows<double><<<grid, 256, 0, G.s[0]>>>(static_cast<const double*>(G.dev[2]), (long long)nrow, num_features, 0, (long long)nrow, num_features, FS, d_feats, d_bounds, G.m->d_bins);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (nb > 0) {
      const int bgrid = (int)std::min<long long>(max_grid, ((long long)nrow * nb + 255) / 256);
      k_bundle_override<<<bgrid, 256, 0, G.s[0]>>>(G.m->d_bins, (long long)nrow, FS, nb, d_bstart, d_bmem);
      g_launches.fetch_add(1, std::memory_order_relaxed);
    }
    CU(cudaGetLastError());
  }
  tick("chunks enqueued");
  CU(cudaStreamSynchronize(G.s[0]));
  CU(cudaStreamSynchronize(G.s[1]));
  tick("drain");
  *out = G.m;
  G.m = nullptr;
  return 0;
}

int lgbcu_binned_copy_to_host(LgbCuBinnedMatrix* m, uint8_t* out) {
  NEED(m && out, "lgbcu_binned_copy_to_host: NULL argument");
  USE_DEVICE(m->device);
  CU(cudaMemcpy(out, m->d_bins, (size_t)m->nrow * m->nfeat, cudaMemcpyDeviceToHost));
  return 0;
}

int lgbcu_binned_device(const LgbCuBinnedMatrix* m, int* device) {
  NEED(m && device, "lgbcu_binned_device: NULL argument");
  *device = m->device;
  return 0;
}

// A2 from a matrix that is already on the device (binned by lgbcu_bin_matrix): no host round trip
int lgbcu_set_dataset_binned(LgbCuContext* c, const LgbCuBinnedMatrix* m, const LgbCuFeatureMeta* meta) {
  NEED(c && m && meta, "lgbcu_set_dataset_binned: NULL argument");
  NEED(m->device == c->device, "lgbcu_set_dataset_binned: the matrix lives on device %d, the learner on device %d", m->device, c->device);
  USE_DEVICE(c->device);
  // (setup_dataset_common derives the byte columns from the metas; they must be the matrix's)
  int fs = 0, nf = 0;
  for (; fs < m->nfeat; ++nf) { const int ns = 1 + meta[nf].num_bin - (meta[nf].most_freq_bin == 0 ? 1 : 0); fs += ns > 256 ? 2 : 1; }
  NEED(fs == m->nfeat, "lgbcu_set_dataset_binned: the feature descriptors do not match the matrix's %d byte columns", m->nfeat);
  if (setup_dataset_common(c, m->nrow, nf, meta)) return -1;
  LAUNCH(k_pad_rows, grid_for(c, (long long)m->nrow * (c->stride / 4), 256), 256, 0, c->stream, (long long)m->nrow,
         m->nfeat, c->stride, m->d_bins, c->d_bins);
  LAUNCH(k_transpose_bins, grid_for(c, m->nrow, 256), 256, 0, c->stream, m->nrow, m->nfeat, c->stride, c->d_bins,
         c->d_bins_col, c->col_stride);
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

}  // extern "C"
