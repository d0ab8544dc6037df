/*
 * oracle/oracle.cpp — TEST INFRASTRUCTURE ONLY.
 *
 * CPU oracle: dependency-free C++17 restatement of the reference's scan-to-map registration hot path.
 * The product (libb200reg.so) never links, loads or executes this file; only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs use it, and only as the checker / reported baseline.
 *
 * PARITY STATUS: the reference cannot be built here (Eigen, boost, yaml-cpp, libnabo, Open3D absent), so this
 * restatement is pinned against the reference's own fixtures (car_cloud400/401 + validT3d, icpSingular, icpIdentity,
 * Conditioning tolerances; see tests/test_oracle_*.py).  Quantities the reference's tests never pin (kNN ids,
 * voxel membership, centroids, normals) are ORACLE-DEFINED ("parity unpinned" beyond the ICP-level fixtures).
 *
 * All citations are relative to /root/reference/.
 * Build: g++ -O3 -std=c++17 -ffp-contract=off -fopenmp -shared -fPIC oracle.cpp -o liboracle.so
 *        (-ffp-contract=off: every fp32 operation is individually rounded, as SURVEY 9.1 defines.)
 */
#include "oracle.h"

#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <numeric>
#include <tuple>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

constexpr float kInf = std::numeric_limits<float>::infinity();

// ------------------------------------------------------------------------------------------------------------
// small dense helpers (row-major, runtime n <= 6)
// ------------------------------------------------------------------------------------------------------------
template <typename S>
struct Mat4 {
  S m[16];
  static Mat4 identity() {
    Mat4 r;
    for (int i = 0; i < 16; ++i) r.m[i] = (i % 5 == 0) ? S(1) : S(0);
    return r;
  }
  S& operator()(int r, int c) { return m[4 * r + c]; }
  S operator()(int r, int c) const { return m[4 * r + c]; }
};

// 4x4 product, depth accumulated sequentially k = 0..3 (Eigen's GEBP order without FMA; SURVEY 9.1 contract)
template <typename S>
Mat4<S> mul(const Mat4<S>& a, const Mat4<S>& b) {
  Mat4<S> r;
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      S acc = a(i, 0) * b(0, j);
      acc = acc + a(i, 1) * b(1, j);
      acc = acc + a(i, 2) * b(2, j);
      acc = acc + a(i, 3) * b(3, j);
      r(i, j) = acc;
    }
  return r;
}

template <typename S>
S det3(const Mat4<S>& T) {
  return T(0, 0) * (T(1, 1) * T(2, 2) - T(1, 2) * T(2, 1)) - T(0, 1) * (T(1, 0) * T(2, 2) - T(1, 2) * T(2, 0)) +
         T(0, 2) * (T(1, 0) * T(2, 1) - T(1, 1) * T(2, 0));
}

// Eigen::Quaternion(Matrix3) (Eigen/src/Geometry/Quaternion.h, Shepperd's method) -> (w,x,y,z)
template <typename S>
void quatFromRot(const Mat4<S>& T, S q[4]) {
  S t = T(0, 0) + T(1, 1) + T(2, 2);
  if (t > S(0)) {
    t = std::sqrt(t + S(1));
    q[0] = S(0.5) * t;
    t = S(0.5) / t;
    q[1] = (T(2, 1) - T(1, 2)) * t;
    q[2] = (T(0, 2) - T(2, 0)) * t;
    q[3] = (T(1, 0) - T(0, 1)) * t;
  } else {
    int i = 0;
    if (T(1, 1) > T(0, 0)) i = 1;
    if (T(2, 2) > T(i, i)) i = 2;
    int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(T(i, i) - T(j, j) - T(k, k) + S(1));
    q[1 + i] = S(0.5) * t;
    t = S(0.5) / t;
    q[0] = (T(k, j) - T(j, k)) * t;
    q[1 + j] = (T(j, i) + T(i, j)) * t;
    q[1 + k] = (T(k, i) + T(i, k)) * t;
  }
}

// Eigen QuaternionBase::angularDistance: d = a * conj(b); 2*atan2(|d.vec|, |d.w|)
template <typename S>
S quatAngularDistance(const S a[4], const S b[4]) {
  const S bw = b[0], bx = -b[1], by = -b[2], bz = -b[3];
  const S w = a[0] * bw - a[1] * bx - a[2] * by - a[3] * bz;
  const S x = a[0] * bx + a[1] * bw + a[2] * bz - a[3] * by;
  const S y = a[0] * by + a[2] * bw + a[3] * bx - a[1] * bz;
  const S z = a[0] * bz + a[3] * bw + a[1] * by - a[2] * bx;
  return S(2) * std::atan2(std::sqrt(x * x + y * y + z * z), std::fabs(w));
}

// ------------------------------------------------------------------------------------------------------------
// fp32 squared distance, fixed operation order ((dx*dx + dy*dy) + dz*dz), every op rounded
// ------------------------------------------------------------------------------------------------------------
inline float dist2f(const float* a, const float* b) {
  const float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
  return (dx * dx + dy * dy) + dz * dz;
}

// sorted (d2,id) result list of capacity k, ascending, ties -> smaller id first
struct KnnList {
  int k, count = 0;
  float* d2;
  int32_t* id;
  KnnList(int k_, float* d, int32_t* i) : k(k_), d2(d), id(i) {
    for (int j = 0; j < k; ++j) { d2[j] = kInf; id[j] = -1; }
  }
  inline bool full() const { return count == k; }
  inline float worst() const { return d2[k - 1]; }
  inline void offer(float d, int32_t i) {
    if (count == k) {
      if (d > d2[k - 1] || (d == d2[k - 1] && i > id[k - 1])) return;
    }
    int pos = (count < k) ? count : k - 1;
    while (pos > 0 && (d2[pos - 1] > d || (d2[pos - 1] == d && id[pos - 1] > i))) {
      d2[pos] = d2[pos - 1];
      id[pos] = id[pos - 1];
      --pos;
    }
    d2[pos] = d;
    id[pos] = i;
    if (count < k) ++count;
  }
};

// ------------------------------------------------------------------------------------------------------------
// k-d tree — restates the *public contract* of libnabo as used at MatchersImpl.cpp:113,129 (libnabo itself is an
// un-vendored dependency, ">= 1.0.7"): squared L2 over the first 3 rows, radius test d2 < maxDist^2,
// (-1, inf) padding, ALLOW_SELF_MATCH, touched-point count returned.  Exact search (epsilon treated as 0), ties ->
// smallest reference index.  Bucket size 8 as libnabo's default.
// ------------------------------------------------------------------------------------------------------------
struct KdNode {
  float lo[3], hi[3];
  int32_t left, right;   // children (internal) or [begin,end) into idx (leaf, left = -1 - begin)
  int32_t begin, end;
};

struct KdTree {
  const float* feat;
  int stride;
  int64_t n;
  std::vector<int32_t> idx;
  std::vector<float> pts;  // leaf-ordered copy: x y z per point (cache-friendly scan)
  std::vector<KdNode> nodes;

  int build(int32_t b, int32_t e) {
    KdNode nd;
    for (int d = 0; d < 3; ++d) { nd.lo[d] = kInf; nd.hi[d] = -kInf; }
    for (int32_t i = b; i < e; ++i) {
      const float* p = feat + (int64_t)idx[i] * stride;
      for (int d = 0; d < 3; ++d) { nd.lo[d] = std::min(nd.lo[d], p[d]); nd.hi[d] = std::max(nd.hi[d], p[d]); }
    }
    nd.begin = b; nd.end = e; nd.left = nd.right = -1;
    const int me = (int)nodes.size();
    nodes.push_back(nd);
    if (e - b > 8) {
      int dim = 0;
      float ext = nd.hi[0] - nd.lo[0];
      for (int d = 1; d < 3; ++d) if (nd.hi[d] - nd.lo[d] > ext) { ext = nd.hi[d] - nd.lo[d]; dim = d; }
      if (ext > 0.f) {
        const int32_t mid = b + (e - b) / 2;
        std::nth_element(idx.begin() + b, idx.begin() + mid, idx.begin() + e, [&](int32_t x, int32_t y) {
          const float fx = feat[(int64_t)x * stride + dim], fy = feat[(int64_t)y * stride + dim];
          return fx < fy || (fx == fy && x < y);
        });
        const int l = build(b, mid);
        const int r = build(mid, e);
        nodes[me].left = l;
        nodes[me].right = r;
      }
    }
    return me;
  }

  static inline float boxLb(const KdNode& nd, const float* q) {
    float g[3];
    for (int d = 0; d < 3; ++d) {
      const float a = nd.lo[d] - q[d], b = q[d] - nd.hi[d];
      g[d] = std::max(0.f, std::max(a, b));
    }
    return (g[0] * g[0] + g[1] * g[1]) + g[2] * g[2];   // same op order as dist2f => monotone lower bound
  }

  void search(int node, const float* q, float maxd2, int64_t selfSkip, KnnList& res, uint64_t& visits) const {
    const KdNode& nd = nodes[node];
    if (nd.left < 0) {
      for (int32_t i = nd.begin; i < nd.end; ++i) {
        if (idx[i] == selfSkip) continue;
        const float d = dist2f(q, &pts[3 * (size_t)i]);
        ++visits;
        if (d < maxd2) res.offer(d, idx[i]);
      }
      return;
    }
    const float ll = boxLb(nodes[nd.left], q), lr = boxLb(nodes[nd.right], q);
    const int first = (ll <= lr) ? nd.left : nd.right, second = (ll <= lr) ? nd.right : nd.left;
    const float lf = std::min(ll, lr), ls = std::max(ll, lr);
    if (lf < maxd2 && !(res.full() && lf > res.worst())) search(first, q, maxd2, selfSkip, res, visits);
    if (ls < maxd2 && !(res.full() && ls > res.worst())) search(second, q, maxd2, selfSkip, res, visits);
  }
};

inline float maxDist2(float maxDist) { return maxDist * maxDist; }   // fp32 product (inf stays inf)

}  // namespace

extern "C" {

void* orc_kdtree_build(const float* feat, int32_t stride, int64_t n) {
  KdTree* t = new KdTree();
  t->feat = feat; t->stride = stride; t->n = n;
  t->idx.resize((size_t)n);
  std::iota(t->idx.begin(), t->idx.end(), 0);
  t->nodes.reserve((size_t)(n / 4 + 16));
  if (n > 0) t->build(0, (int32_t)n);
  t->pts.resize(3 * (size_t)n);
  for (int64_t i = 0; i < n; ++i)
    for (int d = 0; d < 3; ++d) t->pts[3 * i + d] = feat[(int64_t)t->idx[i] * stride + d];
  return t;
}

void orc_kdtree_free(void* tree) { delete (KdTree*)tree; }

uint64_t orc_kdtree_knn(const void* tree, const float* q, int32_t qstride, int64_t nq, int32_t k, float max_dist,
                        int32_t allow_self, int32_t* ids, float* d2, int32_t nthreads) {
  const KdTree* t = (const KdTree*)tree;
  const float md2 = maxDist2(max_dist);
  uint64_t visits = 0;
#pragma omp parallel for schedule(dynamic, 1024) num_threads(nthreads > 0 ? nthreads : 1) reduction(+ : visits)
  for (int64_t i = 0; i < nq; ++i) {
    KnnList res(k, d2 + (int64_t)k * i, ids + (int64_t)k * i);
    uint64_t v = 0;
    if (t->n > 0) t->search(0, q + i * qstride, md2, allow_self ? -1 : i, res, v);
    visits += v;
  }
  return visits;
}

uint64_t orc_knn_brute(const float* ref, int32_t rstride, int64_t m, const float* q, int32_t qstride, int64_t nq, int32_t k,
                       float max_dist, int32_t* ids, float* d2, int32_t nthreads) {
  const float md2 = maxDist2(max_dist);
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : 1)
  for (int64_t i = 0; i < nq; ++i) {
    KnnList res(k, d2 + (int64_t)k * i, ids + (int64_t)k * i);
    for (int64_t j = 0; j < m; ++j) {
      const float d = dist2f(q + i * qstride, ref + j * rstride);
      if (d < md2) res.offer(d, (int32_t)j);
    }
  }
  return (uint64_t)m * (uint64_t)nq;
}

}  // extern "C"

// ================================================================================================================
// RigidTransformation::inPlaceCompute / checkParameters  (libpointmatcher/pointmatcher/TransformationsImpl.cpp:61-114)
// ================================================================================================================
namespace {

// features <- T * features; normals <- R * normals.  fp32, depth accumulated k = 0..3, no FMA.
void transformCloud(const Mat4<float>& T, float* feat, int stride, float* normals, int64_t n) {
  for (int64_t i = 0; i < n; ++i) {
    float* p = feat + i * stride;
    const float x = p[0], y = p[1], z = p[2], w = (stride >= 4) ? p[3] : 1.f;
    for (int r = 0; r < 3; ++r) p[r] = ((T(r, 0) * x + T(r, 1) * y) + T(r, 2) * z) + T(r, 3) * w;
    if (stride >= 4) p[3] = ((T(3, 0) * x + T(3, 1) * y) + T(3, 2) * z) + T(3, 3) * w;
    if (normals) {
      float* nn = normals + 3 * i;
      const float a = nn[0], b = nn[1], c = nn[2];
      for (int r = 0; r < 3; ++r) nn[r] = (T(r, 0) * a + T(r, 1) * b) + T(r, 2) * c;
    }
  }
}

bool checkRigid(const Mat4<float>& T) {   // TransformationsImpl.cpp:100-114, epsilon 0.001
  return !(std::fabs(1.f - det3(T)) > 0.001f);
}

Mat4<float> loadT(const float T[16]) {
  Mat4<float> r;
  std::memcpy(r.m, T, sizeof(r.m));
  return r;
}

}  // namespace

extern "C" int32_t orc_transform(const float T[16], float* feat, int32_t stride, float* normals, int64_t n) {
  const Mat4<float> M = loadT(T);
  if (!checkRigid(M)) return ORC_E_TRANSFORM;
  transformCloud(M, feat, stride, normals, n);
  return ORC_OK;
}

// ================================================================================================================
// Outlier filters
// ================================================================================================================
namespace {

// Matches::getDistsQuantile (Matches.cpp:61-87): finite values only; rank = size_t(float(n) * quantile) (float
// multiply, truncation); quantile == 1 -> max.
int distsQuantile(const float* d2, int64_t count, float quantile, float* out) {
  std::vector<float> values;
  values.reserve((size_t)count);
  for (int64_t i = 0; i < count; ++i)
    if (d2[i] != kInf) values.push_back(d2[i]);
  if (values.empty()) return ORC_E_CONVERGENCE;
  if (quantile < 0.0f || quantile > 1.0f) return ORC_E_CONVERGENCE;
  if (quantile == 1.0f) {
    *out = *std::max_element(values.begin(), values.end());
    return ORC_OK;
  }
  const size_t rank = (size_t)((float)values.size() * quantile);
  std::nth_element(values.begin(), values.begin() + rank, values.end());
  *out = values[rank];
  return ORC_OK;
}

}  // namespace

extern "C" {

// TrimmedDistOutlierFilter::compute (OutlierFiltersImpl.cpp:140-147): w = (d2 <= limit)
// MaxDist / MinDist / MedianDist outlier filters (OutlierFiltersImpl.cpp:64-125): kind 0 w = (d2 <= fl(maxDist^2)), kind 1
// w = (d2 >= fl(minDist^2)) — pow(x, 2) on T = float is evaluated in double and rounded once, i.e. the fp32 product —, kind 2
// w = (d2 <= fl(factor * getDistsQuantile(0.5))).
int32_t orc_dist_outlier(int32_t kind, const float* d2, int64_t count, float param, float* limit_out, float* w_out) {
  float limit = 0.f;
  if (kind == 2) {
    float median = 0.f;
    const int st = distsQuantile(d2, count, 0.5f, &median);
    if (st != ORC_OK) return st;
    limit = param * median;
  } else {
    limit = (float)std::pow((double)param, 2);
  }
  for (int64_t i = 0; i < count; ++i) w_out[i] = ((kind == 1) ? (d2[i] >= limit) : (d2[i] <= limit)) ? 1.f : 0.f;
  if (limit_out) *limit_out = limit;
  return ORC_OK;
}

int32_t orc_trimmed_dist(const float* d2, int64_t count, float ratio, float* tau_out, float* w_out) {
  float tau = 0.f;
  const int st = distsQuantile(d2, count, ratio, &tau);
  if (st != ORC_OK) return st;
  if (tau_out) *tau_out = tau;
  if (w_out)
    for (int64_t i = 0; i < count; ++i) w_out[i] = (d2[i] <= tau) ? 1.f : 0.f;
  return ORC_OK;
}

// SurfaceNormalOutlierFilter::compute (OutlierFiltersImpl.cpp:236-281): eps = cos(maxAngle) in T = float;
// w = 0 if id invalid or dot < eps, else 1; all ones if either cloud lacks normals.
void orc_surface_normal_outlier(const float* nrm_read, const float* nrm_ref, const int32_t* ids, int32_t k, int64_t n,
                                float max_angle, float* w_out) {
  const float eps = std::cos(max_angle);
  for (int64_t i = 0; i < n; ++i)
    for (int j = 0; j < k; ++j) {
      const int64_t o = (int64_t)k * i + j;
      if (!nrm_read || !nrm_ref) { w_out[o] = 1.f; continue; }
      const int32_t id = ids[o];
      if (id == -1) { w_out[o] = 0.f; continue; }
      const float* a = nrm_read + 3 * i;
      const float* b = nrm_ref + 3 * (int64_t)id;
      const float v = (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2];
      w_out[o] = (v < eps) ? 0.f : 1.f;
    }
}

// OutlierFilters::compute with an empty chain (OutlierFilter.cpp:70-85)
void orc_no_filter_weights(const float* d2, int64_t count, float* w_out) {
  for (int64_t i = 0; i < count; ++i) w_out[i] = (d2[i] == kInf) ? 0.f : 1.f;
}

}  // extern "C"

// ================================================================================================================
// Small dense linear algebra used by PointToPlane::solvePossiblyUnderdeterminedLinearSystem (PointToPlane.cpp:185-238)
// — restates the Eigen calls at that site: fullPivHouseholderQr().isInvertible()/rank(), llt().solve(), and the
// rank-reduced minimum-norm solution.  Eigen is an un-vendored dependency; algorithms per its documentation.
// ================================================================================================================
namespace {

template <typename S>
struct FullPivQR {
  int n;
  std::vector<S> R;        // n x n row-major, upper triangular after factorisation
  std::vector<S> Q;        // n x n row-major, accumulated (A * P = Q * R)
  std::vector<int> colPerm;  // P: column j of A*P is column colPerm[j] of A
  int rank;
  bool invertible;

  // Householder QR with full pivoting (largest remaining |entry| moved to the pivot position by a row and a column
  // transposition; Eigen/src/QR/FullPivHouseholderQR.h).  rank = #pivots with |pivot| > eps * n * maxpivot.
  FullPivQR(const std::vector<S>& A, int n_) : n(n_), R(A), Q((size_t)n_ * n_, S(0)), colPerm(n_), rank(0) {
    for (int i = 0; i < n; ++i) { Q[(size_t)i * n + i] = S(1); colPerm[i] = i; }
    std::vector<S> pivots(n, S(0));
    S maxpivot = S(0);
    for (int k = 0; k < n; ++k) {
      int pr = k, pc = k;
      S best = S(-1);
      for (int i = k; i < n; ++i)
        for (int j = k; j < n; ++j)
          if (std::fabs(R[(size_t)i * n + j]) > best) { best = std::fabs(R[(size_t)i * n + j]); pr = i; pc = j; }
      if (best <= S(0)) break;
      if (pr != k) {   // row swap: R rows, and the matching columns of Q (Q * Rowswap * Rowswap * R)
        for (int j = 0; j < n; ++j) std::swap(R[(size_t)pr * n + j], R[(size_t)k * n + j]);
        for (int i = 0; i < n; ++i) std::swap(Q[(size_t)i * n + pr], Q[(size_t)i * n + k]);
      }
      if (pc != k) {
        for (int i = 0; i < n; ++i) std::swap(R[(size_t)i * n + pc], R[(size_t)i * n + k]);
        std::swap(colPerm[pc], colPerm[k]);
      }
      // Householder vector for column k, rows k..n-1
      S norm2 = S(0);
      for (int i = k; i < n; ++i) norm2 += R[(size_t)i * n + k] * R[(size_t)i * n + k];
      const S x0 = R[(size_t)k * n + k];
      const S alpha = (x0 >= S(0)) ? -std::sqrt(norm2) : std::sqrt(norm2);
      std::vector<S> v(n, S(0));
      v[k] = x0 - alpha;
      for (int i = k + 1; i < n; ++i) v[i] = R[(size_t)i * n + k];
      S vnorm2 = S(0);
      for (int i = k; i < n; ++i) vnorm2 += v[i] * v[i];
      if (vnorm2 > S(0)) {
        for (int j = k; j < n; ++j) {   // R <- (I - 2 v v^T / v^T v) R
          S dot = S(0);
          for (int i = k; i < n; ++i) dot += v[i] * R[(size_t)i * n + j];
          const S f = S(2) * dot / vnorm2;
          for (int i = k; i < n; ++i) R[(size_t)i * n + j] -= f * v[i];
        }
        for (int i = 0; i < n; ++i) {   // Q <- Q (I - 2 v v^T / v^T v)
          S dot = S(0);
          for (int j = k; j < n; ++j) dot += Q[(size_t)i * n + j] * v[j];
          const S f = S(2) * dot / vnorm2;
          for (int j = k; j < n; ++j) Q[(size_t)i * n + j] -= f * v[j];
        }
      }
      for (int i = k + 1; i < n; ++i) R[(size_t)i * n + k] = S(0);
      pivots[k] = std::fabs(R[(size_t)k * n + k]);
      maxpivot = std::max(maxpivot, pivots[k]);
    }
    // threshold in the reference's scalar T = float in BOTH precisions (the rank decision is part of the contract)
    const S thr = (S)std::numeric_limits<float>::epsilon() * S(n) * maxpivot;
    for (int k = 0; k < n; ++k)
      if (pivots[k] > thr) ++rank;
    invertible = (rank == n);
  }
};

// x = A.llt().solve(b) for SPD A (n x n row-major)
template <typename S>
void lltSolve(const std::vector<S>& A, const std::vector<S>& b, int n, std::vector<S>& x) {
  std::vector<S> L((size_t)n * n, S(0));
  for (int j = 0; j < n; ++j) {
    S d = A[(size_t)j * n + j];
    for (int k = 0; k < j; ++k) d -= L[(size_t)j * n + k] * L[(size_t)j * n + k];
    d = std::sqrt(d);
    L[(size_t)j * n + j] = d;
    for (int i = j + 1; i < n; ++i) {
      S s = A[(size_t)i * n + j];
      for (int k = 0; k < j; ++k) s -= L[(size_t)i * n + k] * L[(size_t)j * n + k];
      L[(size_t)i * n + j] = s / d;
    }
  }
  std::vector<S> y(n);
  for (int i = 0; i < n; ++i) {
    S s = b[i];
    for (int k = 0; k < i; ++k) s -= L[(size_t)i * n + k] * y[k];
    y[i] = s / L[(size_t)i * n + i];
  }
  x.assign(n, S(0));
  for (int i = n - 1; i >= 0; --i) {
    S s = y[i];
    for (int k = i + 1; k < n; ++k) s -= L[(size_t)k * n + i] * x[k];
    x[i] = s / L[(size_t)i * n + i];
  }
}

// Minimum-norm least-squares solve via one-sided Jacobi eigen-decomposition of the symmetric A in double — stands in
// for the `jacobiSvd(ComputeThinU|ComputeThinV).solve()` fallback at PointToPlane.cpp:221 (double precision there too).
void symPinvSolve(const std::vector<double>& A, const std::vector<double>& b, int n, std::vector<double>& x) {
  std::vector<double> a(A), V((size_t)n * n, 0.0);
  for (int i = 0; i < n; ++i) V[(size_t)i * n + i] = 1.0;
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0;
    for (int p = 0; p < n; ++p)
      for (int q = p + 1; q < n; ++q) off += a[(size_t)p * n + q] * a[(size_t)p * n + q];
    if (off < 1e-300) break;
    for (int p = 0; p < n; ++p)
      for (int q = p + 1; q < n; ++q) {
        const double apq = a[(size_t)p * n + q];
        if (apq == 0.0) continue;
        const double theta = (a[(size_t)q * n + q] - a[(size_t)p * n + p]) / (2.0 * apq);
        const double t = ((theta >= 0) ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
        const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < n; ++k) {
          const double akp = a[(size_t)k * n + p], akq = a[(size_t)k * n + q];
          a[(size_t)k * n + p] = c * akp - s * akq;
          a[(size_t)k * n + q] = s * akp + c * akq;
        }
        for (int k = 0; k < n; ++k) {
          const double apk = a[(size_t)p * n + k], aqk = a[(size_t)q * n + k];
          a[(size_t)p * n + k] = c * apk - s * aqk;
          a[(size_t)q * n + k] = s * apk + c * aqk;
        }
        for (int k = 0; k < n; ++k) {
          const double vkp = V[(size_t)k * n + p], vkq = V[(size_t)k * n + q];
          V[(size_t)k * n + p] = c * vkp - s * vkq;
          V[(size_t)k * n + q] = s * vkp + c * vkq;
        }
      }
  }
  double lmax = 0;
  for (int i = 0; i < n; ++i) lmax = std::max(lmax, std::fabs(a[(size_t)i * n + i]));
  const double thr = std::numeric_limits<double>::epsilon() * n * lmax;
  x.assign(n, 0.0);
  for (int i = 0; i < n; ++i) {
    const double l = a[(size_t)i * n + i];
    if (std::fabs(l) <= thr) continue;
    double proj = 0;
    for (int k = 0; k < n; ++k) proj += V[(size_t)k * n + i] * b[k];
    proj /= l;
    for (int k = 0; k < n; ++k) x[k] += V[(size_t)k * n + i] * proj;
  }
}

// solvePossiblyUnderdeterminedLinearSystem (PointToPlane.cpp:185-238)
template <typename S>
void solvePossiblyUnderdetermined(const std::vector<S>& A, const std::vector<S>& b, int n, std::vector<S>& x) {
  FullPivQR<S> qr(A, n);
  if (!qr.invertible) {
    const int rank = qr.rank;
    x.assign(n, S(0));
    if (rank == 0) return;
    // R1 = top `rank` rows of R (= (Q1^T A P)), upper-trapezoidal; rhs = Q1^T b
    std::vector<S> rhs(rank, S(0));
    for (int i = 0; i < rank; ++i)
      for (int k = 0; k < n; ++k) rhs[i] += qr.Q[(size_t)k * n + i] * b[k];
    // x_p = R1^T (R1 R1^T)^-1 rhs   (minimum-norm solution, :205-207)
    std::vector<S> G((size_t)rank * rank, S(0));
    for (int i = 0; i < rank; ++i)
      for (int j = 0; j < rank; ++j) {
        S s = S(0);
        for (int k = std::max(i, j); k < n; ++k) s += qr.R[(size_t)i * n + k] * qr.R[(size_t)j * n + k];
        G[(size_t)i * rank + j] = s;
      }
    std::vector<S> y;
    lltSolve(G, rhs, rank, y);
    std::vector<S> xp(n, S(0));
    for (int k = 0; k < n; ++k)
      for (int i = 0; i <= std::min(k, rank - 1); ++i) xp[k] += qr.R[(size_t)i * n + k] * y[i];
    for (int j = 0; j < n; ++j) x[qr.colPerm[j]] = xp[j];   // x = P * x_p (:214)
    // accuracy check (:216-232): !b.isApprox(A x, 1e-5) -> double-precision SVD re-solve
    S diff2 = S(0), nb2 = S(0), nax2 = S(0);
    for (int i = 0; i < n; ++i) {
      S ax = S(0);
      for (int k = 0; k < n; ++k) ax += A[(size_t)i * n + k] * x[k];
      diff2 += (b[i] - ax) * (b[i] - ax);
      nb2 += b[i] * b[i];
      nax2 += ax * ax;
    }
    if (!(diff2 <= S(1e-5) * S(1e-5) * std::min(nb2, nax2))) {
      std::vector<double> Ad(A.begin(), A.end()), bd(b.begin(), b.end()), xd;
      symPinvSolve(Ad, bd, n, xd);
      for (int i = 0; i < n; ++i) x[i] = (S)xd[i];
    }
  } else {
    lltSolve(A, b, n, x);   // :236
  }
}

}  // namespace

// ================================================================================================================
// ErrorElements (ErrorMinimizer.cpp:59-193) + PointToPlaneErrorMinimizer::compute (PointToPlane.cpp:240-368)
// S = float  : ORC_REF32  (means, centring, G G^T, solve, Rodrigues all in fp32 like the reference's T = float)
// S = double : ORC_MIXED  (same fp32 inputs, fp64 sums + solve — the contract of the CUDA path)
// ================================================================================================================
namespace {

struct Pair { int64_t read; int32_t ref; float w; };

template <typename S>
struct P2PlaneResult {
  int dim = 6;
  std::vector<S> A, b, x;
  Mat4<S> T;
  int kept = 0;
  float pointUsedRatio = 0, weightedPointUsedRatio = 0;
};

template <typename S>
int pointToPlane(const float* readFeat, int rstride, int64_t n, const float* refFeat, int fstride, const float* refNormals,
                 const int32_t* ids, const float* d2, const float* w, int k, bool force2D, bool force4DOF,
                 P2PlaneResult<S>& out) {
  if (!refNormals) return ORC_E_FIELD;   // getDescriptorViewByName("normals") throws InvalidField (PointToPlane.cpp:115)
  // --- ErrorElements: keep (k,i) pairs with finite dist and non-zero weight, reading-major order (:98-129)
  std::vector<Pair> pairs;
  int64_t nonZero = 0;
  for (int64_t i = 0; i < n * k; ++i) nonZero += (w[i] != 0.f);
  if (nonZero == 0) return ORC_E_CONVERGENCE;   // "no point to minimize" (:76-77)
  pairs.reserve((size_t)nonZero);
  float wsum = 0.f;
  for (int64_t i = 0; i < n; ++i)
    for (int j = 0; j < k; ++j) {
      const int64_t o = (int64_t)k * i + j;
      if (d2[o] == kInf) continue;
      if (w[o] != 0.f) {
        pairs.push_back({i, ids[o], w[o]});
        wsum += w[o];
      }
    }
  const int64_t K = (int64_t)pairs.size();
  out.kept = (int)K;
  out.pointUsedRatio = float(K) / float((int64_t)k * n);          // :139
  out.weightedPointUsedRatio = wsum / float((int64_t)k * n);      // :140
  if (K == 0) return ORC_E_CONVERGENCE;

  // --- means of the matched clouds (PointToPlane.cpp:263-264); force2D overwrites row 2 with ones first (:247-254)
  const int sdim = force2D ? 2 : 3;   // spatial dims that take part
  S meanR[3] = {0, 0, 0}, meanQ[3] = {0, 0, 0};
  for (const Pair& p : pairs)
    for (int d = 0; d < sdim; ++d) {
      meanR[d] += (S)readFeat[p.read * rstride + d];
      meanQ[d] += (S)refFeat[(int64_t)p.ref * fstride + d];
    }
  for (int d = 0; d < sdim; ++d) { meanR[d] /= (S)K; meanQ[d] /= (S)K; }

  // --- G = [cross; n], h = (p' - q') . n ; A = G G^T ; b = -G h^T   (:107-156; weights NOT applied, :153-155)
  const int dim = force2D ? 3 : (force4DOF ? 4 : 6);
  out.dim = dim;
  std::vector<S> A((size_t)dim * dim, S(0)), b(dim, S(0));
  for (const Pair& p : pairs) {
    S r[3] = {0, 0, 0}, q[3] = {0, 0, 0}, nn[3] = {0, 0, 0};
    for (int d = 0; d < sdim; ++d) {
      r[d] = (S)readFeat[p.read * rstride + d] - meanR[d];            // :267
      q[d] = (S)refFeat[(int64_t)p.ref * fstride + d] - meanQ[d];     // :268
      nn[d] = (S)refNormals[3 * (int64_t)p.ref + d];
    }
    S g[6];
    if (force2D) {          // 2D: cross = x*ny - y*nx (ErrorMinimizer.cpp:282-315, dim 3 branch)
      g[0] = r[0] * nn[1] - r[1] * nn[0];
      g[1] = nn[0]; g[2] = nn[1];
    } else if (force4DOF) { // diag((Gamma P)^T N) = -py*nx + px*ny   (PointToPlane.cpp:128-134)
      g[0] = -r[1] * nn[0] + r[0] * nn[1];
      g[1] = nn[0]; g[2] = nn[1]; g[3] = nn[2];
    } else {                // cross = p x n   (ErrorMinimizer.cpp:304-306)
      g[0] = r[1] * nn[2] - r[2] * nn[1];
      g[1] = r[2] * nn[0] - r[0] * nn[2];
      g[2] = r[0] * nn[1] - r[1] * nn[0];
      g[3] = nn[0]; g[4] = nn[1]; g[5] = nn[2];
    }
    S h = S(0);
    for (int d = 0; d < sdim; ++d) h += (r[d] - q[d]) * nn[d];       // :144-148
    for (int a = 0; a < dim; ++a) {
      for (int c = 0; c < dim; ++c) A[(size_t)a * dim + c] += g[a] * g[c];
      b[a] -= g[a] * h;
    }
  }
  std::vector<S> x;
  solvePossiblyUnderdetermined(A, b, dim, x);
  out.A = A; out.b = b; out.x = x;

  // --- parameters -> matrix (:282-366)
  Mat4<S> R = Mat4<S>::identity();
  S t[3] = {0, 0, 0};
  if (!force2D) {
    S axis[3] = {0, 0, 1}, angle;
    if (!force4DOF) {
      angle = std::sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);     // x.head(3).norm()
      // stableNormalized(): scale by the max |coeff| first; zero vector stays zero
      const S wmax = std::max(std::fabs(x[0]), std::max(std::fabs(x[1]), std::fabs(x[2])));
      S v[3] = {x[0], x[1], x[2]};
      if (wmax > S(0)) {
        const S z = (v[0] / wmax) * (v[0] / wmax) + (v[1] / wmax) * (v[1] / wmax) + (v[2] / wmax) * (v[2] / wmax);
        if (z > S(0)) for (int d = 0; d < 3; ++d) v[d] = v[d] / (std::sqrt(z) * wmax);
      }
      for (int d = 0; d < 3; ++d) axis[d] = v[d];
      t[0] = x[3]; t[1] = x[4]; t[2] = x[5];
    } else {
      angle = x[0];
      t[0] = x[1]; t[1] = x[2]; t[2] = x[3];
    }
    // Eigen::AngleAxis::toRotationMatrix
    const S s = std::sin(angle), c = std::cos(angle);
    const S cx = (S(1) - c) * axis[0], cy = (S(1) - c) * axis[1], cz = (S(1) - c) * axis[2];
    const S sx = s * axis[0], sy = s * axis[1], sz = s * axis[2];
    S tmp;
    tmp = cx * axis[1]; R(0, 1) = tmp - sz; R(1, 0) = tmp + sz;
    tmp = cx * axis[2]; R(0, 2) = tmp + sy; R(2, 0) = tmp - sy;
    tmp = cy * axis[2]; R(1, 2) = tmp - sx; R(2, 1) = tmp + sx;
    R(0, 0) = cx * axis[0] + c; R(1, 1) = cy * axis[1] + c; R(2, 2) = cz * axis[2] + c;
  } else {
    const S angle = x[0];
    R(0, 0) = std::cos(angle); R(0, 1) = -std::sin(angle); R(1, 0) = std::sin(angle); R(1, 1) = std::cos(angle);
    t[0] = x[1]; t[1] = x[2];
  }
  // T = Trans(meanQ) * [R t] * Trans(-meanR)   (:316-324 / :341-349)
  Mat4<S> T = R;
  for (int r = 0; r < 3; ++r) {
    S acc = S(0);
    for (int c = 0; c < 3; ++c) acc += R(r, c) * (-meanR[c]);
    T(r, 3) = acc + (t[r] + meanQ[r]);
  }
  bool hasNaN = false;
  for (int i = 0; i < 16; ++i) hasNaN |= (T.m[i] != T.m[i]);
  if (hasNaN)   // :326-332 — rotation block replaced by identity
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) T(r, c) = (r == c) ? S(1) : S(0);
  out.T = T;
  return ORC_OK;
}

template <typename S>
Mat4<float> toFloat(const Mat4<S>& a) {
  Mat4<float> r;
  for (int i = 0; i < 16; ++i) r.m[i] = (float)a.m[i];
  return r;
}

}  // namespace

extern "C" int32_t orc_point_to_plane(const float* read_feat, int32_t rstride, int64_t n, const float* ref_feat, int32_t fstride,
                                      const float* ref_normals, const int32_t* ids, const float* d2, const float* w, int32_t k,
                                      int32_t force2D, int32_t force4DOF, int32_t precision, float T_out[16], double* A_out,
                                      double* b_out, double* x_out, int32_t* kept_out, float* point_used_ratio,
                                      float* weighted_point_used_ratio) {
  auto emit = [&](auto& res) {
    const Mat4<float> Tf = toFloat(res.T);
    if (T_out) std::memcpy(T_out, Tf.m, sizeof(Tf.m));
    const int dim = res.dim;
    if (A_out) for (int i = 0; i < dim * dim; ++i) A_out[i] = (double)res.A[i];
    if (b_out) for (int i = 0; i < dim; ++i) b_out[i] = (double)res.b[i];
    if (x_out) for (int i = 0; i < dim; ++i) x_out[i] = (double)res.x[i];
  };
  int st;
  int kept = 0;
  float pur = 0, wpur = 0;
  if (precision == ORC_REF32) {
    P2PlaneResult<float> res;
    st = pointToPlane<float>(read_feat, rstride, n, ref_feat, fstride, ref_normals, ids, d2, w, k, force2D, force4DOF, res);
    kept = res.kept; pur = res.pointUsedRatio; wpur = res.weightedPointUsedRatio;
    if (st == ORC_OK) emit(res);
  } else {
    P2PlaneResult<double> res;
    st = pointToPlane<double>(read_feat, rstride, n, ref_feat, fstride, ref_normals, ids, d2, w, k, force2D, force4DOF, res);
    kept = res.kept; pur = res.pointUsedRatio; wpur = res.weightedPointUsedRatio;
    if (st == ORC_OK) emit(res);
  }
  if (kept_out) *kept_out = kept;
  if (point_used_ratio) *point_used_ratio = pur;
  if (weighted_point_used_ratio) *weighted_point_used_ratio = wpur;
  return st;
}

// PointToPlaneErrorMinimizer::getResidualError / computeResidualError (PointToPlane.cpp:372-422): over the ErrorElements pairs
// (finite distance, non-zero weight; ErrorMinimizer.cpp:100-118), sum of w ((reading - reference) . n_reference)^2; force2D drops z.
extern "C" int32_t orc_point_to_plane_residual(const float* read_feat, int32_t rstride, int64_t n, const float* ref_feat, int32_t fstride,
                                               const float* ref_normals, const int32_t* ids, const float* d2, const float* w, int32_t k,
                                               int32_t force2D, int32_t precision, double* residual_out) {
  double accD = 0.0;
  float accF = 0.f;
  const int dims = force2D ? 2 : 3;
  for (int64_t i = 0; i < n; ++i)
    for (int j = 0; j < k; ++j) {
      const size_t o = (size_t)k * i + j;
      if (d2[o] == kInf || w[o] == 0.f) continue;
      const float* r = read_feat + i * rstride;
      const float* s = ref_feat + (int64_t)ids[o] * fstride;
      const float* nn = ref_normals + 3 * (int64_t)ids[o];
      if (precision == ORC_REF32) {
        float dot = 0.f;
        for (int c = 0; c < dims; ++c) dot += (r[c] - s[c]) * nn[c];
        accF += w[o] * (dot * dot);
      } else {
        double dot = 0.0;
        for (int c = 0; c < dims; ++c) dot += ((double)r[c] - (double)s[c]) * (double)nn[c];
        accD += (double)w[o] * (dot * dot);
      }
    }
  *residual_out = (precision == ORC_REF32) ? (double)accF : accD;
  return ORC_OK;
}

// ================================================================================================================
// Transformation checkers (TransformationCheckersImpl.cpp:58-225) and the ICP driver (ICP.cpp:258-468)
// ================================================================================================================
namespace {

struct Checkers {
  const orc_icp_config& cfg;
  int counter = 0;
  std::vector<std::array<float, 4>> rotations;
  std::vector<std::array<float, 3>> translations;
  float initQ[4];
  float initT[3];
  explicit Checkers(const orc_icp_config& c) : cfg(c) {}

  void init(const Mat4<float>& T) {
    counter = 0;                                         // Counter::init :58-61
    rotations.clear(); translations.clear();             // Differential::init :103-124
    std::array<float, 4> q;
    quatFromRot(T, q.data());
    rotations.push_back(q);
    translations.push_back({T(0, 3), T(1, 3), T(2, 3)});
    quatFromRot(T, initQ);                               // Bound::init :198-211
    initT[0] = T(0, 3); initT[1] = T(1, 3); initT[2] = T(2, 3);
  }

  // returns ORC_OK or an error; sets iterate / maxIterReached.  Checker order: Differential, Counter, Bound is
  // irrelevant for the outcome except which one fires first within the same iteration.
  int check(const Mat4<float>& T, bool& iterate, bool& maxIterReached) {
    int status = ORC_OK;
    if (cfg.smooth_length >= 0) {                        // Differential::check :127-158
      std::array<float, 4> q;
      quatFromRot(T, q.data());
      rotations.push_back(q);
      translations.push_back({T(0, 3), T(1, 3), T(2, 3)});
      float cv0 = 0.f, cv1 = 0.f;
      const size_t sl = (size_t)cfg.smooth_length;
      if (rotations.size() > sl) {
        for (size_t i = rotations.size() - 1; i >= rotations.size() - sl && sl > 0; --i) {
          cv0 += std::fabs(quatAngularDistance(rotations[i].data(), rotations[i - 1].data()));
          const float dx = translations[i][0] - translations[i - 1][0], dy = translations[i][1] - translations[i - 1][1],
                      dz = translations[i][2] - translations[i - 1][2];
          cv1 += std::fabs(std::sqrt(dx * dx + dy * dy + dz * dz));
        }
        cv0 /= (float)sl; cv1 /= (float)sl;
        if (cv0 < cfg.min_diff_rot && cv1 < cfg.min_diff_trans) iterate = false;
      }
      if (cv0 != cv0 || cv1 != cv1) status = ORC_E_CONVERGENCE;
    }
    if (cfg.bound_rot >= 0.f && status == ORC_OK) {      // Bound::check :198-225
      float q[4];
      quatFromRot(T, q);
      const float rot = quatAngularDistance(q, initQ);
      const float dx = T(0, 3) - initT[0], dy = T(1, 3) - initT[1], dz = T(2, 3) - initT[2];
      const float tr = std::sqrt(dx * dx + dy * dy + dz * dz);
      if (rot > cfg.bound_rot || tr > cfg.bound_trans) status = ORC_E_CONVERGENCE;
    }
    if (cfg.max_iter >= 0 && status == ORC_OK) {          // Counter::check :64-76 (throws, caught at ICP.cpp:441-445)
      ++counter;
      if (counter >= cfg.max_iter) { iterate = false; maxIterReached = true; }
    }
    return status;
  }
};

double nowSec() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

namespace {
struct IcpRef {   // state kept by ICP::initReference: referenceFiltered (centred), T_refIn_refMean, the matcher's tree
  std::vector<float> ref;
  std::vector<float> normals;
  bool hasNormals = false;
  int64_t m = 0;
  float mean[3] = {0, 0, 0};
  void* tree = nullptr;
  double t_reference_s = 0;
};
}  // namespace

// ICP::initReference (ICP.cpp:293-328): copy, (no reference filters here), mean, centre, matcher->init
extern "C" void* orc_icp_ref_create(const orc_icp_config* cfg, const float* ref_feat, int32_t ref_stride, const float* ref_normals, int64_t m) {
  if (m == 0) return nullptr;
  const double t0 = nowSec();
  IcpRef* R = new IcpRef();
  R->m = m;
  R->ref.resize(4 * (size_t)m);
  double sum[3] = {0, 0, 0};
  for (int64_t i = 0; i < m; ++i) {
    for (int d = 0; d < 3; ++d) { R->ref[4 * i + d] = ref_feat[i * ref_stride + d]; sum[d] += R->ref[4 * i + d]; }
    R->ref[4 * i + 3] = 1.f;
  }
  // meanReference = features.rowwise().mean() (:313).  Eigen's fp32 reduction order is unspecified; the contract is the
  // fp64 sum rounded once to fp32.
  for (int d = 0; d < 3; ++d) R->mean[d] = (float)(sum[d] / (double)m);
  for (int64_t i = 0; i < m; ++i)
    for (int d = 0; d < 3; ++d) R->ref[4 * i + d] = R->ref[4 * i + d] - R->mean[d];   // :320
  if (ref_normals) { R->normals.assign(ref_normals, ref_normals + 3 * (size_t)m); R->hasNormals = true; }
  if (cfg->matcher_kind == 0) R->tree = orc_kdtree_build(R->ref.data(), 4, m);
  R->t_reference_s = nowSec() - t0;
  return R;
}

extern "C" void orc_icp_ref_free(void* h) {
  IcpRef* R = (IcpRef*)h;
  if (!R) return;
  if (R->tree) orc_kdtree_free(R->tree);
  delete R;
}

// ICP::computeWithTransformedReference (ICP.cpp:332-468) against a prepared reference
extern "C" int32_t orc_icp_compute(const orc_icp_config* cfg, const void* refHandle, const float* read_feat, int32_t read_stride,
                                   const float* read_normals, int64_t n, const float T_init[16], float T_out[16], orc_icp_stats* stats) {
  orc_icp_stats st;
  std::memset(&st, 0, sizeof(st));
  auto finish = [&](int code) { st.status = code; if (stats) *stats = st; return code; };
  const IcpRef* R = (const IcpRef*)refHandle;
  if (!R) return finish(ORC_E_EMPTY);
  const std::vector<float>& ref = R->ref;
  const float* ref_normals = R->hasNormals ? R->normals.data() : nullptr;
  const float* mean = R->mean;
  void* tree = nullptr;            // owned by the handle; never freed here
  const void* ctree = R->tree;
  for (int d = 0; d < 3; ++d) st.mean_ref[d] = mean[d];
  st.t_reference_s = R->t_reference_s;
  Mat4<float> T_refIn_refMean = Mat4<float>::identity();
  for (int d = 0; d < 3; ++d) T_refIn_refMean(d, 3) = mean[d];
  const double t1 = nowSec();

  // ---- computeWithTransformedReference (ICP.cpp:332-468)
  if (n == 0) { if (tree) orc_kdtree_free(tree); return finish(ORC_E_EMPTY); }
  std::vector<float> reading(4 * (size_t)n), readingN;
  for (int64_t i = 0; i < n; ++i) {
    for (int d = 0; d < 3; ++d) reading[4 * i + d] = read_feat[i * read_stride + d];
    reading[4 * i + 3] = 1.f;
  }
  if (read_normals) readingN.assign(read_normals, read_normals + 3 * (size_t)n);
  // meanReading forced to zero in this fork (:363-364) => T_readIn_readMean = I
  Mat4<float> T_refMean_refIn = Mat4<float>::identity();   // inverse of a pure translation
  for (int d = 0; d < 3; ++d) T_refMean_refIn(d, 3) = -mean[d];
  const Mat4<float> Tinit = loadT(T_init);
  const Mat4<float> T_refMean_readMean = mul(T_refMean_refIn, Tinit);   // :373
  if (!checkRigid(T_refMean_readMean)) { if (tree) orc_kdtree_free(tree); return finish(ORC_E_TRANSFORM); }
  transformCloud(T_refMean_readMean, reading.data(), 4, read_normals ? readingN.data() : nullptr, n);   // :375

  Mat4<float> T_iter = Mat4<float>::identity();
  bool iterate = true, maxIterReached = false;
  Checkers checkers(*cfg);
  checkers.init(T_iter);

  const int k = (cfg->matcher_kind == 1) ? 1 : cfg->knn;
  std::vector<float> step(4 * (size_t)n), stepN, d2((size_t)k * n), w((size_t)k * n), w2;
  std::vector<int32_t> ids((size_t)k * n);
  int status = ORC_OK;
  int iterations = 0;
  while (iterate) {
    step = reading;                                             // :401
    if (read_normals) stepN = readingN;
    if (!checkRigid(T_iter)) { status = ORC_E_TRANSFORM; break; }
    transformCloud(T_iter, step.data(), 4, read_normals ? stepN.data() : nullptr, n);   // :409
    if (cfg->matcher_kind == 1) {                               // MirrorMatcher (MatchersImpl.cpp:65-84)
      for (int64_t i = 0; i < n; ++i) { ids[i] = (int32_t)i; d2[i] = 0.f; }
    } else {                                                    // :413
      st.points_touched += orc_kdtree_knn(ctree, step.data(), 4, n, k, cfg->max_dist, 1, ids.data(), d2.data(), cfg->nthreads);
    }
    // outlier filters (:417, OutlierFilter.cpp:64-103)
    bool any = false;
    if (cfg->trim_ratio > 0.f) {
      status = orc_trimmed_dist(d2.data(), (int64_t)k * n, cfg->trim_ratio, &st.last_tau, w.data());
      if (status != ORC_OK) break;
      any = true;
    }
    if (cfg->max_angle >= 0.f) {
      if (!any) {
        orc_surface_normal_outlier(read_normals ? stepN.data() : nullptr, ref_normals, ids.data(), k, n, cfg->max_angle, w.data());
      } else {
        w2.resize(w.size());
        orc_surface_normal_outlier(read_normals ? stepN.data() : nullptr, ref_normals, ids.data(), k, n, cfg->max_angle, w2.data());
        for (size_t i = 0; i < w.size(); ++i) w[i] = w[i] * w2[i];
      }
      any = true;
    }
    if (!any) orc_no_filter_weights(d2.data(), (int64_t)k * n, w.data());
    // error minimisation (:433)
    Mat4<float> Tstep;
    int kept = 0;
    status = orc_point_to_plane(step.data(), 4, n, ref.data(), 4, ref_normals, ids.data(), d2.data(), w.data(), k, cfg->force2D,
                                cfg->force4DOF, cfg->precision, Tstep.m, nullptr, nullptr, nullptr, &kept, &st.point_used_ratio,
                                &st.weighted_point_used_ratio);
    st.kept_last = kept;
    if (status != ORC_OK) break;
    T_iter = mul(Tstep, T_iter);
    status = checkers.check(T_iter, iterate, maxIterReached);   // :437-445
    ++iterations;
    if (status != ORC_OK) break;
  }
  if (tree) orc_kdtree_free(tree);
  st.iterations = iterations;
  st.max_iterations_reached = maxIterReached ? 1 : 0;
  st.t_converge_s = nowSec() - t1;
  if (status != ORC_OK) return finish(status);
  // :459-465  T_refIn_refMean * (T_iter * T_refMean_readMean) * T_readIn_readMean^-1 (= I)
  const Mat4<float> out = mul(T_refIn_refMean, mul(T_iter, T_refMean_readMean));
  std::memcpy(T_out, out.m, sizeof(out.m));
  return finish(ORC_OK);
}

// ICP::compute (ICP.cpp:258-290) = initReference + computeWithTransformedReference
extern "C" int32_t orc_icp(const orc_icp_config* cfg, const float* ref_feat, int32_t ref_stride, const float* ref_normals, int64_t m,
                           const float* read_feat, int32_t read_stride, const float* read_normals, int64_t n,
                           const float T_init[16], float T_out[16], orc_icp_stats* stats) {
  void* R = orc_icp_ref_create(cfg, ref_feat, ref_stride, ref_normals, m);
  if (!R) {
    if (stats) { std::memset(stats, 0, sizeof(*stats)); stats->status = ORC_E_EMPTY; }
    return ORC_E_EMPTY;
  }
  const int32_t code = orc_icp_compute(cfg, R, read_feat, read_stride, read_normals, n, T_init, T_out, stats);
  orc_icp_ref_free(R);
  return code;
}

// ================================================================================================================
// VoxelGridDataPointsFilter::inPlaceFilter (DataPointsFilters/VoxelGrid.cpp:71-344), 3D clouds (featDim == 4)
// ================================================================================================================
extern "C" int64_t orc_voxelgrid(float* feat, int32_t stride, float* desc, int32_t desc_dim, int64_t n, float vx, float vy, float vz,
                                 int32_t use_centroid, int32_t average_descriptors, uint32_t* cell_of_point_out) {
  if (n == 0) return 0;
  float mn[3] = {kInf, kInf, kInf}, mx[3] = {-kInf, -kInf, -kInf};
  for (int64_t i = 0; i < n; ++i)
    for (int d = 0; d < 3; ++d) { mn[d] = std::min(mn[d], feat[i * stride + d]); mx[d] = std::max(mx[d], feat[i * stride + d]); }
  const float vs[3] = {vx, vy, vz};
  float minB[3], maxB[3];
  for (int d = 0; d < 3; ++d) { minB[d] = mn[d] / vs[d]; maxB[d] = mx[d] / vs[d]; }   // :100-111
  const unsigned numDivX = (unsigned)(1 + maxB[0] - minB[0]);                          // :115-121 (float arithmetic)
  const unsigned numDivY = (unsigned)(1 + maxB[1] - minB[1]);
  const unsigned numDivZ = (unsigned)(1 + maxB[2] - minB[2]);
  const unsigned numVox = numDivX * numDivY * numDivZ;                                 // unsigned wrap like the reference
  if (numVox == 0) return ORC_E_PARAM;                                                 // :127-130
  struct Voxel { unsigned numPoints = 0, firstPoint = 0; };
  // the reference allocates a dense vector<Voxel>(numVox) (:143-153); a map gives the same membership without the allocation
  std::map<unsigned, Voxel> voxels;
  std::vector<unsigned> indices((size_t)n);
  for (int64_t p = 0; p < n; ++p) {                                                    // :155-182
    const unsigned i = (unsigned)std::floor(feat[p * stride + 0] / vx - minB[0]);
    const unsigned j = (unsigned)std::floor(feat[p * stride + 1] / vy - minB[1]);
    const unsigned k = (unsigned)std::floor(feat[p * stride + 2] / vz - minB[2]);
    const unsigned idx = i + j * numDivX + k * numDivX * numDivY;
    Voxel& v = voxels[idx];
    if (++v.numPoints == 1) v.firstPoint = (unsigned)p;
    indices[p] = idx;
    if (cell_of_point_out) cell_of_point_out[p] = idx;
  }
  std::vector<unsigned> keep;
  if (use_centroid) {
    for (int64_t p = 0; p < n; ++p) {                                                  // :192-222 sums in input order, fp32
      const unsigned fp = voxels[indices[p]].firstPoint;
      if (fp != (unsigned)p) {
        for (int f = 0; f < 3; ++f) feat[(int64_t)fp * stride + f] += feat[p * stride + f];
        if (average_descriptors && desc)
          for (int d = 0; d < desc_dim; ++d) desc[(int64_t)fp * desc_dim + d] += desc[p * desc_dim + d];
      }
    }
    for (auto& kv : voxels) {                                                          // :227-243
      const unsigned np = kv.second.numPoints, fp = kv.second.firstPoint;
      for (int f = 0; f < 3; ++f) feat[(int64_t)fp * stride + f] /= (float)np;
      if (average_descriptors && desc)
        for (int d = 0; d < desc_dim; ++d) desc[(int64_t)fp * desc_dim + d] /= (float)np;
      keep.push_back(fp);
    }
  } else {
    // cell-centre mode.  The reference's branch (:285-304) writes rows 1,2,3 and omits the bbox origin;
    // the contract implemented here (and on the GPU) is the corrected one: centre = (minB + idx + 0.5) * vSize.
    if (average_descriptors && desc)
      for (int64_t p = 0; p < n; ++p) {
        const unsigned fp = voxels[indices[p]].firstPoint;
        if (fp != (unsigned)p)
          for (int d = 0; d < desc_dim; ++d) desc[(int64_t)fp * desc_dim + d] += desc[p * desc_dim + d];
      }
    for (auto& kv : voxels) {
      const unsigned idx = kv.first, np = kv.second.numPoints, fp = kv.second.firstPoint;
      const unsigned k = idx / (numDivX * numDivY);
      const unsigned j = (idx - k * numDivX * numDivY) / numDivX;
      const unsigned i = idx - k * numDivX * numDivY - j * numDivX;
      feat[(int64_t)fp * stride + 0] = (minB[0] + (float)i + 0.5f) * vx;
      feat[(int64_t)fp * stride + 1] = (minB[1] + (float)j + 0.5f) * vy;
      feat[(int64_t)fp * stride + 2] = (minB[2] + (float)k + 0.5f) * vz;
      if (average_descriptors && desc)
        for (int d = 0; d < desc_dim; ++d) desc[(int64_t)fp * desc_dim + d] /= (float)np;
      keep.push_back(fp);
    }
  }
  std::sort(keep.begin(), keep.end());                                                 // :322
  for (size_t o = 0; o < keep.size(); ++o) {                                           // :324-334
    const unsigned kq = keep[o];
    for (int f = 0; f < stride; ++f) feat[o * stride + f] = feat[(int64_t)kq * stride + f];
    if (desc)
      for (int d = 0; d < desc_dim; ++d) desc[o * desc_dim + d] = desc[(int64_t)kq * desc_dim + d];
  }
  return (int64_t)keep.size();
}

// ================================================================================================================
// SurfaceNormalDataPointsFilter::inPlaceFilter (DataPointsFilters/SurfaceNormal.cpp:83-330)
// ================================================================================================================
namespace {

// symmetric 3x3 eigen-decomposition (cyclic Jacobi, double) — stands in for Eigen::EigenSolver at SurfaceNormal.cpp:212-214
// (general real solver on a symmetric PSD matrix; eigenvector signs/order are not pinned by any reference test).
// Output: eigenvalues ascending, eigenvectors as columns v[r][c].
void symEig3(const double Cin[9], double val[3], double vec[3][3]) {
  double a[3][3], V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) a[i][j] = Cin[3 * i + j];
  for (int sweep = 0; sweep < 64; ++sweep) {
    const double off = a[0][1] * a[0][1] + a[0][2] * a[0][2] + a[1][2] * a[1][2];
    if (off == 0.0) break;
    for (int p = 0; p < 3; ++p)
      for (int q = p + 1; q < 3; ++q) {
        // converged pair: an off-diagonal below 2^-70 of its diagonals rotates by an angle no fp64 result can see; without this
        // the residue left by each rotation never reaches exactly 0 and most matrices run all 64 sweeps
        if (fabs(a[p][q]) <= 8.470329472543003e-22 * (fabs(a[p][p]) + fabs(a[q][q]))) { a[p][q] = 0.0; a[q][p] = 0.0; continue; }
        const double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
        const double t = ((theta >= 0) ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
        const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < 3; ++k) { const double x = a[k][p], y = a[k][q]; a[k][p] = c * x - s * y; a[k][q] = s * x + c * y; }
        for (int k = 0; k < 3; ++k) { const double x = a[p][k], y = a[q][k]; a[p][k] = c * x - s * y; a[q][k] = s * x + c * y; }
        for (int k = 0; k < 3; ++k) { const double x = V[k][p], y = V[k][q]; V[k][p] = c * x - s * y; V[k][q] = s * x + c * y; }
      }
  }
  int order[3] = {0, 1, 2};
  std::sort(order, order + 3, [&](int x, int y) { return a[x][x] < a[y][y]; });
  for (int c = 0; c < 3; ++c) {
    val[c] = a[order[c]][order[c]];
    for (int r = 0; r < 3; ++r) vec[r][c] = V[r][order[c]];
  }
}

}  // namespace

extern "C" int32_t orc_surface_normals(const float* feat, int32_t stride, int64_t n, int32_t knn, float max_dist, int32_t precision,
                                       int32_t nthreads, const float* orient_origin, float* normals, float* densities,
                                       float* eig_values, float* eig_vectors, float* matched_ids, float* mean_dists,
                                       float* lin_plan_curv, int32_t* degenerate_count) {
  if (knn < 3) return ORC_E_PARAM;   // SurfaceNormal.h:69 (min 3)
  // self-kNN incl. the point itself through an own KDTreeMatcher (:170-179)
  std::vector<int32_t> ids((size_t)knn * n);
  std::vector<float> d2((size_t)knn * n);
  void* tree = orc_kdtree_build(feat, stride, n);
  orc_kdtree_knn(tree, feat, stride, n, knn, max_dist, 1, ids.data(), d2.data(), nthreads);
  orc_kdtree_free(tree);
  int degenerate = 0;
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : 1) reduction(+ : degenerate)
  for (int64_t i = 0; i < n; ++i) {
    // gather neighbours (:190-198)
    std::vector<double> d;   // 3 x realKnn, values are fp32 coordinates
    d.reserve(3 * (size_t)knn);
    int realKnn = 0;
    for (int j = 0; j < knn; ++j)
      if (d2[(size_t)knn * i + j] != kInf) {
        const float* p = feat + (int64_t)ids[(size_t)knn * i + j] * stride;
        d.push_back(p[0]); d.push_back(p[1]); d.push_back(p[2]);
        ++realKnn;
      }
    double mean[3] = {0, 0, 0}, C[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    double maxSq = 0;
    if (precision == ORC_REF32) {   // fp32 mean / NN / covariance (:201-204)
      float mf[3] = {0, 0, 0};
      for (int j = 0; j < realKnn; ++j) for (int c = 0; c < 3; ++c) mf[c] += (float)d[3 * j + c];
      for (int c = 0; c < 3; ++c) mf[c] /= (float)realKnn;
      float Cf[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      float maxSqF = 0.f;
      for (int j = 0; j < realKnn; ++j) {
        float nn[3];
        for (int c = 0; c < 3; ++c) nn[c] = (float)d[3 * j + c] - mf[c];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) Cf[3 * r + c] += nn[r] * nn[c];
        maxSqF = std::max(maxSqF, (nn[0] * nn[0] + nn[1] * nn[1]) + nn[2] * nn[2]);
      }
      for (int q = 0; q < 9; ++q) C[q] = Cf[q] / (float)realKnn;
      for (int c = 0; c < 3; ++c) mean[c] = mf[c];
      maxSq = maxSqF;
    } else {
      for (int j = 0; j < realKnn; ++j) for (int c = 0; c < 3; ++c) mean[c] += d[3 * j + c];
      for (int c = 0; c < 3; ++c) mean[c] /= (double)realKnn;
      for (int j = 0; j < realKnn; ++j) {
        double nn[3];
        for (int c = 0; c < 3; ++c) nn[c] = d[3 * j + c] - mean[c];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) C[3 * r + c] += nn[r] * nn[c];
        maxSq = std::max(maxSq, nn[0] * nn[0] + nn[1] * nn[1] + nn[2] * nn[2]);
      }
      for (int q = 0; q < 9; ++q) C[q] /= (double)realKnn;
    }
    // rank test: C.fullPivHouseholderQr().rank() + 1 >= 3 (:210), in the reference's T = float
    bool isDegenerate;
    {
      std::vector<float> Cf(9);
      for (int q = 0; q < 9; ++q) Cf[q] = (float)C[q];
      FullPivQR<float> qr(Cf, 3);
      isDegenerate = !(qr.rank + 1 >= 3);
      if (realKnn == 0) isDegenerate = true;
    }
    double val[3] = {0, 0, 0}, vec[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    if (!isDegenerate) symEig3(C, val, vec); else ++degenerate;
    if (normals) {   // eigenvector of the smallest eigenvalue (:237-246), clamped to [-1,1]
      double nv[3] = {vec[0][0], vec[1][0], vec[2][0]};
      if (orient_origin && !isDegenerate) {   // SurfaceNormalEstimatorPCA.cpp:69-84 semantics
        const float* p = feat + i * stride;
        const double dot = (orient_origin[0] - p[0]) * nv[0] + (orient_origin[1] - p[1]) * nv[1] + (orient_origin[2] - p[2]) * nv[2];
        if (dot < 0) for (int c = 0; c < 3; ++c) nv[c] = -nv[c];
      }
      for (int c = 0; c < 3; ++c) normals[3 * i + c] = (float)std::min(1.0, std::max(-1.0, nv[c]));
    }
    if (densities) {   // utils.h:105-128
      if (isDegenerate) densities[i] = 0.f;
      else {
        const float m2 = (float)maxSq;
        const float volume = (float)(4. / 3.) * (float)M_PI * (m2 * std::sqrt(m2));
        densities[i] = (volume > 0) ? ((float)realKnn / volume) : 0.f;
      }
    }
    if (eig_values) for (int c = 0; c < 3; ++c) eig_values[3 * i + c] = (float)val[c];
    if (eig_vectors)   // serializeEigVec: row-major (utils.h:90-101)
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) eig_vectors[9 * i + 3 * r + c] = (float)vec[r][c];
    if (mean_dists) {
      if (isDegenerate) mean_dists[i] = (float)std::numeric_limits<std::size_t>::max();
      else {
        const float* p = feat + i * stride;
        const double dx = p[0] - mean[0], dy = p[1] - mean[1], dz = p[2] - mean[2];
        mean_dists[i] = (float)std::sqrt(dx * dx + dy * dy + dz * dz);
      }
    }
    if (lin_plan_curv) {   // SurfaceNormalEstimatorPCA.cpp:86-99 on ascending eigenvalues
      if (isDegenerate) { lin_plan_curv[3 * i] = lin_plan_curv[3 * i + 1] = lin_plan_curv[3 * i + 2] = 0.f; }
      else {
        const float s1 = std::sqrt((float)val[0]), s2 = std::sqrt((float)val[1]), s3 = std::sqrt((float)val[2]);
        lin_plan_curv[3 * i + 0] = (s3 - s2) / s3;
        lin_plan_curv[3 * i + 1] = (s2 - s1) / s3;
        lin_plan_curv[3 * i + 2] = s1 / (s1 + s2 + s3);
      }
    }
    if (matched_ids) for (int j = 0; j < knn; ++j) matched_ids[(size_t)knn * i + j] = (float)ids[(size_t)knn * i + j];   // :294-297
  }
  if (degenerate_count) *degenerate_count = degenerate;
  return ORC_OK;
}

// smoothNormals (SurfaceNormal.cpp:299-323): one sweep in index order that OVERWRITES normals->col(i) as it goes, so point i
// averages the already smoothed normals of its neighbours with a smaller index and the original ones of the others (itself
// included); each neighbour's normal is flipped to the side of the point's own (original) normal; T = float sums in neighbour order.
extern "C" int32_t orc_smooth_normals(const float* feat, int32_t stride, int64_t n, int32_t knn, float max_dist, int32_t nthreads,
                                      float* normals /* in/out, 3 per point */) {
  if (knn < 3) return ORC_E_PARAM;
  std::vector<int32_t> ids((size_t)knn * n);
  std::vector<float> d2((size_t)knn * n);
  void* tree = orc_kdtree_build(feat, stride, n);
  orc_kdtree_knn(tree, feat, stride, n, knn, max_dist, 1, ids.data(), d2.data(), nthreads);
  orc_kdtree_free(tree);
  for (int64_t i = 0; i < n; ++i) {   // sequential by construction
    const float cur[3] = {normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]};
    float mean[3] = {0.f, 0.f, 0.f};
    int cnt = 0;
    for (int j = 0; j < knn; ++j) {
      if (d2[(size_t)knn * i + j] == kInf) continue;
      const float* nb = normals + 3 * (int64_t)ids[(size_t)knn * i + j];
      const float dot = (cur[0] * nb[0] + cur[1] * nb[1]) + cur[2] * nb[2];
      if (dot > 0.f) { mean[0] += nb[0]; mean[1] += nb[1]; mean[2] += nb[2]; }
      else { mean[0] -= nb[0]; mean[1] -= nb[1]; mean[2] -= nb[2]; }
      ++cnt;
    }
    for (int c = 0; c < 3; ++c) normals[3 * i + c] = mean[c] / (float)cnt;
  }
  return ORC_OK;
}

// estimateNormalsOrCovariancesIfNeeded (open3d_slam/src/CloudRegistration.cpp:62-75) -> Open3D v0.15.1 PointCloud::EstimateNormals with
// KDTreeSearchParamHybrid(radius, knn) + OrientNormalsTowardsCameraLocation, in Open3D's scalar (double): the (up to) knn nearest
// points with d2 < radius^2 (the point itself included), covariance of those neighbours, eigenvector of the smallest eigenvalue,
// flipped toward `origin`.  Fewer than 3 neighbours -> zero normal (counted as degenerate; Open3D's own value there is unpinned).
// exact != 0: neighbours by brute force in fp64 (independent of any search structure; small clouds).  exact == 0: the knn + 2 nearest
// of the fp32 k-d tree are re-ranked in fp64 (the same candidate rule the device uses; for clouds too large for brute force).
extern "C" int32_t orc_estimate_normals_f64(const double* xyz, int64_t n, int32_t knn, double radius, const double* origin, int32_t exact,
                                            int32_t nthreads, double* normals, int32_t* degenerate_count) {
  if (knn < 3) return ORC_E_PARAM;
  const double r2 = radius * radius;
  std::vector<int32_t> cand;
  int kc = 0;
  if (!exact) {
    kc = knn + 2;
    std::vector<float> f((size_t)4 * n);
    for (int64_t i = 0; i < n; ++i) { for (int c = 0; c < 3; ++c) f[4 * i + c] = (float)xyz[3 * i + c]; f[4 * i + 3] = 1.f; }
    cand.resize((size_t)kc * n);
    std::vector<float> d2f((size_t)kc * n);
    void* tree = orc_kdtree_build(f.data(), 4, n);
    orc_kdtree_knn(tree, f.data(), 4, n, kc, (float)(radius * 1.00001 + 1e-6), 1, cand.data(), d2f.data(), nthreads);
    orc_kdtree_free(tree);
  }
  int degenerate = 0;
#pragma omp parallel for schedule(dynamic, 64) num_threads(nthreads > 0 ? nthreads : 1) reduction(+ : degenerate)
  for (int64_t i = 0; i < n; ++i) {
    std::vector<std::pair<double, int32_t>> nb;
    const double* p = xyz + 3 * i;
    auto consider = [&](int32_t j) {
      const double* q = xyz + 3 * (int64_t)j;
      const double dx = p[0] - q[0], dy = p[1] - q[1], dz = p[2] - q[2];
      const double d = (dx * dx + dy * dy) + dz * dz;
      if (d < r2) nb.emplace_back(d, j);
    };
    if (exact) for (int64_t j = 0; j < n; ++j) consider((int32_t)j);
    else for (int c = 0; c < kc; ++c) { const int32_t j = cand[(size_t)kc * i + c]; if (j >= 0) consider(j); }
    std::sort(nb.begin(), nb.end());
    if ((int)nb.size() > knn) nb.resize(knn);
    double nv[3] = {0, 0, 0};
    if (nb.size() < 3) ++degenerate;
    else {
      double mean[3] = {0, 0, 0}, C[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (auto& e : nb) for (int c = 0; c < 3; ++c) mean[c] += xyz[3 * (int64_t)e.second + c];
      for (int c = 0; c < 3; ++c) mean[c] /= (double)nb.size();
      for (auto& e : nb) {
        double d[3];
        for (int c = 0; c < 3; ++c) d[c] = xyz[3 * (int64_t)e.second + c] - mean[c];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) C[3 * r + c] += d[r] * d[c];
      }
      for (int q = 0; q < 9; ++q) C[q] /= (double)nb.size();
      double val[3], vec[3][3];
      symEig3(C, val, vec);
      nv[0] = vec[0][0]; nv[1] = vec[1][0]; nv[2] = vec[2][0];
      if (origin) {
        const double dot = (origin[0] - p[0]) * nv[0] + (origin[1] - p[1]) * nv[1] + (origin[2] - p[2]) * nv[2];
        if (dot < 0) for (int c = 0; c < 3; ++c) nv[c] = -nv[c];
      }
    }
    for (int c = 0; c < 3; ++c) normals[3 * i + c] = nv[c];
  }
  if (degenerate_count) *degenerate_count = degenerate;
  return ORC_OK;
}

// ================================================================================================================
// open3d_slam side (fp64): croppers, transform, map voxelisation, Open3D VoxelDownSample, Open3D -> pointmatcher
// citations relative to open3d_slam_rsl/open3d_slam/open3d_slam/
// ================================================================================================================
namespace {

// CroppingVolume::isWithinVolume (src/croppers.cpp:53-55,121-167): only pose_.translation() is used; inclusive tests;
// the cylinder's z range is absolute.
inline bool withinVolume(const orc_crop_spec& s, const double* p) {
  const double dx = p[0] - s.center[0], dy = p[1] - s.center[1], dz = p[2] - s.center[2];
  bool in;
  switch (s.kind) {
    case ORC_CROP_MAX_RADIUS: in = std::sqrt(dx * dx + dy * dy + dz * dz) <= s.rmax; break;
    case ORC_CROP_MIN_RADIUS: in = std::sqrt(dx * dx + dy * dy + dz * dz) >= s.rmin; break;
    case ORC_CROP_MINMAX_RADIUS: {
      const double d = std::sqrt(dx * dx + dy * dy + dz * dz);
      in = d <= s.rmax && d >= s.rmin;
      break;
    }
    default: in = p[2] >= s.zmin && p[2] <= s.zmax && std::sqrt(dx * dx + dy * dy) <= s.rmax; break;
  }
  return s.invert ? !in : in;
}

struct Acc {   // AccumulatedPoint (src/helpers.cpp:30-73)
  int n = 0;
  double p[3] = {0, 0, 0}, nn[3] = {0, 0, 0};
};
typedef std::tuple<int, int, int> Key;   // ordered (x, y, z) — our deterministic emission order

}  // namespace

extern "C" {

// CroppingVolume::crop (src/croppers.cpp:76-106): order-preserving compaction of points (+ normals)
int64_t orc_crop(const orc_crop_spec* spec, const double* xyz, const double* nrm, int64_t n, double* xyz_out, double* nrm_out,
                 uint8_t* mask_out) {
  int64_t o = 0;
  for (int64_t i = 0; i < n; ++i) {
    const bool in = withinVolume(*spec, xyz + 3 * i);
    if (mask_out) mask_out[i] = in ? 1 : 0;
    if (in) {
      if (xyz_out) for (int d = 0; d < 3; ++d) xyz_out[3 * o + d] = xyz[3 * i + d];
      if (nrm_out && nrm) for (int d = 0; d < 3; ++d) nrm_out[3 * o + d] = nrm[3 * i + d];
      ++o;
    }
  }
  return o;
}

// o3d_slam::transform (src/helpers.cpp:283-318): p' = (T [p;1]).head3 / w ; n' = (T [n;0]).head3.
// The reference's duplicate-on-identity quirk (:285-288) is NOT reproduced.
void orc_transform_f64(const double T[16], const double* xyz, const double* nrm, int64_t n, double* xyz_out, double* nrm_out) {
  for (int64_t i = 0; i < n; ++i) {
    const double* p = xyz + 3 * i;
    double q[4];
    for (int r = 0; r < 4; ++r) q[r] = ((T[4 * r + 0] * p[0] + T[4 * r + 1] * p[1]) + T[4 * r + 2] * p[2]) + T[4 * r + 3] * 1.0;
    for (int r = 0; r < 3; ++r) xyz_out[3 * i + r] = q[r] / q[3];
    if (nrm && nrm_out) {
      const double* a = nrm + 3 * i;
      for (int r = 0; r < 3; ++r) nrm_out[3 * i + r] = ((T[4 * r + 0] * a[0] + T[4 * r + 1] * a[1]) + T[4 * r + 2] * a[2]) + T[4 * r + 3] * 0.0;
    }
  }
}

// voxelizeWithinCroppingVolume (src/helpers.cpp:117-192) with key = int(floor(p * (1/v))) (VoxelHashMap.hpp:43-51).
// Output: pass-through points (outside the volume) in input order, then one point per voxel: mean position,
// normalised mean normal (NaN normals skipped in the sum, :35-37), voxels in ascending (x,y,z) key order.
int64_t orc_voxelize_within_volume(double voxel, const orc_crop_spec* spec, const double* xyz, const double* nrm, int64_t n,
                                   double* xyz_out, double* nrm_out, int64_t* n_passthrough) {
  const double inv = 1.0 / voxel;
  std::map<Key, Acc> acc;
  int64_t o = 0;
  for (int64_t i = 0; i < n; ++i) {
    const double* p = xyz + 3 * i;
    if (withinVolume(*spec, p)) {
      const Key key((int)std::floor(p[0] * inv), (int)std::floor(p[1] * inv), (int)std::floor(p[2] * inv));
      Acc& a = acc[key];
      for (int d = 0; d < 3; ++d) a.p[d] += p[d];
      if (nrm) {
        const double* q = nrm + 3 * i;
        if (!std::isnan(q[0]) && !std::isnan(q[1]) && !std::isnan(q[2])) for (int d = 0; d < 3; ++d) a.nn[d] += q[d];
      }
      a.n++;
    } else {
      for (int d = 0; d < 3; ++d) xyz_out[3 * o + d] = p[d];
      if (nrm && nrm_out) for (int d = 0; d < 3; ++d) nrm_out[3 * o + d] = nrm[3 * i + d];
      ++o;
    }
  }
  if (n_passthrough) *n_passthrough = o;
  for (auto& kv : acc) {
    const Acc& a = kv.second;
    for (int d = 0; d < 3; ++d) xyz_out[3 * o + d] = a.p[d] / (double)a.n;
    if (nrm && nrm_out) {
      double m[3];
      for (int d = 0; d < 3; ++d) m[d] = a.nn[d] / (double)a.n;
      const double nrm2 = std::sqrt(m[0] * m[0] + m[1] * m[1] + m[2] * m[2]);   // Eigen normalized(): divides if norm > 0
      for (int d = 0; d < 3; ++d) nrm_out[3 * o + d] = (nrm2 > 0.0) ? m[d] / nrm2 : m[d];
    }
    ++o;
  }
  return o;
}

// Open3D v0.15.1 PointCloud::VoxelDownSample as called at src/helpers.cpp:108-115 (Open3D is an un-vendored dependency,
// pinned by open3d_catkin/CMakeLists.txt:106-107): voxel_min_bound = min_bound - voxel/2, key = floor((p - min)/voxel),
// mean position, mean normal (not re-normalised).  Emission order: ascending key (Open3D's is hash order — unpinned).
int64_t orc_voxel_down_sample(double voxel, const double* xyz, const double* nrm, int64_t n, double* xyz_out, double* nrm_out) {
  if (n == 0) return 0;
  double mn[3] = {xyz[0], xyz[1], xyz[2]};
  for (int64_t i = 1; i < n; ++i) for (int d = 0; d < 3; ++d) mn[d] = std::min(mn[d], xyz[3 * i + d]);
  for (int d = 0; d < 3; ++d) mn[d] = mn[d] - voxel * 0.5;
  std::map<Key, Acc> acc;
  for (int64_t i = 0; i < n; ++i) {
    const double* p = xyz + 3 * i;
    const Key key((int)std::floor((p[0] - mn[0]) / voxel), (int)std::floor((p[1] - mn[1]) / voxel), (int)std::floor((p[2] - mn[2]) / voxel));
    Acc& a = acc[key];
    for (int d = 0; d < 3; ++d) a.p[d] += p[d];
    if (nrm) {
      const double* q = nrm + 3 * i;
      if (!std::isnan(q[0]) && !std::isnan(q[1]) && !std::isnan(q[2])) for (int d = 0; d < 3; ++d) a.nn[d] += q[d];
    }
    a.n++;
  }
  int64_t o = 0;
  for (auto& kv : acc) {
    const Acc& a = kv.second;
    for (int d = 0; d < 3; ++d) xyz_out[3 * o + d] = a.p[d] / (double)a.n;
    if (nrm && nrm_out) for (int d = 0; d < 3; ++d) nrm_out[3 * o + d] = a.nn[d] / (double)a.n;
    ++o;
  }
  return o;
}

// open3dToPointmatcher (open3d_slam_rsl/open3d_utils/open3d_conversions/src/open3d_conversions.cpp:21-118):
// features 4 x N float (x y z pad=1), normals 3 x N float, fp64 -> fp32 round-to-nearest.
void orc_o3d_to_pm(const double* xyz, const double* nrm, int64_t n, float* feat4, float* nrm3) {
  for (int64_t i = 0; i < n; ++i) {
    for (int d = 0; d < 3; ++d) feat4[4 * i + d] = (float)xyz[3 * i + d];
    feat4[4 * i + 3] = 1.f;
    if (nrm && nrm3) for (int d = 0; d < 3; ++d) nrm3[3 * i + d] = (float)nrm[3 * i + d];
  }
}

// Space carving on the map cloud: Submap::carve (src/Submap.cpp:130-144) -> getIdxsOfCarvedPoints (src/helpers.cpp:245-281).
// The map points inside `spec` are binned into a VoxelMap with key = int(floor(p * (1/voxel))) (Voxel.cpp:122-128,
// VoxelHashMap.hpp:48-51,127); every scan ray (scan already in the map frame) is sampled at distance = 0, step, 2 step, ...
// < max(voxel, min(length - truncation, maxRaytracingLength)) with `distance += stepSize` accumulation (:254-277); every map
// point in a sampled voxel is removed if the cloud has no normals or |direction . normal.normalized()| > minDot (:264-269).
// Output: remove_mask[i] = 1 for removed map points (the reference's unordered id set + SelectByIndex(invert) keeps the rest in
// order, helpers.cpp:230-236).  Returns the number of removed points.
int64_t orc_carve(const double* scan_xyz, int64_t n_scan, const double sensor[3], const double* map_xyz, const double* map_nrm,
                  int64_t n_map, const orc_crop_spec* spec, double voxel, double max_ray_length, double truncation, double min_dot,
                  uint8_t* remove_mask) {
  const double inv = 1.0 / voxel;
  std::map<Key, std::vector<int64_t>> vox;
  for (int64_t i = 0; i < n_map; ++i) {
    remove_mask[i] = 0;
    const double* p = map_xyz + 3 * i;
    if (!withinVolume(*spec, p)) continue;   // cropper.getIndicesWithinVolume(*map) (Submap.cpp:138)
    vox[Key((int)std::floor(p[0] * inv), (int)std::floor(p[1] * inv), (int)std::floor(p[2] * inv))].push_back(i);
  }
  for (int64_t i = 0; i < n_scan; ++i) {
    const double* p = scan_xyz + 3 * i;
    const double d0 = p[0] - sensor[0], d1 = p[1] - sensor[1], d2 = p[2] - sensor[2];
    const double length = std::sqrt((d0 * d0 + d1 * d1) + d2 * d2);
    const double dir[3] = {d0 / length, d1 / length, d2 / length};
    const double maxPath = std::max(voxel, std::min(length - truncation, max_ray_length));
    double distance = 0.0;
    while (distance < maxPath) {
      const double c[3] = {distance * dir[0] + sensor[0], distance * dir[1] + sensor[1], distance * dir[2] + sensor[2]};
      const auto it = vox.find(Key((int)std::floor(c[0] * inv), (int)std::floor(c[1] * inv), (int)std::floor(c[2] * inv)));
      if (it != vox.end()) {
        for (const int64_t id : it->second) {
          bool rm = true;
          if (map_nrm) {
            const double* a = map_nrm + 3 * id;
            const double n2 = (a[0] * a[0] + a[1] * a[1]) + a[2] * a[2];   // Eigen normalized(): v / sqrt(squaredNorm) when > 0
            double nx = a[0], ny = a[1], nz = a[2];
            if (n2 > 0.0) { const double nn = std::sqrt(n2); nx /= nn; ny /= nn; nz /= nn; }
            rm = std::fabs((dir[0] * nx + dir[1] * ny) + dir[2] * nz) > min_dot;
          }
          if (rm) remove_mask[id] = 1;
        }
      }
      distance += voxel;
    }
  }
  int64_t removed = 0;
  for (int64_t i = 0; i < n_map; ++i) removed += remove_mask[i];
  return removed;
}


// sensor_msgs/PointCloud2 -> clouds.  RosPointCloud2Deserializer<float>::deserialize (pointmatcher_ros/src/
// RosPointCloud2Deserializer.cpp:9-38,113-143,213-232) and open3d_conversions::rosToOpen3d (open3d_conversions.cpp:203-244) both
// walk the message with sensor_msgs::PointCloud2ConstIterator<float>: element i of a field = the 4 bytes at
// data + i * point_step + offset read as a float (no datatype conversion, row_step unused).  offs = {x, y, z, normal_x, normal_y,
// normal_z}; the normals are taken when all three offsets are >= 0.  feat4 (nullable): x y z 1; xyz64 (nullable): doubles.
int32_t orc_pointcloud2(const uint8_t* data, int64_t n, uint32_t point_step, const int32_t offs[6], float* feat4, float* nrm3,
                        double* xyz64, double* nrm64) {
  const bool hasN = offs[3] >= 0 && offs[4] >= 0 && offs[5] >= 0;
  if (offs[0] < 0 || offs[1] < 0 || offs[2] < 0) return ORC_E_FIELD;
  for (int64_t i = 0; i < n; ++i) {
    const uint8_t* p = data + (size_t)i * point_step;
    float v[6] = {0, 0, 0, 0, 0, 0};
    for (int k = 0; k < (hasN ? 6 : 3); ++k) std::memcpy(&v[k], p + offs[k], 4);
    if (feat4) { feat4[4 * i] = v[0]; feat4[4 * i + 1] = v[1]; feat4[4 * i + 2] = v[2]; feat4[4 * i + 3] = 1.f; }
    if (xyz64) for (int d = 0; d < 3; ++d) xyz64[3 * i + d] = (double)v[d];
    if (hasN && nrm3) for (int d = 0; d < 3; ++d) nrm3[3 * i + d] = v[3 + d];
    if (hasN && nrm64) for (int d = 0; d < 3; ++d) nrm64[3 * i + d] = (double)v[3 + d];
  }
  return ORC_OK;
}

// computeIndicesOfOverlappingPoints (open3d_slam/src/helpers.cpp:320-345): masks of the source / target points whose voxel
// (key = int(floor(p * (1/voxel))), source moved by Open3D's PointCloud::Transform first) holds >= min_pts points of BOTH clouds.
void orc_overlap(const double* src, int64_t n_src, const double* tgt, int64_t n_tgt, const double T[16], double voxel, int64_t min_pts,
                 uint8_t* src_mask, uint8_t* tgt_mask) {
  const double inv = 1.0 / voxel;
  std::map<Key, std::pair<std::vector<int64_t>, std::vector<int64_t>>> vox;
  for (int64_t i = 0; i < n_tgt; ++i) {
    const double* p = tgt + 3 * i;
    vox[Key((int)std::floor(p[0] * inv), (int)std::floor(p[1] * inv), (int)std::floor(p[2] * inv))].second.push_back(i);
    tgt_mask[i] = 0;
  }
  for (int64_t i = 0; i < n_src; ++i) {
    const double* p = src + 3 * i;
    double q[4];
    for (int r = 0; r < 4; ++r) q[r] = ((T[4 * r + 0] * p[0] + T[4 * r + 1] * p[1]) + T[4 * r + 2] * p[2]) + T[4 * r + 3] * 1.0;
    const double x = q[0] / q[3], y = q[1] / q[3], z = q[2] / q[3];
    vox[Key((int)std::floor(x * inv), (int)std::floor(y * inv), (int)std::floor(z * inv))].first.push_back(i);
    src_mask[i] = 0;
  }
  for (const auto& kv : vox) {
    if ((int64_t)kv.second.first.size() >= min_pts && (int64_t)kv.second.second.size() >= min_pts) {
      for (int64_t i : kv.second.first) src_mask[i] = 1;
      for (int64_t i : kv.second.second) tgt_mask[i] = 1;
    }
  }
}

// GetInformationMatrixFromPointClouds (Open3D v0.15.1 pipelines/registration/Registration.cpp, external): source moved by T,
// nearest target point within max_dist (d2 < r2, fp64 brute force here), GTG += G_r^T G_r for the three rows
// [0 z -y 1 0 0], [-z 0 x 0 1 0], [y -x 0 0 0 1] of the matched TARGET point.  Returns the number of correspondences.
int64_t orc_information_matrix(const double* src, int64_t n_src, const double* tgt, int64_t n_tgt, double max_dist, const double T[16],
                               double info[36]) {
  for (int i = 0; i < 36; ++i) info[i] = 0.0;
  std::vector<int64_t> match((size_t)n_src, -1);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n_src; ++i) {
    const double* p = src + 3 * i;
    double q[4];
    for (int r = 0; r < 4; ++r) q[r] = ((T[4 * r + 0] * p[0] + T[4 * r + 1] * p[1]) + T[4 * r + 2] * p[2]) + T[4 * r + 3] * 1.0;
    const double x = q[0] / q[3], y = q[1] / q[3], z = q[2] / q[3];
    double best = max_dist * max_dist;
    for (int64_t j = 0; j < n_tgt; ++j) {
      const double dx = x - tgt[3 * j], dy = y - tgt[3 * j + 1], dz = z - tgt[3 * j + 2];
      const double d2 = (dx * dx + dy * dy) + dz * dz;
      if (d2 < best) { best = d2; match[(size_t)i] = j; }
    }
  }
  int64_t count = 0;
  for (int64_t i = 0; i < n_src; ++i) {
    if (match[(size_t)i] < 0) continue;
    ++count;
    const double* t = tgt + 3 * match[(size_t)i];
    const double rows[3][6] = {{0, t[2], -t[1], 1, 0, 0}, {-t[2], 0, t[0], 0, 1, 0}, {t[1], -t[0], 0, 0, 0, 1}};
    for (int r = 0; r < 3; ++r)
      for (int a = 0; a < 6; ++a)
        for (int b = 0; b < 6; ++b) info[6 * a + b] += rows[r][a] * rows[r][b];
  }
  return count;
}

}  // extern "C"

// ================================================================================================================
// VarTrimmedDistOutlierFilter, PointToPointErrorMinimizer, PointToPlaneWithCov covariance and the
// compacting DataPointsFilters
// ================================================================================================================
#include <random>

// VarTrimmedDistOutlierFilter::optimizeInlierRatio (OutlierFiltersImpl.cpp:178-220), then compute (:166-175).
// The window [minEl, maxEl) is cut at the number of valid distances: beyond it the reference reads vector capacity that
// std::partial_sum never wrote (undefined behaviour).  pow(ratio, lambda) is the fp64 pow rounded once to float.
extern "C" int32_t orc_var_trimmed_dist(const float* d2, int64_t count, float min_ratio, float max_ratio, float lambda, float* ratio_out,
                                        float* tau_out, float* w_out) {
  if (!(min_ratio < max_ratio)) return ORC_E_PARAM;     // :160-163
  std::vector<float> sorted;
  sorted.reserve((size_t)count);
  for (int64_t i = 0; i < count; ++i)
    if (d2[i] != kInf && d2[i] > 0.f) sorted.push_back(d2[i]);   // :189-192
  if (sorted.empty()) return ORC_E_CONVERGENCE;                   // :193-194
  std::sort(sorted.begin(), sorted.end());
  std::vector<float> cum(sorted.size());
  {
    volatile float run = 0.f;   // std::partial_sum on float: one rounded add per element, in order
    for (size_t i = 0; i < sorted.size(); ++i) { run = run + sorted[i]; cum[i] = run; }
  }
  const int points_nbr = (int)count;
  const int minEl = (int)std::floor(min_ratio * points_nbr), maxEl = (int)std::floor(max_ratio * points_nbr);   // :200-201
  const int end = std::min<int64_t>(maxEl, (int64_t)sorted.size());
  int minIndex = -1;
  float best = kInf;
  for (int e = std::max(minEl, 0); e < end; ++e) {
    const float id = (float)(e + 1);                                   // LinSpaced(.., minEl + 1, maxEl)
    volatile float ratio = id / (float)points_nbr;                     // :208
    volatile float deno = (float)std::pow((double)ratio, (double)lambda);   // :209
    volatile float invId = 1.f / id, invD = 1.f / deno;
    volatile float a = cum[e] * invId, b = invD * invD;
    const float frms = a * b;                                          // :211
    if (frms < best) { best = frms; minIndex = e - minEl; }            // minCoeff: first minimum
  }
  if (minIndex < 0) return ORC_E_CONVERGENCE;
  const float ratio = (float)(minIndex + minEl) / (float)points_nbr;   // :214
  if (ratio_out) *ratio_out = ratio;
  if (!tau_out && !w_out) return ORC_OK;
  return orc_trimmed_dist(d2, count, ratio, tau_out, w_out);           // :172-173
}

namespace {

// 3x3 SVD through the symmetric eigen-decomposition of M^T M (fp64): M = U diag(s) V^T, s descending
void svd3ViaEig(const double M[9], double U[3][3], double s[3], double V[3][3]) {
  double MtM[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) MtM[3 * i + j] = M[i] * M[j] + M[3 + i] * M[3 + j] + M[6 + i] * M[6 + j];
  double val[3], vec[3][3];
  symEig3(MtM, val, vec);   // ascending
  for (int c = 0; c < 3; ++c) {
    const int src = 2 - c;
    s[c] = std::sqrt(std::max(val[src], 0.0));
    for (int r = 0; r < 3; ++r) V[r][c] = vec[r][src];
  }
  for (int c = 0; c < 2; ++c) {
    double u[3];
    for (int r = 0; r < 3; ++r) u[r] = M[3 * r] * V[0][c] + M[3 * r + 1] * V[1][c] + M[3 * r + 2] * V[2][c];
    if (c == 1) {   // re-orthogonalise against the first column
      const double d = u[0] * U[0][0] + u[1] * U[1][0] + u[2] * U[2][0];
      for (int r = 0; r < 3; ++r) u[r] -= d * U[r][0];
    }
    const double nn = std::sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
    for (int r = 0; r < 3; ++r) U[r][c] = u[r] / nn;
  }
  // third column: M v3 / s3 when s3 is well above round-off, else the cross product (sign fixed later by the reflection test)
  double u3[3] = {U[1][0] * U[2][1] - U[2][0] * U[1][1], U[2][0] * U[0][1] - U[0][0] * U[2][1], U[0][0] * U[1][1] - U[1][0] * U[0][1]};
  double mv[3];
  for (int r = 0; r < 3; ++r) mv[r] = M[3 * r] * V[0][2] + M[3 * r + 1] * V[1][2] + M[3 * r + 2] * V[2][2];
  const double sgn = (mv[0] * u3[0] + mv[1] * u3[1] + mv[2] * u3[2]) < 0.0 ? -1.0 : 1.0;
  for (int r = 0; r < 3; ++r) U[r][2] = sgn * u3[r];
}

template <typename S>
int pointToPoint(const float* readFeat, int rstride, int64_t n, const float* refFeat, int fstride, const int32_t* ids, const float* d2,
                 const float* w, int k, float T_out[16], int* kept_out) {
  std::vector<Pair> pairs;
  for (int64_t i = 0; i < n; ++i)
    for (int j = 0; j < k; ++j) {
      const int64_t o = (int64_t)k * i + j;
      if (d2[o] == kInf) continue;                                   // ErrorMinimizer.cpp:103-106
      if (w[o] != 0.f) pairs.push_back({i, ids[o], w[o]});           // :108
    }
  if (kept_out) *kept_out = (int)pairs.size();
  if (pairs.empty()) return ORC_E_CONVERGENCE;
  S wsum = 0, mp[3] = {0, 0, 0}, mq[3] = {0, 0, 0};
  for (const Pair& p : pairs) {                                      // PointToPoint.cpp:62-67
    wsum += (S)p.w;
    for (int d = 0; d < 3; ++d) {
      mp[d] += (S)readFeat[p.read * rstride + d] * (S)p.w;
      mq[d] += (S)refFeat[(int64_t)p.ref * fstride + d] * (S)p.w;
    }
  }
  const S inv = S(1) / wsum;
  for (int d = 0; d < 3; ++d) { mp[d] *= inv; mq[d] *= inv; }
  S C[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  for (const Pair& p : pairs)                                        // :70-77  m = ref_c * diag(w) * read_c^T
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        C[3 * i + j] += ((S)refFeat[(int64_t)p.ref * fstride + i] - mq[i]) * (S)p.w * ((S)readFeat[p.read * rstride + j] - mp[j]);
  double M[9], U[3][3], s[3], V[3][3], R[3][3];
  for (int i = 0; i < 9; ++i) M[i] = (double)C[i];
  svd3ViaEig(M, U, s, V);
  auto mul = [&]() {
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) R[i][j] = U[i][0] * V[j][0] + U[i][1] * V[j][1] + U[i][2] * V[j][2];
  };
  mul();
  const double det = R[0][0] * (R[1][1] * R[2][2] - R[1][2] * R[2][1]) - R[0][1] * (R[1][0] * R[2][2] - R[1][2] * R[2][0]) +
                     R[0][2] * (R[1][0] * R[2][1] - R[1][1] * R[2][0]);
  if (det < 0.0) {                                                   // :80-85
    for (int r = 0; r < 3; ++r) V[r][2] = -V[r][2];
    mul();
  }
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) T_out[4 * i + j] = (float)R[i][j];
    T_out[4 * i + 3] = (float)((double)mq[i] - (R[i][0] * (double)mp[0] + R[i][1] * (double)mp[1] + R[i][2] * (double)mp[2]));   // :86
  }
  T_out[12] = T_out[13] = T_out[14] = 0.f;
  T_out[15] = 1.f;
  return ORC_OK;
}

}  // namespace

extern "C" int32_t orc_point_to_point(const float* read_feat, int32_t rstride, int64_t n, const float* ref_feat, int32_t fstride,
                                      const int32_t* ids, const float* d2, const float* w, int32_t k, int32_t precision, float T_out[16],
                                      int32_t* kept_out) {
  int kept = 0;
  const int st = precision == ORC_REF32 ? pointToPoint<float>(read_feat, rstride, n, ref_feat, fstride, ids, d2, w, k, T_out, &kept)
                                        : pointToPoint<double>(read_feat, rstride, n, ref_feat, fstride, ids, d2, w, k, T_out, &kept);
  if (kept_out) *kept_out = kept;
  return st;
}

namespace {

// PointToPlaneWithCovErrorMinimizer::estimateCovariance (PointToPlaneWithCov.cpp:70-160) over the ErrorElements as
// PointToPlaneErrorMinimizer::compute leaves them: mean-centred in place (PointToPlane.cpp:263-268)
template <typename S>
int p2planeCovariance(const float* readFeat, int rstride, int64_t n, const float* refFeat, int fstride, const float* refNormals,
                      const int32_t* ids, const float* d2, const float* w, int k, const float T[16], float sensorStdDev, double cov[36]) {
  if (!refNormals) {                                                 // :91-92
    for (int i = 0; i < 36; ++i) cov[i] = (i % 7 == 0) ? (double)std::numeric_limits<float>::max() : 0.0;
    return ORC_OK;
  }
  std::vector<Pair> pairs;
  for (int64_t i = 0; i < n; ++i)
    for (int j = 0; j < k; ++j) {
      const int64_t o = (int64_t)k * i + j;
      if (d2[o] == kInf) continue;
      if (w[o] != 0.f) pairs.push_back({i, ids[o], w[o]});
    }
  if (pairs.empty()) return ORC_E_CONVERGENCE;
  const S K = (S)pairs.size();
  S mp[3] = {0, 0, 0}, mq[3] = {0, 0, 0};
  for (const Pair& p : pairs)
    for (int d = 0; d < 3; ++d) { mp[d] += (S)readFeat[p.read * rstride + d]; mq[d] += (S)refFeat[(int64_t)p.ref * fstride + d]; }
  for (int d = 0; d < 3; ++d) { mp[d] /= K; mq[d] /= K; }
  const S beta = -std::asin((S)T[8]);                                // :94-99
  const S alpha = std::atan2((S)T[9], (S)T[10]);
  const S gamma = std::atan2((S)T[4] / std::cos(beta), (S)T[0] / std::cos(beta));
  const S tx = (S)T[3], ty = (S)T[7], tz = (S)T[11];
  std::vector<S> H(36, S(0)), C(36, S(0));
  for (const Pair& p : pairs) {
    S r[3], q[3], nn[3];
    for (int d = 0; d < 3; ++d) {
      r[d] = (S)readFeat[p.read * rstride + d] - mp[d];
      q[d] = (S)refFeat[(int64_t)p.ref * fstride + d] - mq[d];
      nn[d] = (S)refNormals[3 * (int64_t)p.ref + d];
    }
    const S rr = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]), fr = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);   // :113-116
    const S rd[3] = {r[0] / rr, r[1] / rr, r[2] / rr}, fd[3] = {q[0] / fr, q[1] / fr, q[2] / fr};
    const S na = nn[2] * rd[1] - nn[1] * rd[2], nb = nn[0] * rd[2] - nn[2] * rd[0], ng = nn[1] * rd[0] - nn[0] * rd[1];         // :118-120
    S E = nn[0] * (r[0] - gamma * r[1] + beta * r[2] + tx - q[0]);                                                            // :122-124
    E += nn[1] * (gamma * r[0] + r[1] - alpha * r[2] + ty - q[1]);
    E += nn[2] * (-beta * r[0] + alpha * r[1] + r[2] + tz - q[2]);
    S Nr = nn[0] * (rd[0] - gamma * rd[1] + beta * rd[2]);                                                                    // :126-128
    Nr += nn[1] * (gamma * rd[0] + rd[1] - alpha * rd[2]);
    Nr += nn[2] * (-beta * rd[0] + alpha * rd[1] + rd[2]);
    const S Nf = -(nn[0] * fd[0] + nn[1] * fd[1] + nn[2] * fd[2]);                                                            // :130
    const S v[6] = {nn[0], nn[1], nn[2], rr * na, rr * nb, rr * ng};                                                          // :133
    const S g = E + rr * Nr;
    const S a[6] = {nn[0] * Nr, nn[1] * Nr, nn[2] * Nr, na * g, nb * g, ng * g};                                              // :137
    const S b[6] = {nn[0] * Nf, nn[1] * Nf, nn[2] * Nf, fr * na * Nf, fr * nb * Nf, fr * ng * Nf};                            // :141
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 6; ++j) { H[6 * i + j] += v[i] * v[j]; C[6 * i + j] += a[i] * a[j] + b[i] * b[j]; }                 // :135, :148-150
  }
  // inverse of J_hessian by solving H X = I column by column (fp64 LLT-free Gauss with partial pivoting)
  double A[6][12];
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) { A[i][j] = (double)H[6 * i + j]; A[i][6 + j] = i == j ? 1.0 : 0.0; }
  for (int c = 0; c < 6; ++c) {
    int piv = c;
    for (int r2 = c + 1; r2 < 6; ++r2) if (std::fabs(A[r2][c]) > std::fabs(A[piv][c])) piv = r2;
    if (!(std::fabs(A[piv][c]) > 0.0)) { for (int i = 0; i < 36; ++i) cov[i] = std::nan(""); return ORC_OK; }
    for (int j = 0; j < 12; ++j) std::swap(A[c][j], A[piv][j]);
    const double pv = A[c][c];
    for (int j = 0; j < 12; ++j) A[c][j] /= pv;
    for (int r2 = 0; r2 < 6; ++r2) {
      if (r2 == c) continue;
      const double f = A[r2][c];
      for (int j = 0; j < 12; ++j) A[r2][j] -= f * A[c][j];
    }
  }
  double t1[36];
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) { double s = 0; for (int l = 0; l < 6; ++l) s += A[i][6 + l] * (double)C[6 * l + j]; t1[6 * i + j] = s; }
  const double s2 = (double)sensorStdDev * (double)sensorStdDev;     // :156
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) { double s = 0; for (int l = 0; l < 6; ++l) s += t1[6 * i + l] * A[l][6 + j]; cov[6 * i + j] = s2 * s; }
  return ORC_OK;
}

}  // namespace

extern "C" int32_t orc_point_to_plane_covariance(const float* read_feat, int32_t rstride, int64_t n, const float* ref_feat, int32_t fstride,
                                                 const float* ref_normals, const int32_t* ids, const float* d2, const float* w, int32_t k,
                                                 int32_t precision, const float T[16], float sensor_std_dev, double cov_out[36]) {
  return precision == ORC_REF32
             ? p2planeCovariance<float>(read_feat, rstride, n, ref_feat, fstride, ref_normals, ids, d2, w, k, T, sensor_std_dev, cov_out)
             : p2planeCovariance<double>(read_feat, rstride, n, ref_feat, fstride, ref_normals, ids, d2, w, k, T, sensor_std_dev, cov_out);
}

// MaxDist.cpp:62-105, MinDist.cpp:62-101, BoundingBox.cpp:66-95, RandomSampling.cpp:60-100 on a 3-D cloud: keep[i] in {0, 1},
// returns the number kept.  kind / fields as b200reg_filter_params (include/b200reg.h).
extern "C" int64_t orc_cloud_filter(const float* feat, int32_t stride, int64_t n, int32_t kind, int32_t dim, float dist, const float box[6],
                                    int32_t remove_inside, int32_t random_method, double prob, uint64_t seed, uint8_t* keep) {
  int64_t j = 0;
  if (kind == 3) {
    std::minstd_rand gen((std::minstd_rand::result_type)seed);       // RandomSampling.cpp:64-65 with a caller-supplied seed
    std::vector<float> rnd((size_t)n);
    if (random_method == 1) {
      std::uniform_real_distribution<float> distribution(0, 1);      // :76-77
      for (int64_t i = 0; i < n; ++i) rnd[i] = distribution(gen);
    } else {
      const float range = static_cast<float>(gen.max() - gen.min()); // :71-72
      for (int64_t i = 0; i < n; ++i) rnd[i] = static_cast<float>(gen() / range);
    }
    const size_t nbPointsOut = (size_t)((double)(size_t)n * prob);   // :88
    for (int64_t i = 0; i < n; ++i) keep[i] = 0;
    for (int64_t i = 0; i < n && (size_t)j <= nbPointsOut; ++i)      // :92
      if (rnd[i] < prob) { keep[i] = 1; ++j; }
    return j;
  }
  for (int64_t i = 0; i < n; ++i) {
    const float* p = feat + i * stride;
    bool k2;
    if (kind == 2) {
      const bool in = (p[0] > box[0] && p[0] < box[1]) && (p[1] > box[2] && p[1] < box[3]) && (p[2] > box[4] && p[2] < box[5]);
      k2 = remove_inside ? !in : in;
    } else {
      float v, lim = dist;
      if (dim == -1) {
        volatile float xx = p[0] * p[0], yy = p[1] * p[1], zz = p[2] * p[2];
        volatile float s = xx + yy;
        s = s + zz;
        v = std::sqrt((float)s);
        lim = std::fabs(lim);
      } else {
        v = p[dim];
      }
      k2 = kind == 0 ? (v < lim) : (v > lim);
    }
    keep[i] = k2 ? 1 : 0;
    j += k2;
  }
  return j;
}
