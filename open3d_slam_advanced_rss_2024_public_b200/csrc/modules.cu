// csrc/modules.cu — the remaining cheap chain modules of SURVEY 8f rank 3, on the device so that a YAML chain that names them
// does not force a download: VarTrimmedDistOutlierFilter (OutlierFiltersImpl.cpp:152-220), PointToPointErrorMinimizer
// (ErrorMinimizers/PointToPoint.cpp:55-94), the covariance of PointToPlaneWithCovErrorMinimizer (PointToPlaneWithCov.cpp:70-160)
// and the compacting DataPointsFilters MaxDist / MinDist / BoundingBox / RandomSampling (DataPointsFilters/*.cpp).
#include <cfloat>
#include <cmath>
#include <cstring>
#include <mutex>

#include <cub/cub.cuh>

#include "common.cuh"

namespace {

// ---------------------------------------------------------------------------------------------------------------------------
// VarTrimmedDistOutlierFilter::optimizeInlierRatio (OutlierFiltersImpl.cpp:178-220)
//   sorted  = ascending squared distances that are finite and > 0
//   cum[i]  = fp32 running sum of sorted[0..i], added in that order (std::partial_sum on float)
//   FRMS[j] = cum[minEl + j] * (1 / id) * (1 / ratio^lambda)^2,  id = minEl + 1 + j (as float), ratio = id / points_nbr
//   optRatio = (first argmin + minEl) / points_nbr,  points_nbr = ALL entries of the distance matrix (invalid ones included)
// The running sum is inherently sequential in fp32 (the rounding of every add depends on the sum so far), so one warp walks
// the sorted array: 32 lanes stage 1024 values in shared memory, lane 0 adds them in order, then all lanes evaluate FRMS for
// that chunk.  pow is taken in fp64 and rounded once to fp32 (= the correctly rounded powf; oracle.cpp does the same).
// Entries of the window [minEl, maxEl) beyond the last valid distance are skipped: the reference reads reserved-but-unwritten
// vector storage there (undefined behaviour), nothing to restate.
// ---------------------------------------------------------------------------------------------------------------------------
__global__ void vtKeyKernel(const float* __restrict__ d2, int64_t count, unsigned* __restrict__ keys, unsigned* __restrict__ nValid) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool ok = false;
  if (i < count) {
    const float d = d2[i];
    ok = (d != CUDART_INF_F) && (d > 0.f);   // NaN fails d > 0
    keys[i] = ok ? __float_as_uint(d) : 0xffffffffu;   // positive floats order like their bit patterns
  }
  const unsigned m = __ballot_sync(0xffffffffu, ok);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(nValid, (unsigned)__popc(m));
}

struct VtResult {
  unsigned minIndex;   // first argmin of FRMS (relative to minEl)
  unsigned nValid;
  float frmsMin;
  int empty;           // window holds no valid entry
};

__global__ void __launch_bounds__(32) vtScanKernel(const unsigned* __restrict__ sorted, const unsigned* __restrict__ nValidPtr, int64_t pointsNbr,
                                                   int64_t minEl, int64_t maxEl, float lambda, VtResult* __restrict__ out) {
  constexpr int CH = 1024;
  __shared__ float s[CH];
  const int lane = threadIdx.x;
  const int64_t nValid = *nValidPtr;
  const int64_t end = maxEl < nValid ? maxEl : nValid;
  float run = 0.f;
  float best = CUDART_INF_F;
  long long bestIdx = -1;
  const float nbr = (float)pointsNbr;
  for (int64_t base = 0; base < end; base += CH) {
    const int len = (int)((end - base) < CH ? (end - base) : CH);
    for (int t = lane; t < len; t += 32) s[t] = __uint_as_float(sorted[base + t]);
    __syncwarp();
    if (lane == 0) {
      float r = run;
#pragma unroll 8
      for (int t = 0; t < len; ++t) { r = __fadd_rn(r, s[t]); s[t] = r; }
      run = r;
    }
    __syncwarp();
    run = __shfl_sync(0xffffffffu, run, 0);
    for (int t = lane; t < len; t += 32) {
      const int64_t e = base + t;
      if (e < minEl) continue;
      const float id = (float)(e + 1);                       // LinSpaced(minEl + 1 .. maxEl): exact integers in float up to 2^24
      const float ratio = __fdiv_rn(id, nbr);
      const float deno = (float)pow((double)ratio, (double)lambda);
      const float invD = __fdiv_rn(1.f, deno);
      const float f = __fmul_rn(__fmul_rn(s[t], __fdiv_rn(1.f, id)), __fmul_rn(invD, invD));
      if (f < best) { best = f; bestIdx = e; }               // strict: the first minimum wins inside a lane (ascending e)
    }
    __syncwarp();
  }
  // first argmin over the warp: smaller value, then smaller index
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const long long oi = __shfl_xor_sync(0xffffffffu, bestIdx, o);
    if (oi >= 0 && (bestIdx < 0 || ob < best || (ob == best && oi < bestIdx))) { best = ob; bestIdx = oi; }
  }
  if (lane == 0) {
    out->nValid = (unsigned)nValid;
    out->empty = bestIdx < 0;
    out->minIndex = bestIdx < 0 ? 0u : (unsigned)(bestIdx - minEl);
    out->frmsMin = best;
  }
}

// ---------------------------------------------------------------------------------------------------------------------------
// fp64 moment reductions over matched pairs (the ErrorElements of ErrorMinimizer.cpp:60-150 without materialising them)
// ---------------------------------------------------------------------------------------------------------------------------
struct PairArgs {
  const float4* reading;
  const float4* refPts;
  const float4* refNrm;
  const int* ids;
  const float* d2;
  const float* w;
  int64_t total;
  int k;
  // covariance mode
  double meanRead[3], meanRef[3];
  double alpha, beta, gamma, tx, ty, tz;
};

template <int N>
__device__ __forceinline__ void blockReduceStore(double (&m)[N], double* __restrict__ partials) {
  __shared__ double sh[8][N];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < N; ++j) {
    double v = m[j];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) sh[warp][j] = v;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < N; j += blockDim.x) {
    double v = 0.0;
    for (int w8 = 0; w8 < (int)(blockDim.x >> 5); ++w8) v += sh[w8][j];
    partials[(size_t)blockIdx.x * N + j] = v;
  }
}
// fixed-order sum of the per-block partials (deterministic for a given grid)
__global__ void sumPartialsKernel(const double* __restrict__ partials, int blocks, int n, double* __restrict__ out) {
  const int j = threadIdx.x;
  if (j >= n) return;
  double v = 0.0;
  for (int b = 0; b < blocks; ++b) v += partials[(size_t)b * n + j];
  out[j] = v;
}

__device__ __forceinline__ bool pairOf(const PairArgs& A, int64_t o, double& wgt, int& id) {
  if (A.d2[o] == CUDART_INF_F) return false;   // ErrorMinimizer.cpp:103-106
  const float w = A.w[o];
  if (w == 0.f) return false;                  // ErrorMinimizer.cpp:108
  id = A.ids[o];
  wgt = (double)w;
  return true;
}

// PointToPoint: 0 sum w | 1-3 sum w p | 4-6 sum w q | 7-15 sum w q p^T (row-major) | 16 kept pairs | 17 pairs with a finite distance
constexpr int P2P_N = 18;
__global__ void __launch_bounds__(256) p2pMomentsKernel(PairArgs A, double* __restrict__ partials) {
  double m[P2P_N];
#pragma unroll
  for (int j = 0; j < P2P_N; ++j) m[j] = 0.0;
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < A.total; o += (int64_t)gridDim.x * blockDim.x) {
    double w;
    int id;
    if (A.d2[o] != CUDART_INF_F) m[17] += 1.0;
    if (!pairOf(A, o, w, id)) continue;
    const float4 p = A.reading[o / A.k], q = A.refPts[id];
    const double px = p.x, py = p.y, pz = p.z, qx = q.x, qy = q.y, qz = q.z;
    m[0] += w;
    m[1] += w * px; m[2] += w * py; m[3] += w * pz;
    m[4] += w * qx; m[5] += w * qy; m[6] += w * qz;
    m[7] += w * qx * px; m[8] += w * qx * py; m[9] += w * qx * pz;
    m[10] += w * qy * px; m[11] += w * qy * py; m[12] += w * qy * pz;
    m[13] += w * qz * px; m[14] += w * qz * py; m[15] += w * qz * pz;
    m[16] += 1.0;
  }
  blockReduceStore<P2P_N>(m, partials);
}

// PointToPlaneErrorMinimizer::computeResidualError (PointToPlane.cpp:372-407): 0 sum w ((p - q) . n)^2 | 1 kept pairs.  A.alpha != 0: force2D
constexpr int RES_N = 2;
__global__ void __launch_bounds__(256) residualKernel(PairArgs A, double* __restrict__ partials) {
  double m[RES_N] = {0.0, 0.0};
  const bool twoD = A.alpha != 0.0;
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < A.total; o += (int64_t)gridDim.x * blockDim.x) {
    double w;
    int id;
    if (!pairOf(A, o, w, id)) continue;
    const float4 p = A.reading[o / A.k], q = A.refPts[id], n = A.refNrm[id];
    double dot = ((double)p.x - (double)q.x) * (double)n.x + ((double)p.y - (double)q.y) * (double)n.y;
    if (!twoD) dot += ((double)p.z - (double)q.z) * (double)n.z;
    m[0] += w * (dot * dot);
    m[1] += 1.0;
  }
  blockReduceStore<RES_N>(m, partials);
}

// PointToPlaneWithCov::estimateCovariance (PointToPlaneWithCov.cpp:70-160): J_hessian = sum v v^T and d2J_dZdX d2J_dZdX^T =
// sum (a a^T + b b^T), upper triangles (21 + 21 moments) | 42 pair count.  The pairs are the mean-centred ErrorElements that
// PointToPlaneErrorMinimizer::compute leaves behind (PointToPlane.cpp:263-268 centres mPts in place before the covariance runs).
constexpr int COV_N = 43;
__global__ void __launch_bounds__(256) covMomentsKernel(PairArgs A, double* __restrict__ partials) {
  double m[COV_N];
#pragma unroll
  for (int j = 0; j < COV_N; ++j) m[j] = 0.0;
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < A.total; o += (int64_t)gridDim.x * blockDim.x) {
    double w;
    int id;
    if (!pairOf(A, o, w, id)) continue;
    const float4 p = A.reading[o / A.k], q = A.refPts[id], nn = A.refNrm[id];
    const double r0 = p.x - A.meanRead[0], r1 = p.y - A.meanRead[1], r2 = p.z - A.meanRead[2];
    const double s0 = q.x - A.meanRef[0], s1 = q.y - A.meanRef[1], s2 = q.z - A.meanRef[2];
    const double n0 = nn.x, n1 = nn.y, n2 = nn.z;
    const double rr = sqrt(r0 * r0 + r1 * r1 + r2 * r2), fr = sqrt(s0 * s0 + s1 * s1 + s2 * s2);
    const double d0 = r0 / rr, d1 = r1 / rr, d2 = r2 / rr;
    const double f0 = s0 / fr, f1 = s1 / fr, f2 = s2 / fr;
    const double na = n2 * d1 - n1 * d2, nb = n0 * d2 - n2 * d0, ng = n1 * d0 - n0 * d1;
    const double al = A.alpha, be = A.beta, ga = A.gamma;
    const double E = n0 * (r0 - ga * r1 + be * r2 + A.tx - s0) + n1 * (ga * r0 + r1 - al * r2 + A.ty - s1) +
                     n2 * (-be * r0 + al * r1 + r2 + A.tz - s2);
    const double Nr = n0 * (d0 - ga * d1 + be * d2) + n1 * (ga * d0 + d1 - al * d2) + n2 * (-be * d0 + al * d1 + d2);
    const double Nf = -(n0 * f0 + n1 * f1 + n2 * f2);
    const double v[6] = {n0, n1, n2, rr * na, rr * nb, rr * ng};
    const double g = E + rr * Nr;
    const double a[6] = {n0 * Nr, n1 * Nr, n2 * Nr, na * g, nb * g, ng * g};
    const double b[6] = {n0 * Nf, n1 * Nf, n2 * Nf, fr * na * Nf, fr * nb * Nf, fr * ng * Nf};
    int t = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = i; j < 6; ++j, ++t) {
        m[t] += v[i] * v[j];
        m[21 + t] += a[i] * a[j] + b[i] * b[j];
      }
    m[42] += 1.0;
  }
  blockReduceStore<COV_N>(m, partials);
}

template <int N, typename K>
int reducePairs(b200reg_ctx* ctx, K kernel, const PairArgs& A, double* hostOut) {
  const int blocks = (int)fmax(1.0, fmin((double)divUp(A.total, 1024), (double)(2 * ctx->sm_count)));
  TRY(ensureBuf(ctx, ctx->partials, sizeof(double) * N * (size_t)(blocks + 1)));
  double* part = (double*)ctx->partials.p;
  double* res = part + (size_t)blocks * N;
  kernel<<<blocks, 256, 0, ctx->stream>>>(A, part);
  LAUNCH_CHECK(ctx);
  sumPartialsKernel<<<1, 64, 0, ctx->stream>>>(part, blocks, N, res);
  LAUNCH_CHECK(ctx);
  double* h = (double*)ctx->pinned;
  CUDA_TRY(ctx, cudaMemcpyAsync(h, res, sizeof(double) * N, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(hostOut, h, sizeof(double) * N);
  return B200REG_OK;
}

// explicit matches + weights (HOST arrays) -> device scratch, as b200reg_point_to_plane does
int stagePairs(b200reg_ctx* ctx, const b200reg_cloud* reading, const b200reg_cloud* reference, const int32_t* ids, const float* d2,
               const float* w, int32_t k, PairArgs& A) {
  const int64_t total = reading->n * k;
  TRY(ensureBuf(ctx, ctx->matchIds, sizeof(int) * (size_t)total));
  TRY(ensureBuf(ctx, ctx->matchD2, sizeof(float) * (size_t)total));
  TRY(ensureBuf(ctx, ctx->misc, sizeof(float) * (size_t)total));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->matchIds.p, ids, sizeof(int) * (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->matchD2.p, d2, sizeof(float) * (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->misc.p, w, sizeof(float) * (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  memset(&A, 0, sizeof(A));
  A.reading = reading->xyz;
  A.refPts = reference->xyz;
  A.refNrm = reference->nrm;
  A.ids = (const int*)ctx->matchIds.p;
  A.d2 = (const float*)ctx->matchD2.p;
  A.w = (const float*)ctx->misc.p;
  A.total = total;
  A.k = k;
  return B200REG_OK;
}

// ---- small fp64 host linear algebra ------------------------------------------------------------------------------------------
// one-sided Jacobi SVD of a 3x3 (M = U diag(s) V^T, s descending like Eigen::JacobiSVD)
void svd3(const double M[9], double U[9], double S[3], double V[9]) {
  double A[9];
  memcpy(A, M, sizeof(A));
  for (int i = 0; i < 9; ++i) V[i] = (i % 4 == 0) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        double apq = 0, app = 0, aqq = 0;
        for (int r = 0; r < 3; ++r) { apq += A[3 * r + p] * A[3 * r + q]; app += A[3 * r + p] * A[3 * r + p]; aqq += A[3 * r + q] * A[3 * r + q]; }
        off = fmax(off, fabs(apq) / (sqrt(app * aqq) + 1e-300));
        if (fabs(apq) <= 1e-300) continue;
        const double zeta = (aqq - app) / (2.0 * apq);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
        for (int r = 0; r < 3; ++r) {
          const double x = A[3 * r + p], y = A[3 * r + q];
          A[3 * r + p] = c * x - s * y; A[3 * r + q] = s * x + c * y;
          const double vx = V[3 * r + p], vy = V[3 * r + q];
          V[3 * r + p] = c * vx - s * vy; V[3 * r + q] = s * vx + c * vy;
        }
      }
    if (off < 1e-15) break;
  }
  int order[3] = {0, 1, 2};
  double nrm[3];
  for (int c = 0; c < 3; ++c) nrm[c] = sqrt(A[c] * A[c] + A[3 + c] * A[3 + c] + A[6 + c] * A[6 + c]);
  for (int i = 0; i < 2; ++i)
    for (int j = i + 1; j < 3; ++j)
      if (nrm[order[j]] > nrm[order[i]]) { const int t = order[i]; order[i] = order[j]; order[j] = t; }
  double Vs[9];
  for (int c = 0; c < 3; ++c) {
    const int o = order[c];
    S[c] = nrm[o];
    for (int r = 0; r < 3; ++r) { U[3 * r + c] = nrm[o] > 0 ? A[3 * r + o] / nrm[o] : 0.0; Vs[3 * r + c] = V[3 * r + o]; }
  }
  memcpy(V, Vs, sizeof(Vs));
  // complete U when the matrix is rank deficient (columns of zero singular value): Gram-Schmidt against the others
  for (int c = 0; c < 3; ++c) {
    if (S[c] > 1e-300 * (S[0] + 1.0) && S[c] > 0) continue;
    double best[3] = {0, 0, 0}, bestN = -1;
    for (int e = 0; e < 3; ++e) {
      double v[3] = {e == 0 ? 1.0 : 0.0, e == 1 ? 1.0 : 0.0, e == 2 ? 1.0 : 0.0};
      for (int o = 0; o < 3; ++o) {
        if (o == c || (o > c && !(S[o] > 0))) continue;
        const double d = v[0] * U[o] + v[1] * U[3 + o] + v[2] * U[6 + o];
        for (int r = 0; r < 3; ++r) v[r] -= d * U[3 * r + o];
      }
      const double n2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
      if (n2 > bestN) { bestN = n2; memcpy(best, v, sizeof(best)); }
    }
    const double inv = 1.0 / sqrt(bestN);
    for (int r = 0; r < 3; ++r) U[3 * r + c] = best[r] * inv;
  }
}
double det3(const double R[9]) {
  return R[0] * (R[4] * R[8] - R[5] * R[7]) - R[1] * (R[3] * R[8] - R[5] * R[6]) + R[2] * (R[3] * R[7] - R[4] * R[6]);
}
// Gauss-Jordan inverse with partial pivoting of a 6x6 (Eigen's general inverse is PartialPivLU); false when singular
bool inv6(const double A[36], double X[36]) {
  double M[6][12];
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) { M[i][j] = A[6 * i + j]; M[i][6 + j] = i == j ? 1.0 : 0.0; }
  for (int c = 0; c < 6; ++c) {
    int piv = c;
    for (int r = c + 1; r < 6; ++r) if (fabs(M[r][c]) > fabs(M[piv][c])) piv = r;
    if (!(fabs(M[piv][c]) > 0.0)) return false;
    if (piv != c) for (int j = 0; j < 12; ++j) { const double t = M[c][j]; M[c][j] = M[piv][j]; M[piv][j] = t; }
    const double inv = 1.0 / M[c][c];
    for (int j = 0; j < 12; ++j) M[c][j] *= inv;
    for (int r = 0; r < 6; ++r) {
      if (r == c) continue;
      const double f = M[r][c];
      if (f != 0.0) for (int j = 0; j < 12; ++j) M[r][j] -= f * M[c][j];
    }
  }
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) X[6 * i + j] = M[i][6 + j];
  return true;
}

// ---------------------------------------------------------------------------------------------------------------------------
// compacting DataPointsFilters: one predicate kernel -> exclusive scan -> order-preserving scatter (setColFrom(j, cloud, i))
// ---------------------------------------------------------------------------------------------------------------------------
// std::minstd_rand: x_{i+1} = 48271 x_i mod (2^31 - 1); the i-th draw by square-and-multiply, so every point has its own
__device__ __forceinline__ unsigned minstdAt(unsigned x0, unsigned long long steps) {
  const unsigned long long M = 2147483647ull;
  unsigned long long a = 48271ull, r = x0;
  while (steps) {
    if (steps & 1ull) r = (r * a) % M;
    a = (a * a) % M;
    steps >>= 1;
  }
  return (unsigned)r;
}

__global__ void filterFlagKernel(const float4* __restrict__ xyz, int64_t n, b200reg_filter_params P, unsigned x0, unsigned* __restrict__ flags) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = xyz[i];
  bool keep = false;
  switch (P.kind) {
    case B200REG_FILTER_MAX_DIST:
    case B200REG_FILTER_MIN_DIST: {
      float v, lim = P.dist;
      if (P.dim == -1) {   // col.head(3).norm(): ((x^2 + y^2) + z^2), every operation rounded, then sqrt; |dist|
        v = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(p.x, p.x), __fmul_rn(p.y, p.y)), __fmul_rn(p.z, p.z)));
        lim = fabsf(lim);
      } else {
        v = P.dim == 0 ? p.x : (P.dim == 1 ? p.y : p.z);
      }
      keep = P.kind == B200REG_FILTER_MAX_DIST ? (v < lim) : (v > lim);   // MaxDist.cpp:82,93 / MinDist.cpp:78,90
      break;
    }
    case B200REG_FILTER_BOUNDING_BOX: {   // BoundingBox.cpp:78-88, strict on both sides
      const bool in = (p.x > P.box[0] && p.x < P.box[1]) && (p.y > P.box[2] && p.y < P.box[3]) && (p.z > P.box[4] && p.z < P.box[5]);
      keep = P.remove_inside ? !in : in;
      break;
    }
    default: {   // RandomSampling.cpp:62-80: draw i of minstd_rand(seed)
      const unsigned g = minstdAt(x0, (unsigned long long)i + 1ull);
      float r;
      if (P.random_method == 1) {   // std::uniform_real_distribution<float>(0, 1) = libstdc++ generate_canonical<float, 24>: one draw
        r = __fdiv_rn((float)(g - 1u), 2147483648.f);   // float(range + 1) = float(2147483646) rounds to 2^31
        if (r >= 1.f) r = 0.99999994f;
      } else {
        r = __fdiv_rn((float)g, 2147483648.f);          // float(max - min) = float(2147483645) rounds to 2^31
      }
      keep = (double)r < P.prob;
      break;
    }
  }
  flags[i] = keep ? 1u : 0u;
}
// RandomSampling.cpp:92: the loop stops once j > nbPointsOut, i.e. a point is kept only while at most nbPointsOut were kept before it
__global__ void filterScatterKernel(const unsigned* __restrict__ flags, const unsigned* __restrict__ offs, int64_t n, unsigned maxBefore,
                                    const float4* __restrict__ xyz, const float4* __restrict__ nrm, float4* __restrict__ oxyz,
                                    float4* __restrict__ onrm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !flags[i]) return;
  const unsigned o = offs[i];
  if (o > maxBefore) return;
  oxyz[o] = xyz[i];
  if (nrm) onrm[o] = nrm[i];
}

}  // namespace

extern "C" {

int32_t b200reg_cloud_filter(b200reg_ctx* ctx, b200reg_cloud* inout, const b200reg_filter_params* P) {
  if (!ctx || !inout || !P || P->kind < B200REG_FILTER_MAX_DIST || P->kind > B200REG_FILTER_RANDOM_SAMPLING)
    return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud_filter: bad arguments");
  if (P->kind <= B200REG_FILTER_MIN_DIST && (P->dim < -1 || P->dim > 2))   // MaxDist.cpp:68-72: dim >= features.rows() - 1
    return setErr(ctx, B200REG_E_INVALID_PARAM, "%s: Error, filtering on dimension number %d, larger than authorized axis id 2",
                  P->kind == B200REG_FILTER_MAX_DIST ? "MaxDistDataPointsFilter" : "MinDistDataPointsFilter", P->dim);
  if (P->kind == B200REG_FILTER_RANDOM_SAMPLING && !(P->prob >= 0.0 && P->prob <= 1.0))
    return setErr(ctx, B200REG_E_INVALID_PARAM, "RandomSamplingDataPointsFilter: prob must lie in [0, 1]");
  const int64_t n = inout->n;
  if (n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const size_t f4 = sizeof(float4) * (size_t)n, u4 = sizeof(unsigned) * (size_t)n;
  TRY(ensureBuf(ctx, ctx->misc, 2 * f4 + 2 * u4 + 64));
  float4* tmpXyz = (float4*)ctx->misc.p;
  float4* tmpNrm = tmpXyz + n;
  unsigned* flags = (unsigned*)(tmpNrm + n);
  unsigned* offs = flags + n;
  // minstd_rand(seed): state = seed mod (2^31 - 1), 0 -> 1 (std::linear_congruential_engine::seed)
  unsigned x0 = (unsigned)(P->seed % 2147483647ull);
  if (x0 == 0u) x0 = 1u;
  filterFlagKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(inout->xyz, n, *P, x0, flags);
  LAUNCH_CHECK(ctx);
  size_t tb = 0;
  CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tb, flags, offs, (int)n, ctx->stream));
  TRY(ensureBuf(ctx, ctx->cubtmp, tb));
  tb = ctx->cubtmp.cap;
  CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(ctx->cubtmp.p, tb, flags, offs, (int)n, ctx->stream));
  ctx->launches += 1;
  unsigned maxBefore = 0xffffffffu;
  if (P->kind == B200REG_FILTER_RANDOM_SAMPLING) maxBefore = (unsigned)(size_t)((double)(size_t)n * P->prob);   // nbPointsOut
  filterScatterKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(flags, offs, n, maxBefore, inout->xyz, inout->nrm, tmpXyz, tmpNrm);
  LAUNCH_CHECK(ctx);
  unsigned* h = (unsigned*)ctx->pinned;
  CUDA_TRY(ctx, cudaMemcpyAsync(h, offs + (n - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(h + 1, flags + (n - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  int64_t kept = (int64_t)h[0] + (int64_t)h[1];
  if (kept > (int64_t)maxBefore + 1) kept = (int64_t)maxBefore + 1;
  CUDA_TRY(ctx, cudaMemcpyAsync(inout->xyz, tmpXyz, sizeof(float4) * (size_t)kept, cudaMemcpyDeviceToDevice, ctx->stream));
  if (inout->nrm) CUDA_TRY(ctx, cudaMemcpyAsync(inout->nrm, tmpNrm, sizeof(float4) * (size_t)kept, cudaMemcpyDeviceToDevice, ctx->stream));
  inout->n = kept;
  return B200REG_OK;
}

int32_t b200reg_var_trimmed_dist(b200reg_ctx* ctx, const float* d2, int64_t count, float min_ratio, float max_ratio, float lambda,
                                 float* ratio_out, float* tau_out, float* w_out) {
  if (!ctx || count < 0 || (count > 0 && !d2)) return setErr(ctx, B200REG_E_INVALID_PARAM, "var_trimmed_dist: bad arguments");
  if (!(min_ratio < max_ratio))   // OutlierFiltersImpl.cpp:160-163
    return setErr(ctx, B200REG_E_INVALID_PARAM, "VarTrimmedDistOutlierFilter: minRatio (%g) should be smaller than maxRatio (%g)", (double)min_ratio, (double)max_ratio);
  if (count == 0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "Inlier ratio optimization failed due to absence of matches");
  if (count >= (1ll << 31)) return setErr(ctx, B200REG_E_INVALID_PARAM, "var_trimmed_dist: more than 2^31 - 1 matches");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  TRY(ensureBuf(ctx, ctx->matchD2, sizeof(float) * (size_t)count));
  TRY(ensureBuf(ctx, ctx->keys0, sizeof(unsigned) * (size_t)count));
  TRY(ensureBuf(ctx, ctx->keys1, sizeof(unsigned) * (size_t)count));
  float* dd = (float*)ctx->matchD2.p;
  unsigned* k0 = (unsigned*)ctx->keys0.p;
  unsigned* k1 = (unsigned*)ctx->keys1.p;
  unsigned* nValid = (unsigned*)((char*)ctx->dstate + 32768);   // scratch window of ctx->dstate (common.cuh)
  VtResult* res = (VtResult*)(nValid + 4);
  CUDA_TRY(ctx, cudaMemcpyAsync(dd, d2, sizeof(float) * (size_t)count, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(nValid, 0, 64, ctx->stream));
  vtKeyKernel<<<divUp(count, 256), 256, 0, ctx->stream>>>(dd, count, k0, nValid);
  LAUNCH_CHECK(ctx);
  size_t tb = 0;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortKeys(nullptr, tb, k0, k1, (int)count, 0, 32, ctx->stream));
  TRY(ensureBuf(ctx, ctx->cubtmp, tb));
  tb = ctx->cubtmp.cap;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortKeys(ctx->cubtmp.p, tb, k0, k1, (int)count, 0, 32, ctx->stream));
  ctx->launches += 5;
  // floor(ratio * points_nbr): float * int -> float product, then floor -> int (OutlierFiltersImpl.cpp:200-201)
  const int64_t minEl = (int64_t)floorf(min_ratio * (float)count), maxEl = (int64_t)floorf(max_ratio * (float)count);
  vtScanKernel<<<1, 32, 0, ctx->stream>>>(k1, nValid, count, minEl, maxEl, lambda, res);
  LAUNCH_CHECK(ctx);
  VtResult* h = (VtResult*)ctx->pinned;
  CUDA_TRY(ctx, cudaMemcpyAsync(h, res, sizeof(VtResult), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (h->nValid == 0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "Inlier ratio optimization failed due to absence of matches");
  if (h->empty) return setErr(ctx, B200REG_E_NOT_CONVERGED, "VarTrimmedDistOutlierFilter: no valid match inside [minRatio, maxRatio)");
  const float ratio = (float)((int64_t)h->minIndex + minEl) / (float)count;   // :217
  if (ratio_out) *ratio_out = ratio;
  if (!tau_out && !w_out) return B200REG_OK;
  return b200reg_trimmed_dist(ctx, d2, count, ratio, tau_out, w_out);   // limit = getDistsQuantile(tunedRatio), w = (d2 <= limit) (:172-173)
}

int32_t b200reg_point_to_point(b200reg_ctx* ctx, const b200reg_cloud* reading, const b200reg_cloud* reference, const int32_t* ids,
                               const float* d2, const float* w, int32_t k, float T_out[16], b200reg_step_stats* stats) {
  if (!ctx || !reading || !reference || !ids || !d2 || !w || !T_out || k < 1) return setErr(ctx, B200REG_E_INVALID_PARAM, "point_to_point: bad arguments");
  if (reading->n * k == 0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "ErrorMnimizer: no point to minimize");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  PairArgs A;
  TRY(stagePairs(ctx, reading, reference, ids, d2, w, k, A));
  double m[P2P_N];
  TRY((reducePairs<P2P_N>(ctx, p2pMomentsKernel, A, m)));
  if (stats) {
    memset(stats, 0, sizeof(*stats));
    stats->kept = (int64_t)m[16];
    stats->n_finite = (int64_t)m[17];
    stats->weight_sum = m[0];
    stats->tau = INFINITY;
    stats->point_used_ratio = (float)m[16] / (float)A.total;
    stats->weighted_point_used_ratio = (float)m[0] / (float)A.total;
  }
  if (m[16] < 1.0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "ErrorMnimizer: no point to minimize");
  // PointToPoint.cpp:62-73: weighted means, cross-covariance of the centred pairs  m = sum w (q - mq)(p - mp)^T
  const double W = m[0];
  double mp[3], mq[3], C[9];
  for (int d = 0; d < 3; ++d) { mp[d] = m[1 + d] / W; mq[d] = m[4 + d] / W; }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C[3 * i + j] = m[7 + 3 * i + j] - W * mq[i] * mp[j];
  double U[9], S[3], V[9], R[9];
  svd3(C, U, S, V);
  auto mulUVt = [&](const double* Vm) {
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) R[3 * i + j] = U[3 * i] * Vm[3 * j] + U[3 * i + 1] * Vm[3 * j + 1] + U[3 * i + 2] * Vm[3 * j + 2];
  };
  mulUVt(V);
  if (det3(R) < 0.0) {   // :80-85: reflection -> flip the last row of V^T (= last column of V)
    for (int r = 0; r < 3; ++r) V[3 * r + 2] = -V[3 * r + 2];
    mulUVt(V);
  }
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) T_out[4 * i + j] = (float)R[3 * i + j];
    T_out[4 * i + 3] = (float)(mq[i] - (R[3 * i] * mp[0] + R[3 * i + 1] * mp[1] + R[3 * i + 2] * mp[2]));   // :86
  }
  T_out[12] = T_out[13] = T_out[14] = 0.f;
  T_out[15] = 1.f;
  if (stats) for (int d = 0; d < 3; ++d) { stats->mean_reading[d] = mp[d]; stats->mean_reference[d] = mq[d]; }
  return B200REG_OK;
}

int32_t b200reg_point_to_plane_residual(b200reg_ctx* ctx, const b200reg_cloud* reading, const b200reg_cloud* reference, const int32_t* ids,
                                        const float* d2, const float* w, int32_t k, int32_t force2D, double* residual_out, int64_t* kept_out) {
  if (!ctx || !reading || !reference || !ids || !d2 || !w || !residual_out || k < 1) return setErr(ctx, B200REG_E_INVALID_PARAM, "point_to_plane_residual: bad arguments");
  if (!reference->nrm) return setErr(ctx, B200REG_E_INVALID_FIELD, "PointToPlaneErrorMinimizer: reference cloud has no 'normals' descriptor");
  *residual_out = 0.0;
  if (kept_out) *kept_out = 0;
  if (reading->n * k == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  PairArgs A;
  TRY(stagePairs(ctx, reading, reference, ids, d2, w, k, A));
  A.alpha = force2D ? 1.0 : 0.0;
  double m[RES_N];
  TRY((reducePairs<RES_N>(ctx, residualKernel, A, m)));
  *residual_out = m[0];
  if (kept_out) *kept_out = (int64_t)m[1];
  return B200REG_OK;
}

int32_t b200reg_point_to_plane_covariance(b200reg_ctx* ctx, const b200reg_cloud* reading, const b200reg_cloud* reference, const int32_t* ids,
                                          const float* d2, const float* w, int32_t k, const double mean_reading[3],
                                          const double mean_reference[3], const float T[16], float sensor_std_dev, double cov_out[36]) {
  if (!ctx || !reading || !reference || !ids || !d2 || !w || !mean_reading || !mean_reference || !T || !cov_out || k < 1)
    return setErr(ctx, B200REG_E_INVALID_PARAM, "point_to_plane_covariance: bad arguments");
  if (!reference->nrm) {   // PointToPlaneWithCov.cpp:91-92: max * Identity
    for (int i = 0; i < 36; ++i) cov_out[i] = (i % 7 == 0) ? (double)FLT_MAX : 0.0;
    return B200REG_OK;
  }
  if (reading->n * k == 0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "ErrorMnimizer: no point to minimize");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  PairArgs A;
  TRY(stagePairs(ctx, reading, reference, ids, d2, w, k, A));
  for (int d = 0; d < 3; ++d) { A.meanRead[d] = mean_reading[d]; A.meanRef[d] = mean_reference[d]; }
  // :94-99, the small-angle Euler extraction from the step transform
  A.beta = -asin((double)T[8]);
  A.alpha = atan2((double)T[9], (double)T[10]);
  A.gamma = atan2((double)T[4] / cos(A.beta), (double)T[0] / cos(A.beta));
  A.tx = T[3]; A.ty = T[7]; A.tz = T[11];
  double m[COV_N];
  TRY((reducePairs<COV_N>(ctx, covMomentsKernel, A, m)));
  if (m[42] < 1.0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "ErrorMnimizer: no point to minimize");
  double H[36], C[36], Hi[36], tmp[36];
  int t = 0;
  for (int i = 0; i < 6; ++i)
    for (int j = i; j < 6; ++j, ++t) { H[6 * i + j] = H[6 * j + i] = m[t]; C[6 * i + j] = C[6 * j + i] = m[21 + t]; }
  if (!inv6(H, Hi)) {
    for (int i = 0; i < 36; ++i) cov_out[i] = NAN;   // what inverting a singular J_hessian yields in the reference (:152)
    return B200REG_OK;
  }
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) { double s = 0; for (int l = 0; l < 6; ++l) s += Hi[6 * i + l] * C[6 * l + j]; tmp[6 * i + j] = s; }
  const double s2 = (double)sensor_std_dev * (double)sensor_std_dev;
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) { double s = 0; for (int l = 0; l < 6; ++l) s += tmp[6 * i + l] * Hi[6 * l + j]; cov_out[6 * i + j] = s2 * s; }
  return B200REG_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------------
// The allocation-free variant of b200reg_point_to_plane_residual (no counterpart in the reference; added
// for callers that may not call cudaMalloc).  Its scratch — the staged matches, distances and weights, the
// per-block partial sums and the result — is this file-scope device array, which every GPU that loads the library holds until
// exit (12 MiB + 8 KiB).  In host code `g_noallocScratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, taken once per device with that device current.  All calls share the array: calls that overlap on the
// GPU (different streams / contexts of one device) race.
// ---------------------------------------------------------------------------------------------------------------------------
constexpr size_t NOALLOC_PAIRS = B200REG_NOALLOC_MAX_PAIRS;
constexpr int NOALLOC_MAX_BLOCKS = 511;                                      // partials: (blocks + 1) x RES_N doubles <= 8 KiB
constexpr size_t NOALLOC_TAIL = sizeof(double) * RES_N * (NOALLOC_MAX_BLOCKS + 1);
constexpr size_t NOALLOC_BYTES = NOALLOC_PAIRS * 12 + NOALLOC_TAIL;
static __device__ __align__(256) unsigned char g_noallocScratch[NOALLOC_BYTES];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noallocMutex;
unsigned char* g_noallocAddr[NOALLOC_MAX_DEVICES] = {nullptr};

// device address of the scratch on ctx's device (that device must be current)
int noallocScratch(b200reg_ctx* ctx, unsigned char** out) {
  if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES) return setErr(ctx, B200REG_E_INVALID_PARAM, "noalloc: device ordinal %d out of range", ctx->device);
  std::lock_guard<std::mutex> lock(g_noallocMutex);
  if (!g_noallocAddr[ctx->device]) {
    unsigned char* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noallocScratch);
    if (e != cudaSuccess || !ptr) return setErr(ctx, B200REG_E_CUDA, "noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
    g_noallocAddr[ctx->device] = ptr;
  }
  *out = g_noallocAddr[ctx->device];
  return B200REG_OK;
}
}  // namespace

extern "C" int32_t b200reg_point_to_plane_residual_noalloc(b200reg_ctx* ctx, const b200reg_cloud* reading, const b200reg_cloud* reference,
                                                           const int32_t* ids, const float* d2, const float* w, int32_t k, int32_t force2D,
                                                           double* residual_out, int64_t* kept_out) {
  if (!ctx || !reading || !reference || !ids || !d2 || !w || !residual_out || k < 1)
    return setErr(ctx, B200REG_E_INVALID_PARAM, "point_to_plane_residual_noalloc: bad arguments");
  if (!reference->nrm) return setErr(ctx, B200REG_E_INVALID_FIELD, "PointToPlaneErrorMinimizer: reference cloud has no 'normals' descriptor");
  const int64_t total = reading->n * (int64_t)k;
  if (total > (int64_t)NOALLOC_PAIRS)
    return setErr(ctx, B200REG_E_INVALID_PARAM, "point_to_plane_residual_noalloc: %lld pairs exceed the static scratch (B200REG_NOALLOC_MAX_PAIRS = %lld)",
                  (long long)total, (long long)NOALLOC_PAIRS);
  *residual_out = 0.0;
  if (kept_out) *kept_out = 0;
  if (total == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  unsigned char* base = nullptr;
  TRY(noallocScratch(ctx, &base));
  int* dIds = (int*)base;
  float* dD2 = (float*)(base + NOALLOC_PAIRS * 4);
  float* dW = (float*)(base + NOALLOC_PAIRS * 8);
  double* part = (double*)(base + NOALLOC_PAIRS * 12);
  const int blocks = (int)fmax(1.0, fmin((double)divUp(total, 1024), fmin((double)(2 * ctx->sm_count), (double)NOALLOC_MAX_BLOCKS)));   // reducePairs' grid
  double* res = part + (size_t)blocks * RES_N;
  // the array keeps the previous call's contents: everything read below is overwritten first (the partials and the result are
  // cleared as well, although every block stores its own)
  CUDA_TRY(ctx, cudaMemsetAsync(part, 0, NOALLOC_TAIL, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(dIds, ids, sizeof(int) * (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(dD2, d2, sizeof(float) * (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(dW, w, sizeof(float) * (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  PairArgs A;
  memset(&A, 0, sizeof(A));
  A.reading = reading->xyz; A.refPts = reference->xyz; A.refNrm = reference->nrm;
  A.ids = dIds; A.d2 = dD2; A.w = dW;
  A.total = total; A.k = k;
  A.alpha = force2D ? 1.0 : 0.0;
  residualKernel<<<blocks, 256, 0, ctx->stream>>>(A, part);
  LAUNCH_CHECK(ctx);
  sumPartialsKernel<<<1, 64, 0, ctx->stream>>>(part, blocks, RES_N, res);
  LAUNCH_CHECK(ctx);
  double m[RES_N] = {0.0, 0.0};
  CUDA_TRY(ctx, cudaMemcpyAsync(m, res, sizeof(m), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  *residual_out = m[0];
  if (kept_out) *kept_out = (int64_t)m[1];
  return B200REG_OK;
}
