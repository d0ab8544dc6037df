// csrc/normals.cu — SurfaceNormalDataPointsFilter::inPlaceFilter (DataPointsFilters/SurfaceNormal.cpp:83-330):
// self-kNN (incl. the point itself) over a voxel-grid index of the cloud, per-point mean / covariance in fp64 over the
// fp32 neighbour coordinates, rank test, symmetric 3x3 eigen-decomposition, normal = eigenvector of the smallest
// eigenvalue, optional densities / eigenvalues / linearity-planarity-curvature, optional orientation toward a point.
//
// This file is compiled with --fmad=false: every fp operation is individually rounded, so the arithmetic below is the
// same sequence of IEEE operations the oracle's mixed-precision mode performs.
#include <cmath>

#include "common.cuh"

int runKnn(b200reg_ctx* ctx, const b200reg_index* ix, const float4* queries, int64_t n, const float* T, int k, float maxDist, int allowSelf,
           int seeded);

namespace {

// rank of a 3x3 matrix by Householder QR with full pivoting in fp32 — the test `C.fullPivHouseholderQr().rank() + 1 >= 3`
// of SurfaceNormal.cpp:210 in the reference's scalar T = float (threshold eps * size * max pivot).
__device__ int rank3Fp32(const double* C) {
  float R[3][3];
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) R[i][j] = (float)C[3 * i + j];
  float piv[3] = {0.f, 0.f, 0.f};
  float maxpivot = 0.f;
  for (int k = 0; k < 3; ++k) {
    int pr = k, pc = k;
    float best = -1.f;
    for (int i = k; i < 3; ++i)
      for (int j = k; j < 3; ++j)
        if (fabsf(R[i][j]) > best) { best = fabsf(R[i][j]); pr = i; pc = j; }
    if (best <= 0.f) break;
    if (pr != k) for (int j = 0; j < 3; ++j) { const float t = R[pr][j]; R[pr][j] = R[k][j]; R[k][j] = t; }
    if (pc != k) for (int i = 0; i < 3; ++i) { const float t = R[i][pc]; R[i][pc] = R[i][k]; R[i][k] = t; }
    float norm2 = 0.f;
    for (int i = k; i < 3; ++i) norm2 += R[i][k] * R[i][k];
    const float x0 = R[k][k];
    const float alpha = (x0 >= 0.f) ? -sqrtf(norm2) : sqrtf(norm2);
    float v[3] = {0.f, 0.f, 0.f};
    v[k] = x0 - alpha;
    for (int i = k + 1; i < 3; ++i) v[i] = R[i][k];
    float vnorm2 = 0.f;
    for (int i = k; i < 3; ++i) vnorm2 += v[i] * v[i];
    if (vnorm2 > 0.f) {
      for (int j = k; j < 3; ++j) {
        float dot = 0.f;
        for (int i = k; i < 3; ++i) dot += v[i] * R[i][j];
        const float f = 2.f * dot / vnorm2;
        for (int i = k; i < 3; ++i) R[i][j] -= f * v[i];
      }
    }
    for (int i = k + 1; i < 3; ++i) R[i][k] = 0.f;
    piv[k] = fabsf(R[k][k]);
    maxpivot = fmaxf(maxpivot, piv[k]);
  }
  const float thr = 1.1920928955078125e-7f * 3.f * maxpivot;
  int rank = 0;
  for (int k = 0; k < 3; ++k) rank += (piv[k] > thr) ? 1 : 0;
  return rank;
}

// symmetric 3x3 eigen-decomposition, cyclic Jacobi in fp64; eigenvalues ascending, eigenvectors in columns
__device__ void symEig3(const double* Cin, double val[3], double vec[3][3]) {
  double a[3][3], V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) a[i][j] = Cin[3 * i + j];
  for (int sweep = 0; sweep < 64; ++sweep) {
    const double off = a[0][1] * a[0][1] + a[0][2] * a[0][2] + a[1][2] * a[1][2];
    if (off == 0.0) break;
#pragma unroll
    for (int p = 0; p < 3; ++p)
#pragma unroll
      for (int q = p + 1; q < 3; ++q) {
        // converged pair: an off-diagonal below 2^-70 of its diagonals rotates by an angle no fp64 result can see; without this
        // the residue left by each rotation never reaches exactly 0 and most matrices run all 64 sweeps
        if (fabs(a[p][q]) <= 8.470329472543003e-22 * (fabs(a[p][p]) + fabs(a[q][q]))) { a[p][q] = 0.0; a[q][p] = 0.0; continue; }
        const double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
        const double t = ((theta >= 0) ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
#pragma unroll
        for (int k = 0; k < 3; ++k) { const double x = a[k][p], y = a[k][q]; a[k][p] = c * x - s * y; a[k][q] = s * x + c * y; }
#pragma unroll
        for (int k = 0; k < 3; ++k) { const double x = a[p][k], y = a[q][k]; a[p][k] = c * x - s * y; a[q][k] = s * x + c * y; }
#pragma unroll
        for (int k = 0; k < 3; ++k) { const double x = V[k][p], y = V[k][q]; V[k][p] = c * x - s * y; V[k][q] = s * x + c * y; }
      }
  }
  // ascending order; stable for equal values (matches std::sort on distinct diagonals; ties are measure-zero)
  int o0 = 0, o1 = 1, o2 = 2;
  if (a[o1][o1] < a[o0][o0]) { const int t = o0; o0 = o1; o1 = t; }
  if (a[o2][o2] < a[o1][o1]) { const int t = o1; o1 = o2; o2 = t; }
  if (a[o1][o1] < a[o0][o0]) { const int t = o0; o0 = o1; o1 = t; }
  const int order[3] = {o0, o1, o2};
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    double v = 0.0;
#pragma unroll
    for (int r = 0; r < 3; ++r) v = (order[c] == r) ? a[r][r] : v;
    val[c] = v;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      double e = 0.0;
#pragma unroll
      for (int cc = 0; cc < 3; ++cc) e = (order[c] == cc) ? V[r][cc] : e;
      vec[r][c] = e;
    }
  }
}

__global__ void __launch_bounds__(128) normalsKernel(const float4* __restrict__ xyz, const float4* __restrict__ sortedPts,
                                                     const int* __restrict__ pos, const float* __restrict__ d2, int64_t n, int k,
                                                     int orient, float ox, float oy, float oz, float4* __restrict__ outN,
                                                     float* __restrict__ dens, float* __restrict__ eigv, float* __restrict__ lpc,
                                                     float* __restrict__ eigVec, float* __restrict__ meanDist, const int* __restrict__ ids,
                                                     float* __restrict__ matchedIds, int* __restrict__ degenerate) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int* myPos = pos + (int64_t)k * i;
  const float* myD = d2 + (int64_t)k * i;
  // mean of the valid neighbours (SurfaceNormal.cpp:190-201), neighbour order = ascending (d2, id)
  double mean[3] = {0, 0, 0};
  int realKnn = 0;
  for (int j = 0; j < k; ++j) {
    if (myD[j] == CUDART_INF_F) continue;
    const float4 p = sortedPts[myPos[j]];
    mean[0] += (double)p.x; mean[1] += (double)p.y; mean[2] += (double)p.z;
    ++realKnn;
  }
  double C[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  double maxSq = 0.0;
  if (realKnn > 0) {
    for (int c = 0; c < 3; ++c) mean[c] /= (double)realKnn;
    for (int j = 0; j < k; ++j) {
      if (myD[j] == CUDART_INF_F) continue;
      const float4 p = sortedPts[myPos[j]];
      const double nn[3] = {(double)p.x - mean[0], (double)p.y - mean[1], (double)p.z - mean[2]};
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) C[3 * r + c] += nn[r] * nn[c];   // :203-204
      maxSq = fmax(maxSq, nn[0] * nn[0] + nn[1] * nn[1] + nn[2] * nn[2]);
    }
    for (int q = 0; q < 9; ++q) C[q] /= (double)realKnn;
  }
  const bool isDeg = (realKnn == 0) || !(rank3Fp32(C) + 1 >= 3);   // :210
  double val[3] = {0, 0, 0}, vec[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
  if (!isDeg) symEig3(C, val, vec);
  else atomicAdd(degenerate, 1);
  const float4 p = xyz[i];
  double nv[3] = {vec[0][0], vec[1][0], vec[2][0]};   // eigenvector of the smallest eigenvalue (:237-243)
  if (orient && !isDeg) {                              // SurfaceNormalEstimatorPCA.cpp:69-84
    const double dot = (double)(ox - p.x) * nv[0] + (double)(oy - p.y) * nv[1] + (double)(oz - p.z) * nv[2];
    if (dot < 0) { nv[0] = -nv[0]; nv[1] = -nv[1]; nv[2] = -nv[2]; }
  }
  outN[i] = make_float4((float)fmin(1.0, fmax(-1.0, nv[0])), (float)fmin(1.0, fmax(-1.0, nv[1])), (float)fmin(1.0, fmax(-1.0, nv[2])), 0.f);   // :245
  if (dens) {   // utils.h:105-128
    float dd = 0.f;
    if (!isDeg) {
      const float m2 = (float)maxSq;
      const float volume = (float)(4. / 3.) * 3.14159274101257324f * (m2 * sqrtf(m2));
      dd = (volume > 0.f) ? ((float)realKnn / volume) : 0.f;
    }
    dens[i] = dd;
  }
  if (eigv) { eigv[3 * i] = (float)val[0]; eigv[3 * i + 1] = (float)val[1]; eigv[3 * i + 2] = (float)val[2]; }
  if (eigVec)   // serializeEigVec: row-major, eigenvectors in columns (utils.h:90-101; SurfaceNormal.cpp:256-257); zeros when degenerate
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) eigVec[9 * i + 3 * r + c] = (float)vec[r][c];
  if (meanDist) {   // SurfaceNormal.cpp:258-268: |point - mean of the neighbours|; numeric_limits<size_t>::max() when degenerate
    if (isDeg) meanDist[i] = 18446744073709551615.0f;
    else {
      const double dx = (double)p.x - mean[0], dy = (double)p.y - mean[1], dz = (double)p.z - mean[2];
      meanDist[i] = (float)sqrt(dx * dx + dy * dy + dz * dz);
    }
  }
  if (matchedIds)   // matches.ids cast to T (:294-297); -1 where there is no neighbour
    for (int j = 0; j < k; ++j) matchedIds[(int64_t)k * i + j] = (float)ids[(int64_t)k * i + j];
  if (lpc) {   // SurfaceNormalEstimatorPCA.cpp:86-99
    float l = 0.f, pl = 0.f, cu = 0.f;
    if (!isDeg) {
      const float s1 = sqrtf((float)val[0]), s2 = sqrtf((float)val[1]), s3 = sqrtf((float)val[2]);
      l = (s3 - s2) / s3;
      pl = (s2 - s1) / s3;
      cu = s1 / (s1 + s2 + s3);
    }
    lpc[3 * i] = l; lpc[3 * i + 1] = pl; lpc[3 * i + 2] = cu;
  }
}

// smoothNormals (SurfaceNormal.cpp:299-323).  The reference overwrites normals->col(i) inside the loop, so point i averages the
// ALREADY SMOOTHED normals of its neighbours with a smaller index and the original normals of the others (itself included) — a
// sequential sweep.  Reproduced exactly by a dependency-ordered pipeline: warps take points in index order from a ticket counter
// (so every smaller index is held by a running or finished warp: no deadlock), wait for the done flags of their lower-index
// neighbours, and read `newN` for those and `oldN` for the rest.  The fp32 sum runs in neighbour order (ascending (d2, index)),
// the sign test is the reference's currentNormal.dot(normal) > 0 on individually rounded operations.
__global__ void __launch_bounds__(256) smoothNormalsKernel(const float4* __restrict__ oldN, float4* newN, const int* __restrict__ ids,
                                                           const float* __restrict__ d2, long long n, int k, unsigned long long* ticket,
                                                           int* done) {
  const int lane = threadIdx.x & 31;
  for (;;) {
    long long i = 0;
    if (lane == 0) i = (long long)atomicAdd(ticket, 1ull);
    i = __shfl_sync(0xffffffffu, i, 0);
    if (i >= n) return;
    const float4 cur = oldN[i];
    float mx = 0.f, my = 0.f, mz = 0.f;
    int cnt = 0;
    for (int j0 = 0; j0 < k; j0 += 32) {
      const int j = j0 + lane;
      bool valid = false;
      float cx = 0.f, cy = 0.f, cz = 0.f;
      if (j < k && d2[(long long)k * i + j] != CUDART_INF_F) {
        const int id = ids[(long long)k * i + j];
        float4 nb;
        if (id < i) {   // already smoothed in the reference's sweep
          while (atomicAdd(&done[id], 0) == 0) __nanosleep(20);
          __threadfence();
          nb = __ldcg(newN + id);
        } else nb = oldN[id];
        const float dot = __fadd_rn(__fadd_rn(__fmul_rn(cur.x, nb.x), __fmul_rn(cur.y, nb.y)), __fmul_rn(cur.z, nb.z));
        const float sgn = (dot > 0.f) ? 1.f : -1.f;   // mean += normal, else mean -= normal (flip)
        cx = sgn * nb.x; cy = sgn * nb.y; cz = sgn * nb.z;
        valid = true;
      }
      const unsigned vm = __ballot_sync(0xffffffffu, valid);
      for (int jj = 0; jj < 32; ++jj) {   // neighbour order, every lane keeps the same running sum
        const float vx = __shfl_sync(0xffffffffu, cx, jj), vy = __shfl_sync(0xffffffffu, cy, jj), vz = __shfl_sync(0xffffffffu, cz, jj);
        if (vm & (1u << jj)) { mx = __fadd_rn(mx, vx); my = __fadd_rn(my, vy); mz = __fadd_rn(mz, vz); ++cnt; }
      }
    }
    if (lane == 0) {
      const float fn = (float)cnt;   // mean / T(n): 0 / 0 = NaN when the point has no neighbour at all, like the reference
      newN[i] = make_float4(__fdiv_rn(mx, fn), __fdiv_rn(my, fn), __fdiv_rn(mz, fn), 0.f);
      __threadfence();
      atomicExch(&done[i], 1);
    }
  }
}

}  // namespace

// ---- Open3D EstimateNormals(Hybrid(radius, knn)) + OrientNormalsTowardsCameraLocation on an fp64 cloud (CloudRegistration.cpp:62-75),
// in Open3D's scalar.  The neighbour CANDIDATES come from the fp32 index (knn + 2 nearest of the rounded copy, radius widened by 1e-5);
// this kernel re-measures them in fp64, keeps the (up to) knn nearest with d2 < radius^2 by (d2, index), and forms mean, covariance
// and the eigenvector of the smallest eigenvalue from the fp64 coordinates — at 80 m an fp32 coordinate carries ~4e-6 m of rounding,
// ~1e-4 rad of normal noise against the 1e-5 rad bar, which is why the points are not rounded first.
#define N64_MAXC 32
__global__ void __launch_bounds__(128) normals64Kernel(const double* __restrict__ xyz, const int* __restrict__ ids, int64_t n, int kc, int knn,
                                                       double r2, int orient, double ox, double oy, double oz, double* __restrict__ outN,
                                                       int* __restrict__ degenerate) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double px = xyz[3 * i], py = xyz[3 * i + 1], pz = xyz[3 * i + 2];
  double cd[N64_MAXC];
  int ci[N64_MAXC];
  int cnt = 0;
  for (int c = 0; c < kc; ++c) {   // insertion sort by (d2, index) over at most 32 candidates
    const int j = ids[(int64_t)kc * i + c];
    if (j < 0) continue;
    const double dx = px - xyz[3 * (int64_t)j], dy = py - xyz[3 * (int64_t)j + 1], dz = pz - xyz[3 * (int64_t)j + 2];
    const double d = (dx * dx + dy * dy) + dz * dz;
    if (!(d < r2)) continue;
    int pos = cnt;
    while (pos > 0 && (cd[pos - 1] > d || (cd[pos - 1] == d && ci[pos - 1] > j))) { cd[pos] = cd[pos - 1]; ci[pos] = ci[pos - 1]; --pos; }
    cd[pos] = d; ci[pos] = j;
    ++cnt;
  }
  if (cnt > knn) cnt = knn;
  double nv[3] = {0.0, 0.0, 0.0};
  if (cnt < 3) atomicAdd(degenerate, 1);
  else {
    double mean[3] = {0, 0, 0}, C[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int t = 0; t < cnt; ++t) { const int64_t j = ci[t]; mean[0] += xyz[3 * j]; mean[1] += xyz[3 * j + 1]; mean[2] += xyz[3 * j + 2]; }
    for (int c = 0; c < 3; ++c) mean[c] /= (double)cnt;
    for (int t = 0; t < cnt; ++t) {
      const int64_t j = ci[t];
      const double d[3] = {xyz[3 * j] - mean[0], xyz[3 * j + 1] - mean[1], xyz[3 * j + 2] - mean[2]};
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) C[3 * r + c] += d[r] * d[c];
    }
    for (int q = 0; q < 9; ++q) C[q] /= (double)cnt;
    double val[3], vec[3][3];
    symEig3(C, val, vec);
    nv[0] = vec[0][0]; nv[1] = vec[1][0]; nv[2] = vec[2][0];
    if (orient) {
      const double dot = (ox - px) * nv[0] + (oy - py) * nv[1] + (oz - pz) * nv[2];
      if (dot < 0) { nv[0] = -nv[0]; nv[1] = -nv[1]; nv[2] = -nv[2]; }
    }
  }
  outN[3 * i] = nv[0]; outN[3 * i + 1] = nv[1]; outN[3 * i + 2] = nv[2];
}

// `rounded`: the fp32 copy of `cloud` (same order) the index is built over; writes cloud->nrm (fp64)
int estimateNormals64(b200reg_ctx* ctx, b200reg_cloud64* cloud, const b200reg_cloud* rounded, int knn, double radius, const double* origin,
                      int32_t* degenerate_count) {
  const int64_t n = cloud->n;
  if (knn < 3) return setErr(ctx, B200REG_E_INVALID_PARAM, "estimate_normals: knn must be >= 3");
  if (!(radius >= 0.0)) return setErr(ctx, B200REG_E_INVALID_PARAM, "estimate_normals: radius must be >= 0");
  const int kc = (knn + 2 <= N64_MAXC) ? knn + 2 : knn;   // two spare candidates: an fp32 / fp64 order swap at the k-th place needs a near-tie (~1e-3 of the points), two in a row ~1e-6
  if (kc > 64) return setErr(ctx, B200REG_E_INVALID_PARAM, "estimate_normals: knn above 64 is not supported");
  if (kc > N64_MAXC) return setErr(ctx, B200REG_E_INVALID_PARAM, "estimate_normals: knn above %d is not supported on the fp64 path", N64_MAXC);
  b200reg_index* ix = nullptr;
  TRY(buildIndex(ctx, rounded, 0, 0.f, &ix));
  const float wide = (float)(radius * 1.00001 + 1e-6);
  int st = runKnn(ctx, ix, ctx->knnLanes == 1 ? rounded->xyz : nullptr, n, nullptr, kc, wide, 1, 0);
  if (st != B200REG_OK) { b200reg_index_free(ctx, ix); return st; }
  int* dDeg = (int*)((char*)ctx->dstate + 32768);
  cudaMemsetAsync(dDeg, 0, sizeof(int), ctx->stream);
  normals64Kernel<<<divUp(n, 128), 128, 0, ctx->stream>>>(cloud->xyz, (const int*)ctx->matchIds.p, n, kc, knn, radius * radius, origin ? 1 : 0,
                                                          origin ? origin[0] : 0.0, origin ? origin[1] : 0.0, origin ? origin[2] : 0.0, cloud->nrm, dDeg);
  ctx->launches++;
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(ctx->pinned, dDeg, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  const bool indexOk = (e != cudaSuccess) || indexHeader(ctx, ix) != nullptr;
  b200reg_index_free(ctx, ix);
  if (e != cudaSuccess) return setErr(ctx, B200REG_E_CUDA, "estimate_normals: %s", cudaGetErrorString(e));
  if (!indexOk) return setErr(ctx, B200REG_E_INVALID_PARAM, "estimate_normals: non-finite coordinates in the cloud");
  if (degenerate_count) *degenerate_count = *(const int*)ctx->pinned;
  return B200REG_OK;
}

extern "C" int32_t b200reg_surface_normals_ex(b200reg_ctx* ctx, b200reg_cloud* inout, const b200reg_normals_options* opt, int32_t* degenerate_count) {
  if (!ctx || !inout || !opt) return setErr(ctx, B200REG_E_INVALID_PARAM, "surface_normals: bad arguments");
  const int32_t knn = opt->knn;
  if (knn < 3) return setErr(ctx, B200REG_E_INVALID_PARAM, "SurfaceNormalDataPointsFilter: knn must be >= 3");   // SurfaceNormal.h:69
  if (!(opt->epsilon >= 0.f) || !(opt->max_dist >= 0.f)) return setErr(ctx, B200REG_E_INVALID_PARAM, "SurfaceNormalDataPointsFilter: epsilon/maxDist must be >= 0");
  const int64_t n = inout->n;
  if (degenerate_count) *degenerate_count = 0;
  if (n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  // own matcher over the cloud itself (SurfaceNormal.cpp:170-179): un-centred index
  b200reg_index* ix = nullptr;
  TRY(buildIndex(ctx, inout, 0, 0.f, &ix));
  // the list kernel walks the cloud in the index's sorted order (null queries); the one-thread kernel takes it as it lies
  const float4* qs = (ctx->knnLanes == 1) ? inout->xyz : nullptr;
  int st = runKnn(ctx, ix, qs, n, nullptr, knn, opt->max_dist, 1, 0);
  if (st != B200REG_OK) { b200reg_index_free(ctx, ix); return st; }
  if (!inout->nrm) {
    cudaError_t e = devAlloc(ctx, &inout->nrm, sizeof(float4) * (size_t)inout->cap);
    if (e != cudaSuccess) { b200reg_index_free(ctx, ix); return setErr(ctx, B200REG_E_OOM, "surface_normals: %s", cudaGetErrorString(e)); }
  }
  // device staging of the optional outputs: [densities n | eigValues 3n | linPlanCurv 3n | eigVectors 9n | meanDists n | matchedIds knn n]
  const size_t extra = sizeof(float) * (size_t)n * (17 + (size_t)knn) + 64;
  st = ensureBuf(ctx, ctx->misc, extra);
  if (st != B200REG_OK) { b200reg_index_free(ctx, ix); return st; }
  float* base = (float*)ctx->misc.p;
  float* dDens = opt->densities ? base : nullptr;
  float* dEig = opt->eig_values ? base + n : nullptr;
  float* dLpc = opt->lin_plan_curv ? base + 4 * n : nullptr;
  float* dVec = opt->eig_vectors ? base + 7 * n : nullptr;
  float* dMd = opt->mean_dists ? base + 16 * n : nullptr;
  float* dMi = opt->matched_ids ? base + 17 * n : nullptr;
  int* dDeg = (int*)((char*)ctx->dstate + 32768);
  cudaMemsetAsync(dDeg, 0, sizeof(int), ctx->stream);
  const float* oo = opt->orient_origin;
  const float ox = oo ? oo[0] : 0.f, oy = oo ? oo[1] : 0.f, oz = oo ? oo[2] : 0.f;
  normalsKernel<<<divUp(n, 128), 128, 0, ctx->stream>>>(inout->xyz, ix->pts, (const int*)ctx->matchPos.p, (const float*)ctx->matchD2.p, n, knn,
                                                        oo ? 1 : 0, ox, oy, oz, inout->nrm, dDens, dEig, dLpc, dVec, dMd, (const int*)ctx->matchIds.p, dMi, dDeg);
  ctx->launches++;
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && opt->smooth_normals) {
    // old normals -> scratch, smoothed normals -> the cloud; flags + ticket in a scratch of their own
    e = (ensureBuf(ctx, ctx->keys1, sizeof(float4) * (size_t)n) == B200REG_OK && ensureBuf(ctx, ctx->vals1, sizeof(int) * (size_t)n + 64) == B200REG_OK)
            ? cudaSuccess : cudaErrorMemoryAllocation;
    if (e == cudaSuccess) e = cudaMemcpyAsync(ctx->keys1.p, inout->nrm, sizeof(float4) * (size_t)n, cudaMemcpyDeviceToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(ctx->vals1.p, 0, sizeof(int) * (size_t)n + 64, ctx->stream);
    if (e == cudaSuccess) {
      int* done = (int*)ctx->vals1.p;
      unsigned long long* ticket = (unsigned long long*)((char*)ctx->vals1.p + ((sizeof(int) * (size_t)n + 15) / 16) * 16);
      const int blocks = (int)fmin((double)(4 * ctx->sm_count), (double)divUp(n * 32, 256));   // resident grid: waiting warps must not starve running ones
      smoothNormalsKernel<<<blocks, 256, 0, ctx->stream>>>((const float4*)ctx->keys1.p, inout->nrm, (const int*)ctx->matchIds.p,
                                                           (const float*)ctx->matchD2.p, (long long)n, knn, ticket, done);
      ctx->launches++;
      e = cudaGetLastError();
    }
  }
  auto d2h = [&](float* host, const float* dev, size_t count) {
    if (e == cudaSuccess && host) e = cudaMemcpyAsync(host, dev, sizeof(float) * count, cudaMemcpyDeviceToHost, ctx->stream);
  };
  d2h(opt->densities, dDens, (size_t)n);
  d2h(opt->eig_values, dEig, 3 * (size_t)n);
  d2h(opt->lin_plan_curv, dLpc, 3 * (size_t)n);
  d2h(opt->eig_vectors, dVec, 9 * (size_t)n);
  d2h(opt->mean_dists, dMd, (size_t)n);
  d2h(opt->matched_ids, dMi, (size_t)knn * (size_t)n);
  if (e == cudaSuccess) e = cudaMemcpyAsync(ctx->pinned, dDeg, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  const bool indexOk = (e != cudaSuccess) || indexHeader(ctx, ix) != nullptr;   // non-finite coordinates: the header carries the verdict
  b200reg_index_free(ctx, ix);
  if (e != cudaSuccess) return setErr(ctx, B200REG_E_CUDA, "surface_normals: %s", cudaGetErrorString(e));
  if (!indexOk) return setErr(ctx, B200REG_E_INVALID_PARAM, "SurfaceNormalDataPointsFilter: non-finite coordinates in the cloud");
  if (degenerate_count) *degenerate_count = *(const int*)ctx->pinned;
  return B200REG_OK;
}

extern "C" int32_t b200reg_surface_normals(b200reg_ctx* ctx, b200reg_cloud* inout, int32_t knn, float max_dist, float epsilon,
                                           const float* orient_origin, float* densities_out, float* eig_values_out,
                                           float* lin_plan_curv_out, int32_t* degenerate_count) {
  b200reg_normals_options o;
  memset(&o, 0, sizeof(o));
  o.knn = knn; o.max_dist = max_dist; o.epsilon = epsilon; o.orient_origin = orient_origin;
  o.densities = densities_out; o.eig_values = eig_values_out; o.lin_plan_curv = lin_plan_curv_out;
  return b200reg_surface_normals_ex(ctx, inout, &o, degenerate_count);
}
