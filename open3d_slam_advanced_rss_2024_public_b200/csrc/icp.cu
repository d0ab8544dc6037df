// csrc/icp.cu — host driver of the ICP iterate loop (ICP.cpp:258-468) over the device kernels, the 6x6 solve
// (PointToPlane.cpp:185-368) and the TransformationCheckers (TransformationCheckersImpl.cpp:58-225).
#include <chrono>
#include <cmath>
#include <cstdlib>

#include "icp_kernels.cuh"

int launchKnn1Loop(b200reg_ctx* ctx, const IcpDevParams* P, const IcpDevState* S, unsigned long long* counters, int blocks);
int runKnn(b200reg_ctx* ctx, const b200reg_index* ix, const float4* queries, int64_t n, const float* T, int k, float maxDist, int allowSelf,
           int seeded);

namespace {

inline SelState* selState(b200reg_ctx* ctx) { return (SelState*)((char*)ctx->dstate + DSTATE_SEL_OFF); }
inline unsigned* histPtr(b200reg_ctx* ctx) { return (unsigned*)((char*)ctx->dstate + DSTATE_HIST_OFF); }
#define LOOP_PARAMS_OFF 64
#define LOOP_STATE_OFF 512
inline double* resultPtr(b200reg_ctx* ctx) { return (double*)((char*)ctx->dstate + DSTATE_RESULT_OFF); }

// Trimmed-distance selection (Matches::getDistsQuantile): three multi-block radix passes.  DP != nullptr: arguments from the
// loop block.  (A single-launch variant that kept the distances in the distributed shared memory of one 8-CTA cluster was
// measured at 19.6 us against 16 us for these three launches — same-bank shared atomics serialise at 32 cycles per warp and
// only 8 SMs work — and was dropped; profiles/r1_notes.md.)
int launchSelect(b200reg_ctx* ctx, const float* d2, int64_t count, float ratio, const IcpDevParams* DP) {
  SelState* st = selState(ctx);
  const int blocks = DP ? 2 * ctx->sm_count : (int)fmax(1.0, fmin((double)divUp(count, 256 * 4), (double)(2 * ctx->sm_count)));
  selectPassKernel<0><<<blocks, 256, 0, ctx->stream>>>(d2, count, ratio, st, histPtr(ctx), DP);
  LAUNCH_CHECK(ctx);
  selectPassKernel<1><<<blocks, 256, 0, ctx->stream>>>(d2, count, ratio, st, histPtr(ctx), DP);
  LAUNCH_CHECK(ctx);
  selectPassKernel<2><<<blocks, 256, 0, ctx->stream>>>(d2, count, ratio, st, histPtr(ctx), DP);
  LAUNCH_CHECK(ctx);
  return B200REG_OK;
}

struct HostStep {   // what one iteration brings back over PCIe: NUM_MOM doubles + the select state + the visit counter
  double mom[NUM_MOM];
  SelState sel;
  unsigned long long visits, far;   // dstate[0], dstate[1]: touched points, queries the 3x3x3 neighbourhood could not settle
};

// device part of one iteration: match -> select -> reduce -> async D2H into pinned memory (no sync here)
int enqueueStep(b200reg_ctx* ctx, const b200reg_index* ix, const float4* reading, const float4* readingNrm, int64_t n, const float* T,
                int k, float maxDist, float trimRatio, float maxAngle, int force2D, int seeded = 0) {
  if (!ix->nrm) return setErr(ctx, B200REG_E_INVALID_FIELD, "PointToPlaneErrorMinimizer: reference cloud has no 'normals' descriptor");
  TRY(runKnn(ctx, ix, reading, n, T, k, maxDist, 1, (seeded && k == 1) ? 1 : 0));
  const int64_t total = n * k;
  SelState* st = selState(ctx);
  const bool useTrim = trimRatio > 0.f;
  if (useTrim) TRY(launchSelect(ctx, (const float*)ctx->matchD2.p, total, trimRatio, nullptr));
  ReduceParams P;
  P.reading = reading;
  P.readingNrm = readingNrm;
  P.refPts = ix->pts;
  P.refNrm = ix->nrm;
  P.ids = (const int*)ctx->matchIds.p;
  P.d2 = (const float*)ctx->matchD2.p;
  P.pos = (const int*)ctx->matchPos.p;
  P.n = n;
  P.k = k;
  P.hasT = T ? 1 : 0;
  memset(&P.T, 0, sizeof(P.T));
  if (T) memcpy(P.T.m, T, sizeof(P.T.m));
  P.useTrim = useTrim ? 1 : 0;
  P.useAngle = (maxAngle >= 0.f) ? 1 : 0;
  P.cosAngle = cosf(maxAngle);
  P.force2D = force2D;
  P.w = nullptr;
  const int rblocks = (int)fmax(1.0, fmin((double)divUp(total, 2 * RED_THREADS), (double)((512 / RED_THREADS) * ctx->sm_count)));
  TRY(ensureBuf(ctx, ctx->partials, sizeof(double) * NUM_MOM * (size_t)rblocks));
  reduceKernel<<<rblocks, RED_THREADS, 0, ctx->stream>>>(P, st, (double*)ctx->partials.p, &st->ticket[3], resultPtr(ctx), nullptr, nullptr);
  LAUNCH_CHECK(ctx);
  HostStep* h = (HostStep*)ctx->pinned;
  CUDA_TRY(ctx, cudaMemcpyAsync(h->mom, resultPtr(ctx), sizeof(double) * NUM_MOM, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(&h->sel, st, sizeof(SelState), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(&h->visits, ctx->dstate, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
  return B200REG_OK;
}

__host__ __device__ inline void cross3(const double a[3], const double b[3], double o[3]) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}

// un-centred moments -> A = G G^T, b = -G h^T on the matched-mean-centred pairs (PointToPlane.cpp:263-268,107-156)
__host__ __device__ void momentsToSystem(const double* m, double A[36], double b[6], double meanR[3], double meanS[3]) {
  const double K = m[M_CNT];
  for (int d = 0; d < 3; ++d) { meanR[d] = m[M_R + d] / K; meanS[d] = m[M_S + d] / K; }
  double Suu[3][3] = {{m[M_UU], m[M_UU + 1], m[M_UU + 2]}, {m[M_UU + 1], m[M_UU + 3], m[M_UU + 4]}, {m[M_UU + 2], m[M_UU + 4], m[M_UU + 5]}};
  double Snn[3][3] = {{m[M_NN], m[M_NN + 1], m[M_NN + 2]}, {m[M_NN + 1], m[M_NN + 3], m[M_NN + 4]}, {m[M_NN + 2], m[M_NN + 4], m[M_NN + 5]}};
  double Sun[3][3];
  for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) Sun[r][c] = m[M_UN + 3 * r + c];
  const double Sue[3] = {m[M_UE], m[M_UE + 1], m[M_UE + 2]}, Sne[3] = {m[M_NE], m[M_NE + 1], m[M_NE + 2]};
  // X = [meanR]_x  (X v = meanR x v)
  const double X[3][3] = {{0, -meanR[2], meanR[1]}, {meanR[2], 0, -meanR[0]}, {-meanR[1], meanR[0], 0}};
  double XSnn[3][3], SunXt[3][3], XSnnXt[3][3];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      double a = 0, s = 0;
      for (int k = 0; k < 3; ++k) { a += X[r][k] * Snn[k][c]; s += Sun[r][k] * X[c][k]; }
      XSnn[r][c] = a;
      SunXt[r][c] = s;
    }
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      double a = 0;
      for (int k = 0; k < 3; ++k) a += XSnn[r][k] * X[c][k];
      XSnnXt[r][c] = a;
    }
  double d[3];
  for (int k = 0; k < 3; ++k) d[k] = meanR[k] - meanS[k];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      A[6 * r + c] = Suu[r][c] - SunXt[r][c] - SunXt[c][r] + XSnnXt[r][c];   // sum c c^T
      A[6 * r + 3 + c] = Sun[r][c] - XSnn[r][c];                             // sum c n^T
      A[6 * (3 + c) + r] = A[6 * r + 3 + c];
      A[6 * (3 + r) + 3 + c] = Snn[r][c];
    }
  double Sund[3], Snnd[3], XSne[3], XSnnd[3];
  for (int r = 0; r < 3; ++r) {
    Sund[r] = Sun[r][0] * d[0] + Sun[r][1] * d[1] + Sun[r][2] * d[2];
    Snnd[r] = Snn[r][0] * d[0] + Snn[r][1] * d[1] + Snn[r][2] * d[2];
  }
  cross3(meanR, Sne, XSne);
  cross3(meanR, Snnd, XSnnd);
  for (int r = 0; r < 3; ++r) {
    b[r] = -(Sue[r] - Sund[r] - XSne[r] + XSnnd[r]);
    b[3 + r] = -(Sne[r] - Snnd[r]);
  }
}

// ---- dense helpers for the host solve (fp64) -------------------------------------------------------------------
__host__ __device__ bool cholSolve(const double* A, const double* b, int n, double* x) {
  double L[36];
  for (int j = 0; j < n; ++j) {
    double d = A[j * n + j];
    for (int k = 0; k < j; ++k) d -= L[j * n + k] * L[j * n + k];
    if (!(d > 0.0)) return false;
    d = sqrt(d);
    L[j * n + j] = d;
    for (int i = j + 1; i < n; ++i) {
      double s = A[i * n + j];
      for (int k = 0; k < j; ++k) s -= L[i * n + k] * L[j * n + k];
      L[i * n + j] = s / d;
    }
  }
  double y[6];
  for (int i = 0; i < n; ++i) {
    double s = b[i];
    for (int k = 0; k < i; ++k) s -= L[i * n + k] * y[k];
    y[i] = s / L[i * n + i];
  }
  for (int i = n - 1; i >= 0; --i) {
    double s = y[i];
    for (int k = i + 1; k < n; ++k) s -= L[k * n + i] * x[k];
    x[i] = s / L[i * n + i];
  }
  return true;
}

// symmetric eigen-decomposition, cyclic Jacobi; eigenvalues on the diagonal of a, eigenvectors in columns of V
__host__ __device__ void jacobiEig(double* a, double* V, int n) {
  for (int i = 0; i < n * n; ++i) V[i] = 0.0;
  for (int i = 0; i < n; ++i) V[i * n + i] = 1.0;
  for (int sweep = 0; sweep < 80; ++sweep) {
    double off = 0;
    for (int p = 0; p < n; ++p) for (int q = p + 1; q < n; ++q) off += a[p * n + q] * a[p * n + q];
    if (off == 0.0) break;
    for (int p = 0; p < n; ++p)
      for (int q = p + 1; q < n; ++q) {
        const double apq = a[p * n + q];
        if (apq == 0.0) continue;
        const double theta = (a[q * n + q] - a[p * n + p]) / (2.0 * apq);
        const double t = ((theta >= 0) ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < n; ++k) { const double x = a[k * n + p], y = a[k * n + q]; a[k * n + p] = c * x - s * y; a[k * n + q] = s * x + c * y; }
        for (int k = 0; k < n; ++k) { const double x = a[p * n + k], y = a[q * n + k]; a[p * n + k] = c * x - s * y; a[q * n + k] = s * x + c * y; }
        for (int k = 0; k < n; ++k) { const double x = V[k * n + p], y = V[k * n + q]; V[k * n + p] = c * x - s * y; V[k * n + q] = s * x + c * y; }
      }
  }
}

// solvePossiblyUnderdeterminedLinearSystem (PointToPlane.cpp:185-238): the invertibility decision uses the
// reference's scalar (float) epsilon x size relative to the largest eigenvalue; singular -> minimum-norm solution.
//
// Fast decision (what makes the solve affordable for ONE device thread): for a symmetric positive definite A,
// lambda_min >= 1 / ||A^-1||_F and lambda_max <= trace(A).  If 1 / ||A^-1||_F > eps * n * trace(A) the eigenvalue test of
// the slow path would certainly report full rank, so the Cholesky solution it would return is returned directly; anything
// closer to singular (or not positive definite) takes the eigen-decomposition path.  Sizes are template parameters so
// the device compiler keeps the small matrices in registers.
template <int N>
__host__ __device__ __forceinline__ bool cholFactorT(const double* A, double* L, double* invDiag) {
#pragma unroll
  for (int j = 0; j < N; ++j) {
    double d = A[j * N + j];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= L[j * N + k] * L[j * N + k];
    if (!(d > 0.0)) return false;
    d = sqrt(d);
    L[j * N + j] = d;
    const double inv = 1.0 / d;   // one division per column; the column and both substitutions multiply by it
    invDiag[j] = inv;
#pragma unroll
    for (int i = j + 1; i < N; ++i) {
      double s = A[i * N + j];
#pragma unroll
      for (int k = 0; k < j; ++k) s -= L[i * N + k] * L[j * N + k];
      L[i * N + j] = s * inv;
    }
  }
  return true;
}
template <int N>
__host__ __device__ __forceinline__ void cholSolveT(const double* L, const double* invDiag, const double* b, double* x) {
  double y[N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double s = b[i];
#pragma unroll
    for (int k = 0; k < i; ++k) s -= L[i * N + k] * y[k];
    y[i] = s * invDiag[i];
  }
#pragma unroll
  for (int i = N - 1; i >= 0; --i) {
    double s = y[i];
#pragma unroll
    for (int k = i + 1; k < N; ++k) s -= L[k * N + i] * x[k];
    x[i] = s * invDiag[i];
  }
}
template <int N>
__host__ __device__ __forceinline__ bool certainlyFullRankT(const double* A, const double* L, const double* invDiag) {
  double Li[N * N];   // L^-1 (lower triangular); A^-1 = Li^T Li
#pragma unroll
  for (int c = 0; c < N; ++c)
#pragma unroll
    for (int r = c; r < N; ++r) {
      double s = (r == c) ? 1.0 : 0.0;
#pragma unroll
      for (int k = c; k < r; ++k) s -= L[r * N + k] * Li[k * N + c];
      Li[r * N + c] = s * invDiag[r];
    }
  double fro2 = 0.0, tr = 0.0;
#pragma unroll
  for (int r = 0; r < N; ++r) {
    tr += A[r * N + r];
#pragma unroll
    for (int c = 0; c <= r; ++c) {   // symmetric: off-diagonal entries counted twice
      double v = 0.0;
#pragma unroll
      for (int k = r; k < N; ++k) v += Li[k * N + r] * Li[k * N + c];
      fro2 += (c == r) ? v * v : 2.0 * (v * v);
    }
  }
  if (!(fro2 > 0.0) || !(tr > 0.0) || !(fro2 < 1.0e300)) return false;
  const double rhs = 4.0 * 1.1920928955078125e-7 * N * tr;   // factor 4: margin over the rounding of this estimate
  return 1.0 > rhs * rhs * fro2;
}

__host__ __device__ __noinline__ void solveSymSlow(const double* A, const double* b, int n, double* x) {
  double a[36], V[36];
  for (int i = 0; i < n * n; ++i) a[i] = A[i];
  jacobiEig(a, V, n);
  double lmax = 0;
  for (int i = 0; i < n; ++i) lmax = fmax(lmax, fabs(a[i * n + i]));
  const double thr = 1.1920928955078125e-7 * n * lmax;
  int rank = 0;
  for (int i = 0; i < n; ++i) rank += (fabs(a[i * n + i]) > thr);
  if (rank == n && cholSolve(A, b, n, x)) return;
  for (int i = 0; i < n; ++i) x[i] = 0.0;
  for (int i = 0; i < n; ++i) {
    const double l = a[i * n + i];
    if (!(fabs(l) > thr)) continue;
    double proj = 0;
    for (int k = 0; k < n; ++k) proj += V[k * n + i] * b[k];
    proj /= l;
    for (int k = 0; k < n; ++k) x[k] += V[k * n + i] * proj;
  }
}

template <int N>
__host__ __device__ __forceinline__ void solveSym(const double* A, const double* b, double* x) {
  double L[N * N], invDiag[N];
  if (cholFactorT<N>(A, L, invDiag) && certainlyFullRankT<N>(A, L, invDiag)) { cholSolveT<N>(L, invDiag, b, x); return; }
  solveSymSlow(A, b, N, x);
}

// ---- fp32 4x4 helpers (row-major), depth accumulated k = 0..3 ----------------------------------------------------
__host__ __device__ void mul4(const float* a, const float* b, float* r) {
  float t[16];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      float acc = a[4 * i] * b[j];
      acc = acc + a[4 * i + 1] * b[4 + j];
      acc = acc + a[4 * i + 2] * b[8 + j];
      acc = acc + a[4 * i + 3] * b[12 + j];
      t[4 * i + j] = acc;
    }
  for (int i = 0; i < 16; ++i) r[i] = t[i];
}
__host__ __device__ bool rigidOk(const float* T) {
  const float det = T[0] * (T[5] * T[10] - T[6] * T[9]) - T[1] * (T[4] * T[10] - T[6] * T[8]) + T[2] * (T[4] * T[9] - T[5] * T[8]);
  return !(fabsf(1.f - det) > 0.001f);
}
__host__ __device__ void quatOf(const float* T, float q[4]) {   // Eigen::Quaternion(Matrix3): (w, x, y, z)
  float t = T[0] + T[5] + T[10];
  if (t > 0.f) {
    t = sqrtf(t + 1.f);
    q[0] = 0.5f * t;
    t = 0.5f / t;
    q[1] = (T[9] - T[6]) * t; q[2] = (T[2] - T[8]) * t; q[3] = (T[4] - T[1]) * t;
  } else {
    int i = 0;
    if (T[5] > T[0]) i = 1;
    if (T[10] > T[5 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = sqrtf(T[5 * i] - T[5 * j] - T[5 * k] + 1.f);
    q[1 + i] = 0.5f * t;
    t = 0.5f / t;
    q[0] = (T[4 * k + j] - T[4 * j + k]) * t;
    q[1 + j] = (T[4 * j + i] + T[4 * i + j]) * t;
    q[1 + k] = (T[4 * k + i] + T[4 * i + k]) * t;
  }
}
__host__ __device__ float quatDist(const float a[4], const float b[4]) {   // QuaternionBase::angularDistance
  const float bw = b[0], bx = -b[1], by = -b[2], bz = -b[3];
  const float w = a[0] * bw - a[1] * bx - a[2] * by - a[3] * bz;
  const float x = a[0] * bx + a[1] * bw + a[2] * bz - a[3] * by;
  const float y = a[0] * by + a[2] * bw + a[3] * bx - a[1] * bz;
  const float z = a[0] * bz + a[3] * bw + a[1] * by - a[2] * bx;
  return 2.f * atan2f(sqrtf(x * x + y * y + z * z), fabsf(w));
}

// Counter / Differential / Bound checkers (TransformationCheckersImpl.cpp:58-225) on a POD state, shared by the host-driven
// loop and the device-resident one (icpTailKernel).
struct CheckerCfg {
  int maxIter, smoothLength;   // maxIter < 0: no Counter; smoothLength < 0: no Differential
  float minDiffRot, minDiffTrans, boundRot, boundTrans;   // boundRot < 0: no Bound
  int order[3];                // evaluation order (B200REG_CHECKER_*; 0 ends the list); all 0 = Differential, Bound, Counter
};
// no configuration may spin the device forever (a chain without checkers loops forever in the reference, ICP.cpp:399)
#define ICP_HARD_ITERATION_CAP 10000
__host__ __device__ void checkersPush(CheckerState& c, const float* T) {
  const int slot = c.count % CHK_HIST;
  quatOf(T, c.rot[slot]);
  c.tr[slot][0] = T[3]; c.tr[slot][1] = T[7]; c.tr[slot][2] = T[11];
  ++c.count;
}
__host__ __device__ void checkersInit(CheckerState& c, const float* T) {
  c.counter = 0;
  c.count = 0;
  checkersPush(c, T);
  quatOf(T, c.q0);
  c.t0[0] = T[3]; c.t0[1] = T[7]; c.t0[2] = T[11];
}
// returns ICPERR_*; may clear `iterate` / set `maxIter`; err[] = the values the reference puts in its message.  The checkers run
// in the chain's order (ICP.cpp:437-446): Counter's MaxNumIterationsReached is a throw, so the checkers listed after it are
// skipped on the iteration it fires (TransformationCheckersImpl.cpp:66-77).
__host__ __device__ int checkersCheck(const CheckerCfg& cfg, CheckerState& c, const float* T, bool& iterate, bool& maxIter, float err[4]) {
  const bool legacy = cfg.order[0] == 0;
  for (int slot = 0; slot < 3; ++slot) {
    const int kind = legacy ? (slot == 0 ? B200REG_CHECKER_DIFFERENTIAL : slot == 1 ? B200REG_CHECKER_BOUND : B200REG_CHECKER_COUNTER) : cfg.order[slot];
    if (kind == 0) break;
    if (kind == B200REG_CHECKER_DIFFERENTIAL && cfg.smoothLength >= 0) {
      checkersPush(c, T);
      const int sl = cfg.smoothLength, cnt = c.count;
      float cv0 = 0.f, cv1 = 0.f;
      if (cnt > sl) {
        for (int i = cnt - 1; sl > 0 && i >= cnt - sl; --i) {
          const int a = i % CHK_HIST, b = (i - 1) % CHK_HIST;
          cv0 += fabsf(quatDist(c.rot[a], c.rot[b]));
          const float dx = c.tr[a][0] - c.tr[b][0], dy = c.tr[a][1] - c.tr[b][1], dz = c.tr[a][2] - c.tr[b][2];
          cv1 += fabsf(sqrtf(dx * dx + dy * dy + dz * dz));
        }
        cv0 /= (float)sl; cv1 /= (float)sl;
        if (cv0 < cfg.minDiffRot && cv1 < cfg.minDiffTrans) iterate = false;
      }
      if (cv0 != cv0) return ICPERR_ROT_NAN;
      if (cv1 != cv1) return ICPERR_TRANS_NAN;
    } else if (kind == B200REG_CHECKER_BOUND && cfg.boundRot >= 0.f) {
      float q[4];
      quatOf(T, q);
      const float r = quatDist(q, c.q0);
      const float dx = T[3] - c.t0[0], dy = T[7] - c.t0[1], dz = T[11] - c.t0[2];
      const float t = sqrtf(dx * dx + dy * dy + dz * dz);
      if (r > cfg.boundRot || t > cfg.boundTrans) {
        err[0] = r; err[1] = cfg.boundRot; err[2] = t; err[3] = cfg.boundTrans;
        return ICPERR_BOUND;
      }
    } else if (kind == B200REG_CHECKER_COUNTER && cfg.maxIter >= 0) {
      ++c.counter;   // maxIterationCount 0 stops after the first iteration, like 1 (counter >= limit, :70)
      if (c.counter >= cfg.maxIter) { iterate = false; maxIter = true; return ICPERR_NONE; }   // MaxNumIterationsReached, caught at ICP.cpp:441-445
    }
  }
  return ICPERR_NONE;
}
int icpErrToStatus(b200reg_ctx* ctx, int kind, const float err[4]) {
  switch (kind) {
    case ICPERR_NONE: return B200REG_OK;
    case ICPERR_ROT_NAN: return setErr(ctx, B200REG_E_NOT_CONVERGED, "abs rotation norm not a number");
    case ICPERR_TRANS_NAN: return setErr(ctx, B200REG_E_NOT_CONVERGED, "abs translation norm not a number");
    case ICPERR_BOUND:
      return setErr(ctx, B200REG_E_NOT_CONVERGED, "limit out of bounds: rot: %g/%g tr: %g/%g", err[0], err[1], err[2], err[3]);
    case ICPERR_NO_QUANTILE: return setErr(ctx, B200REG_E_NOT_CONVERGED, "No matches available for computing distance quantiles");
    case ICPERR_NO_POINTS: return setErr(ctx, B200REG_E_NOT_CONVERGED, "ErrorMnimizer: no point to minimize");
    case ICPERR_NOT_RIGID: return setErr(ctx, B200REG_E_TRANSFORM, "RigidTransformation: Error, rotation matrix is not orthogonal.");
    case ICPERR_ITER_CAP: return setErr(ctx, B200REG_E_NOT_CONVERGED, "ICP: no transformation checker stopped the loop within %d iterations", ICP_HARD_ITERATION_CAP);
  }
  return setErr(ctx, B200REG_E_CUDA, "ICP: unknown device error %d", kind);
}

__host__ __device__ void solveP2P(const double* A, const double* b, const double* mean_reading, const double* mean_reference, int force2D,
                                  int force4DOF, float* T_out, double* x_out);

double nowSec() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

}  // namespace

struct b200reg_icp {
  b200reg_ctx* ctx = nullptr;
  b200reg_icp_config cfg{};
  b200reg_index* index = nullptr;   // over the centred, (un)filtered reference — `referenceFiltered` + the matcher
  bool ownsIndex = true;            // false: the index belongs to a b200reg_batch (shared by its lanes, read-only)
  double refPrepS = 0;
  DevBuf reading, readingNrm;       // `reading` of ICP.cpp:352-375 (in the refMean frame)
  // device-resident iterate loop (k = 1): a CUDA graph whose body is a `while` conditional node
  char* dBlock = nullptr;           // ONE device block: [visit / far counters, 64 B | IcpDevParams, 448 B | IcpDevState]: one H2D copy per registration
  IcpDevParams* dP = nullptr;
  IcpDevState* dS = nullptr;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  int graphState = 0;               // 0 not built yet, 1 ready, -1 unavailable (host-driven loop is used)
  int graphNodes = 0;               // kernels per iteration of the body
  // a registration enqueued on the ctx stream and not yet collected (icpEnqueue / icpFinish): the batch lanes keep one in flight each
  bool pending = false;
  float pendTinit[16];
  int64_t pendN = 0;
  double pendT1 = 0;
  b200reg_icp_stats pendStats{};
};

extern "C" {

int32_t b200reg_solve_point_to_plane(const double A[36], const double b[6], const double mean_reading[3],
                                     const double mean_reference[3], int32_t force2D, int32_t force4DOF, float T_out[16],
                                     double x_out[6]) {
  if (!A || !b || !mean_reading || !mean_reference || !T_out) return B200REG_E_INVALID_PARAM;
  solveP2P(A, b, mean_reading, mean_reference, force2D, force4DOF, T_out, x_out);
  return B200REG_OK;
}

}  // extern "C"

namespace {
__host__ __device__ void solveP2P(const double* A, const double* b, const double* mean_reading, const double* mean_reference, int force2D,
                                  int force4DOF, float* T_out, double* x_out) {
  double x[6] = {0, 0, 0, 0, 0, 0};
  double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, t[3] = {0, 0, 0};
  if (force2D) {   // unknowns [gamma, tx, ty] = rows/cols {2,3,4} (PointToPlane.cpp:247-254, 334-366)
    const int sel[3] = {2, 3, 4};
    double As[9], bs[3], xs[3];
    for (int r = 0; r < 3; ++r) { bs[r] = b[sel[r]]; for (int c = 0; c < 3; ++c) As[3 * r + c] = A[6 * sel[r] + sel[c]]; }
    solveSym<3>(As, bs, xs);
    x[0] = xs[0]; x[1] = xs[1]; x[2] = xs[2];
    R[0] = cos(xs[0]); R[1] = -sin(xs[0]); R[3] = sin(xs[0]); R[4] = cos(xs[0]);
    t[0] = xs[1]; t[1] = xs[2];
  } else {
    double axis[3] = {0, 0, 1}, angle;
    if (force4DOF) {   // unknowns [gamma, tx, ty, tz] = rows/cols {2,3,4,5} (PointToPlane.cpp:128-134,297-313)
      const int sel[4] = {2, 3, 4, 5};
      double As[16], bs[4], xs[4];
      for (int r = 0; r < 4; ++r) { bs[r] = b[sel[r]]; for (int c = 0; c < 4; ++c) As[4 * r + c] = A[6 * sel[r] + sel[c]]; }
      solveSym<4>(As, bs, xs);
      for (int i = 0; i < 4; ++i) x[i] = xs[i];
      angle = xs[0];
      t[0] = xs[1]; t[1] = xs[2]; t[2] = xs[3];
    } else {
      solveSym<6>(A, b, x);
      angle = sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
      const double wmax = fmax(fabs(x[0]), fmax(fabs(x[1]), fabs(x[2])));   // stableNormalized()
      double v[3] = {x[0], x[1], x[2]};
      if (wmax > 0.0) {
        const double z = (v[0] / wmax) * (v[0] / wmax) + (v[1] / wmax) * (v[1] / wmax) + (v[2] / wmax) * (v[2] / wmax);
        if (z > 0.0) for (int d = 0; d < 3; ++d) v[d] = v[d] / (sqrt(z) * wmax);
      }
      for (int d = 0; d < 3; ++d) axis[d] = v[d];
      t[0] = x[3]; t[1] = x[4]; t[2] = x[5];
    }
    const double s = sin(angle), c = cos(angle);   // AngleAxis::toRotationMatrix
    const double cx = (1 - c) * axis[0], cy = (1 - c) * axis[1], cz = (1 - c) * axis[2];
    const double sx = s * axis[0], sy = s * axis[1], sz = s * axis[2];
    double tmp;
    tmp = cx * axis[1]; R[1] = tmp - sz; R[3] = tmp + sz;
    tmp = cx * axis[2]; R[2] = tmp + sy; R[6] = tmp - sy;
    tmp = cy * axis[2]; R[5] = tmp - sx; R[7] = tmp + sx;
    R[0] = cx * axis[0] + c; R[4] = cy * axis[1] + c; R[8] = cz * axis[2] + c;
  }
  // T = Trans(meanRef) * [R t] * Trans(-meanRead)   (PointToPlane.cpp:316-324)
  const int sd = force2D ? 2 : 3;
  double T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  bool nan = false;
  for (int r = 0; r < 3; ++r) {
    double acc = 0;
    for (int c = 0; c < 3; ++c) { T[4 * r + c] = R[3 * r + c]; acc += R[3 * r + c] * (c < sd ? -mean_reading[c] : 0.0); }
    T[4 * r + 3] = acc + (t[r] + (r < sd ? mean_reference[r] : 0.0));
  }
  for (int i = 0; i < 16; ++i) nan |= (T[i] != T[i]);
  if (nan)   // :326-332 — rotation block replaced by identity
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) T[4 * r + c] = (r == c) ? 1.0 : 0.0;
  for (int i = 0; i < 16; ++i) T_out[i] = (float)T[i];
  if (x_out) for (int i = 0; i < 6; ++i) x_out[i] = x[i];
}
}  // namespace

extern "C" {

int32_t b200reg_icp_step(b200reg_ctx* ctx, const b200reg_index* index, const b200reg_cloud* reading, const float* T_iter,
                         int32_t k, float max_dist, float trim_ratio, float max_angle, int32_t force2D, double A_out[36],
                         double b_out[6], b200reg_step_stats* stats) {
  if (!ctx || !index || !reading) return setErr(ctx, B200REG_E_INVALID_PARAM, "icp_step: bad arguments");
  if (reading->n == 0) return setErr(ctx, B200REG_E_NO_POINTS, "The reading point cloud is empty.");
  if (trim_ratio > 1.0f) return setErr(ctx, B200REG_E_NOT_CONVERGED, "Distance quantile of matches must lie in the range [0,1]");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->dstate, 0, sizeof(unsigned long long), ctx->stream));
  TRY(enqueueStep(ctx, index, reading->xyz, reading->nrm, reading->n, T_iter, k, max_dist, trim_ratio, max_angle, force2D));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  const HostStep* h = (const HostStep*)ctx->pinned;
  const double total = (double)reading->n * k;
  if (stats) {
    memset(stats, 0, sizeof(*stats));
    stats->kept = (int64_t)h->mom[M_CNT];
    stats->n_finite = (int64_t)h->mom[M_VALID];
    stats->weight_sum = h->mom[M_CNT];
    stats->tau = (trim_ratio > 0.f) ? *(const float*)&h->sel.tauBits : INFINITY;
    stats->point_used_ratio = (float)h->mom[M_CNT] / (float)total;
    stats->weighted_point_used_ratio = (float)h->mom[M_CNT] / (float)total;
    stats->visit_count = h->visits;
  }
  if (trim_ratio > 0.f && h->sel.error) return setErr(ctx, B200REG_E_NOT_CONVERGED, "No matches available for computing distance quantiles");
  if (h->mom[M_CNT] < 1.0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "ErrorMnimizer: no point to minimize");
  double A[36], b[6], mr[3], ms[3];
  momentsToSystem(h->mom, A, b, mr, ms);
  if (A_out) memcpy(A_out, A, sizeof(A));
  if (b_out) memcpy(b_out, b, sizeof(b));
  if (stats) for (int d = 0; d < 3; ++d) { stats->mean_reading[d] = mr[d]; stats->mean_reference[d] = ms[d]; }
  return B200REG_OK;
}

}  // extern "C"

// =================================================================================================================
// b200reg_icp — PointMatcher<float>::ICP with the CUDA chain inside (ICP.cpp:258-468)
// =================================================================================================================
namespace {

// reading <- T_refMean_readMean * reading with T_refMean_readMean = T_refIn_refMean^-1 * T_init (ICP.cpp:363-375; meanReading is
// forced to 0 in this fork).  The reference mean is read from the index header the build left in HBM, so the host never waits
// for it; every thread forms the 4x4 product with the host's own mul4 (same rounded operations, this file is --fmad=false).
struct Mat16 { float m[16]; };
__global__ void prepReadingKernel(const float4* __restrict__ xyz, const float4* __restrict__ nrm, int64_t n, Mat16 Tinit,
                                  const IndexHeader* __restrict__ hdr, float4* __restrict__ oxyz, float4* __restrict__ onrm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float TrefInv[16] = {1, 0, 0, -hdr->mean[0], 0, 1, 0, -hdr->mean[1], 0, 0, 1, -hdr->mean[2], 0, 0, 0, 1};
  float Tpre[16];
  mul4(TrefInv, Tinit.m, Tpre);
  Rigid12 T;
#pragma unroll
  for (int j = 0; j < 12; ++j) T.m[j] = Tpre[j];
  const float4 p = xyz[i];
  const float3 r = xformPoint(T, p.x, p.y, p.z, p.w);
  const float w = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(Tpre[12], p.x), __fmul_rn(Tpre[13], p.y)), __fmul_rn(Tpre[14], p.z)), __fmul_rn(Tpre[15], p.w));
  oxyz[i] = make_float4(r.x, r.y, r.z, w);
  if (nrm) {
    const float4 q = nrm[i];
    const float3 s = xformNormal(T, q.x, q.y, q.z);
    onrm[i] = make_float4(s.x, s.y, s.z, 0.f);
  }
}

}  // namespace

namespace {

// Everything the host-driven loop does between two iterations (ICP.cpp:418-446), by one device thread: ErrorElements
// checks, A/b from the moments, the 6x6 solve, T_iter <- T_step * T_iter, the transformation checkers, and the loop
// condition of the graph's while node.  Same __host__ __device__ arithmetic as the host loop (this file is built with
// --fmad=false), so both loops take the same decisions.
__global__ void __launch_bounds__(32) icpTailKernel(const IcpDevParams* __restrict__ P, IcpDevState* __restrict__ S,
                                                    const double* __restrict__ result, const SelState* __restrict__ st,
                                                    unsigned long long* __restrict__ visitCounter,
                                                    cudaGraphConditionalHandle handle) {
  if (threadIdx.x != 0) return;
  traceStamp(P->trace, 6);
  // every global read up front and independent of the others: one memory round trip instead of a chain of them
  const IcpDevParams p = *P;
  double mom[NUM_MOM];
#pragma unroll
  for (int j = 0; j < NUM_MOM; ++j) mom[j] = __ldcg(result + j);
  const int selError = __ldcg(&st->error);
  const unsigned tauBits = __ldcg(&st->tauBits);
  const unsigned long long visits = __ldcg(visitCounter);
  const unsigned long long farNow = __ldcg(visitCounter + 1);
  const unsigned long long farPrev = S->farPrev;
  float Told[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) Told[i] = S->T[i];
  int err = ICPERR_NONE;
  bool iterate = true, maxIter = false;
  float errv[4] = {0.f, 0.f, 0.f, 0.f};
  if (p.useTrim && selError) err = ICPERR_NO_QUANTILE;
  else if (mom[M_CNT] < 1.0) err = ICPERR_NO_POINTS;
  else {
    double A[36], b[6], mr[3], ms[3];
    momentsToSystem(mom, A, b, mr, ms);
    float Tstep[16], Tnew[16];
    solveP2P(A, b, mr, ms, p.force2D, p.force4DOF, Tstep, nullptr);
    mul4(Tstep, Told, Tnew);   // T_iter = T_step * T_iter (ICP.cpp:433)
#pragma unroll
    for (int i = 0; i < 16; ++i) S->T[i] = Tnew[i];
    const CheckerCfg ccfg{p.maxIter, p.smoothLength, p.minDiffRot, p.minDiffTrans, p.boundRot, p.boundTrans,
                          {p.checkerOrder[0], p.checkerOrder[1], p.checkerOrder[2]}};
    err = checkersCheck(ccfg, S->chk, Tnew, iterate, maxIter, errv);
    const int its = S->iterations + 1;
    S->iterations = its;
    if (err == ICPERR_NONE && iterate && its >= ICP_HARD_ITERATION_CAP) err = ICPERR_ITER_CAP;
    if (err == ICPERR_NONE && iterate && !rigidOk(Tnew)) err = ICPERR_NOT_RIGID;   // the next iteration's RigidTransformation check
  }
  if (err != ICPERR_NONE) iterate = false;
  S->kept = mom[M_CNT];
  S->valid = mom[M_VALID];
  S->tau = p.useTrim ? __uint_as_float(tauBits) : CUDART_INF_F;
  S->visits = visits;
  S->farPrev = farNow;
  visitCounter[2] = 0ull;   // the match kernel's work counter (knn1LoopBody), for the next iteration
  S->lanes = p.lanesForced ? p.lanesForced : (((farNow - farPrev) * 16ull < (unsigned long long)p.n) ? 4 : 8);
  S->errKind = err;
  S->errA = errv[0]; S->errB = errv[1]; S->errC = errv[2]; S->errD = errv[3];
  S->maxIterReached = maxIter ? 1 : 0;
  S->iterate = iterate ? 1 : 0;
  cudaGraphSetConditional(handle, iterate ? 1u : 0u);
  traceStamp(P->trace, 7);   // end of the tail
}

#define GRAPH_TRY(expr)                                                    \
  do {                                                                     \
    cudaError_t _e = (expr);                                               \
    if (_e != cudaSuccess) { why = #expr; code = _e; goto fail; }          \
  } while (0)

// Build the while-graph once per ICP object.  Every kernel of the body takes its arguments from icp->dP / icp->dS and
// runs on a fixed grid (a multiple of the SM count), so the graph never has to be updated or re-instantiated.
int buildLoopGraph(b200reg_icp* icp) {
  b200reg_ctx* ctx = icp->ctx;
  const char* why = "";
  cudaError_t code = cudaSuccess;
  cudaGraph_t g = nullptr, body = nullptr, captured = nullptr;
  cudaGraphExec_t exec = nullptr;
  bool capturing = false;
  const int debugSync = ctx->debugSync;
  ctx->debugSync = 0;   // no synchronisation inside a stream capture
  struct Restore { b200reg_ctx* c; int v; ~Restore() { c->debugSync = v; } } restore{ctx, debugSync};
  const bool useTrim = icp->cfg.trim_ratio > 0.f;
  SelState* st = selState(ctx);
  // the graph launches the match kernel on the full resident grid; IcpDevParams::matchBlocks (icpEnqueue) says how many of its blocks work
  int perSm = KNN1_LOOP_MINB;
  if (const char* e = getenv("B200REG_KNN_GRID_PER_SM")) { const int v = atoi(e); if (v >= 1 && v <= KNN1_LOOP_MINB) perSm = v; }   // A/B
  const int knnBlocks = perSm * ctx->sm_count /* the resident grid of knn1LoopKernel: its warps pull work from a counter */, redBlocks = ((getenv("B200REG_RED_GRID_PER_SM") && atoi(getenv("B200REG_RED_GRID_PER_SM")) == 2) ? 2 : 1) * ctx->sm_count;   /* one 256-thread block per SM: fewer 32 K-register blocks to place between the lanes' match kernels (3667 -> 3805 registrations/s; env = 2: the old grid, A/B) */
  if (ensureBuf(ctx, ctx->partials, sizeof(double) * NUM_MOM * (size_t)redBlocks) != B200REG_OK) return -1;
  {
    cudaGraphConditionalHandle handle;
    cudaGraphNodeParams np = {cudaGraphNodeTypeConditional};
    cudaGraphNode_t node;
    GRAPH_TRY(cudaGraphCreate(&g, 0));
    GRAPH_TRY(cudaGraphConditionalHandleCreate(&handle, g, 1, cudaGraphCondAssignDefault));
    np.type = cudaGraphNodeTypeConditional;
    np.conditional.handle = handle;
    np.conditional.type = cudaGraphCondTypeWhile;
    np.conditional.size = 1;
    GRAPH_TRY(cudaGraphAddNode(&node, g, nullptr, 0, &np));
    body = np.conditional.phGraph_out[0];
    GRAPH_TRY(cudaStreamBeginCaptureToGraph(ctx->stream, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
    capturing = true;
    int nodes = 0;
    launchKnn1Loop(ctx, icp->dP, icp->dS, (unsigned long long*)icp->dBlock, knnBlocks);
    ++nodes;
    const uint64_t l0 = ctx->launches;
    if (useTrim && !ctx->select3) {
      int selBlocks = ctx->sm_count;
      if (const char* e = getenv("B200REG_SEL_BLOCKS")) { const int v = atoi(e); if (v >= 1 && v <= 4 * ctx->sm_count) selBlocks = v; }   // A/B
      selectFusedKernel<<<selBlocks, SELF_THREADS, 0, ctx->stream>>>(icp->dP, st);
      LAUNCH_CHECK(ctx);
    } else if (useTrim && launchSelect(ctx, nullptr, 0, 0.f, icp->dP) != B200REG_OK) { why = "select launch"; code = cudaErrorUnknown; goto fail; }
    nodes += (int)(ctx->launches - l0);
    ctx->launches = l0;
    ReduceParams RP;
    memset(&RP, 0, sizeof(RP));
    reduceKernel<<<redBlocks, RED_THREADS, 0, ctx->stream>>>(RP, st, nullptr, &st->ticket[3], resultPtr(ctx), icp->dP, icp->dS);
    icpTailKernel<<<1, 32, 0, ctx->stream>>>(icp->dP, icp->dS, resultPtr(ctx), st, (unsigned long long*)icp->dBlock, handle);
    nodes += 2;
    GRAPH_TRY(cudaGetLastError());
    capturing = false;
    GRAPH_TRY(cudaStreamEndCapture(ctx->stream, &captured));
    GRAPH_TRY(cudaGraphInstantiate(&exec, g, 0));
    icp->graph = g;
    icp->exec = exec;
    icp->graphNodes = nodes;
    return 0;
  }
fail:
  if (capturing) cudaStreamEndCapture(ctx->stream, &captured);
  if (g) cudaGraphDestroy(g);
  cudaGetLastError();
  if (getenv("B200REG_VERBOSE")) fprintf(stderr, "b200reg: device-resident loop unavailable (%s: %s); using the host-driven loop\n", why, cudaGetErrorString(code));
  return -1;
}

}  // namespace

extern "C" {

int32_t b200reg_icp_create(b200reg_ctx* ctx, const b200reg_icp_config* cfg, b200reg_icp** out) {
  if (!ctx || !cfg || !out) return setErr(ctx, B200REG_E_INVALID_PARAM, "icp_create: bad arguments");
  // parameter bounds of the CPU twins (Parametrizable.cpp:183-186 -> InvalidParameter)
  if (cfg->knn < 1) return setErr(ctx, B200REG_E_INVALID_PARAM, "KDTreeMatcher: knn must be >= 1");                    // MatchersImpl.h:97
  if (!(cfg->epsilon >= 0.f)) return setErr(ctx, B200REG_E_INVALID_PARAM, "KDTreeMatcher: epsilon must be >= 0");       // :98
  if (!(cfg->max_dist >= 0.f)) return setErr(ctx, B200REG_E_INVALID_PARAM, "KDTreeMatcher: maxDist must be >= 0");      // :100
  if (cfg->trim_ratio > 1.0f) return setErr(ctx, B200REG_E_INVALID_PARAM, "TrimmedDistOutlierFilter: ratio must be in [1e-7, 1]");   // OutlierFiltersImpl.h:137
  if (cfg->max_angle > 3.1416f) return setErr(ctx, B200REG_E_INVALID_PARAM, "SurfaceNormalOutlierFilter: maxAngle must be in [0, 3.1416]");
  if (cfg->smooth_length > CHK_HIST - 1)
    return setErr(ctx, B200REG_E_INVALID_PARAM, "DifferentialTransformationChecker: smoothLength above %d is not supported", CHK_HIST - 1);
  b200reg_icp* icp = new b200reg_icp();
  icp->ctx = ctx;
  icp->cfg = *cfg;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  static_assert(sizeof(IcpDevParams) <= LOOP_STATE_OFF - LOOP_PARAMS_OFF, "loop block layout");
  if (devAlloc(icp->ctx, &icp->dBlock, LOOP_STATE_OFF + sizeof(IcpDevState)) != cudaSuccess) {
    delete icp;
    return setErr(ctx, B200REG_E_OOM, "icp_create: allocation failed");
  }
  icp->dP = (IcpDevParams*)(icp->dBlock + LOOP_PARAMS_OFF);
  icp->dS = (IcpDevState*)(icp->dBlock + LOOP_STATE_OFF);
  *out = icp;
  return B200REG_OK;
}

void b200reg_icp_destroy(b200reg_icp* icp) {
  if (!icp) return;
  b200reg_ctx* ctx = icp->ctx;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (icp->index && icp->ownsIndex) b200reg_index_free(ctx, icp->index);
  if (icp->reading.p) devFree(icp->ctx, icp->reading.p);
  if (icp->readingNrm.p) devFree(icp->ctx, icp->readingNrm.p);
  if (icp->exec) cudaGraphExecDestroy(icp->exec);
  if (icp->graph) cudaGraphDestroy(icp->graph);
  if (icp->dBlock) devFree(icp->ctx, icp->dBlock);
  delete icp;
}

// ICP::initReference (ICP.cpp:293-328): mean of the reference, centring, matcher->init.  Enqueue only: mean, bbox and grid are
// derived on the device (index.cu), the host never waits.  A reference with non-finite coordinates is reported by the first call
// that needs the result (b200reg_icp_register, b200reg_icp_reference_mean, b200reg_icp_index).
int32_t b200reg_icp_set_reference(b200reg_icp* icp, const b200reg_cloud* reference) {
  if (!icp || !reference) return B200REG_E_INVALID_PARAM;
  b200reg_ctx* ctx = icp->ctx;
  const double t0 = nowSec();
  if (icp->pending) return setErr(ctx, B200REG_E_INVALID_PARAM, "icp_set_reference: a registration is still in flight on this object");
  b200reg_index* old = (icp->index && icp->ownsIndex) ? icp->index : nullptr;   // matcherIsInitialized = false
  icp->index = nullptr;
  icp->ownsIndex = true;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (reference->n == 0) {   // :295-298
    if (old) b200reg_index_free(ctx, old);
    return setErr(ctx, B200REG_E_NO_POINTS, "initReference: reference cloud is empty");
  }
  TRY(buildIndex(ctx, reference, 1, icp->cfg.cell, &icp->index, old));   // the previous index's storage is reused when it fits
  icp->refPrepS = nowSec() - t0;
  return B200REG_OK;
}

const b200reg_index* b200reg_icp_index(const b200reg_icp* icp) { return icp ? icp->index : nullptr; }

int32_t b200reg_icp_reference_mean(const b200reg_icp* icp, float mean_out[3]) {
  if (!icp || !mean_out) return B200REG_E_INVALID_PARAM;
  for (int d = 0; d < 3; ++d) mean_out[d] = 0.f;
  if (!icp->index) return B200REG_OK;
  const IndexHeader* h = indexHeader(icp->ctx, icp->index);
  if (!h) return B200REG_E_INVALID_PARAM;
  for (int d = 0; d < 3; ++d) mean_out[d] = h->mean[d];
  return B200REG_OK;
}

}  // extern "C"

namespace {

// T_refMean_readMean = T_refIn_refMean^-1 * T_init (meanReading forced to 0 in this fork, ICP.cpp:363-373), on the host
void preTransform(const float mean[3], const float T_init[16], float Tpre[16]) {
  const float TrefInv[16] = {1, 0, 0, -mean[0], 0, 1, 0, -mean[1], 0, 0, 1, -mean[2], 0, 0, 0, 1};
  mul4(TrefInv, T_init, Tpre);
}

// RigidTransformation::checkParameters on T_refMean_readMean (ICP.cpp:375).  With a last row of (0 0 0 1) its rotation block is
// T_init's, entry for entry, so the usual case needs no reference mean — and therefore no wait for the index build.
int checkPreTransform(b200reg_icp* icp, const float T_init[16]) {
  b200reg_ctx* ctx = icp->ctx;
  bool ok;
  if (T_init[12] == 0.f && T_init[13] == 0.f && T_init[14] == 0.f && T_init[15] == 1.f) ok = rigidOk(T_init);
  else {
    const IndexHeader* h = indexHeader(ctx, icp->index);
    if (!h) return B200REG_E_INVALID_PARAM;
    float Tpre[16];
    preTransform(h->mean, T_init, Tpre);
    ok = rigidOk(Tpre);
  }
  return ok ? B200REG_OK : setErr(ctx, B200REG_E_TRANSFORM, "RigidTransformation: Error, rotation matrix is not orthogonal.");
}

int enqueuePrepReading(b200reg_icp* icp, const b200reg_cloud* reading, const float T_init[16], bool needNrm) {
  b200reg_ctx* ctx = icp->ctx;
  const int64_t n = reading->n;
  TRY(ensureBuf(ctx, icp->reading, sizeof(float4) * (size_t)n));
  if (needNrm) TRY(ensureBuf(ctx, icp->readingNrm, sizeof(float4) * (size_t)n));
  Mat16 Ti;
  memcpy(Ti.m, T_init, sizeof(Ti.m));
  prepReadingKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(reading->xyz, needNrm ? reading->nrm : nullptr, n, Ti, icp->index->dHdr,
                                                            (float4*)icp->reading.p, (float4*)icp->readingNrm.p);
  LAUNCH_CHECK(ctx);
  return B200REG_OK;
}

// icpCorrected_T_refIn_readIn = T_refIn_refMean * (T_iter * T_refMean_readMean) * I   (ICP.cpp:459-465)
int composeResult(b200reg_icp* icp, const float T_iter[16], const float T_init[16], float T_out[16]) {
  const IndexHeader* h = indexHeader(icp->ctx, icp->index);
  if (!h) return B200REG_E_INVALID_PARAM;
  float Tpre[16], tmp[16];
  preTransform(h->mean, T_init, Tpre);
  const float Tref[16] = {1, 0, 0, h->mean[0], 0, 1, 0, h->mean[1], 0, 0, 1, h->mean[2], 0, 0, 0, 1};
  mul4(T_iter, Tpre, tmp);
  mul4(Tref, tmp, T_out);
  return B200REG_OK;
}

inline bool wantGraphLoop(const b200reg_icp* icp) { return icp->cfg.knn == 1 && !icp->ctx->profile && !icp->ctx->hostLoop; }

// First half of a registration through the device-resident loop: everything up to the read-back of the final state is enqueued
// on the ctx stream; nothing waits.  The host image of the loop block and the result slot live in the ctx's pinned staging, so
// ONE registration per ctx can be in flight (the batch lanes own a ctx each).
int icpEnqueue(b200reg_icp* icp, const b200reg_cloud* reading, const float T_init[16]) {
  b200reg_ctx* ctx = icp->ctx;
  const b200reg_icp_config& cfg = icp->cfg;
  b200reg_icp_stats& st = icp->pendStats;
  memset(&st, 0, sizeof(st));
  if (icp->pending) return setErr(ctx, B200REG_E_INVALID_PARAM, "icp_register: a registration is still in flight on this object");
  if (!icp->index) return setErr(ctx, B200REG_E_NO_POINTS, "ICP: matcher not initialised (empty reference)");
  if (!icp->index->nrm) return setErr(ctx, B200REG_E_INVALID_FIELD, "PointToPlaneErrorMinimizer: reference cloud has no 'normals' descriptor");
  const int64_t n = reading->n;
  st.reference_point_count = icp->index->m;
  st.reading_point_count = n;
  st.reference_preprocessing_s = icp->refPrepS;
  if (n == 0) return setErr(ctx, B200REG_E_NO_POINTS, "The reading point cloud is empty.");   // ICP.cpp:357-359
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const double t0 = nowSec();
  TRY(checkPreTransform(icp, T_init));
  const bool needNrm = reading->nrm && cfg.max_angle >= 0.f;
  TRY(enqueuePrepReading(icp, reading, T_init, needNrm));
  st.reading_preprocessing_s = nowSec() - t0;
  icp->pendT1 = nowSec();
  const float T_iter[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  TRY(ensureBuf(ctx, ctx->matchIds, sizeof(int) * (size_t)n));
  TRY(ensureBuf(ctx, ctx->matchD2, sizeof(float) * (size_t)n));
  TRY(ensureBuf(ctx, ctx->matchPos, sizeof(int) * (size_t)n));
  char* pin = (char*)ctx->pinned;
  char* hBlock = pin + 8192;   // host image of the loop block
  IcpDevParams* hP = (IcpDevParams*)(hBlock + LOOP_PARAMS_OFF);
  IcpDevState* hS = (IcpDevState*)(hBlock + LOOP_STATE_OFF);
  IcpDevState* hR = (IcpDevState*)(pin + 16384);
  memset(hBlock, 0, LOOP_PARAMS_OFF);   // visit / far counters
  memset(hP, 0, sizeof(*hP));
  hP->pts = icp->index->pts; hP->nrm = icp->index->nrm; hP->cellStart = icp->index->cellStart; hP->gp = &icp->index->dHdr->g;
  hP->reading = (const float4*)icp->reading.p;
  hP->readingNrm = needNrm ? (const float4*)icp->readingNrm.p : nullptr;
  hP->n = n;
  hP->ids = (int*)ctx->matchIds.p; hP->d2 = (float*)ctx->matchD2.p; hP->pos = (int*)ctx->matchPos.p;
  hP->partials = (double*)ctx->partials.p;
  hP->maxd2 = cfg.max_dist * cfg.max_dist;
  hP->trimRatio = cfg.trim_ratio;
  hP->cosAngle = cosf(cfg.max_angle);
  hP->useTrim = cfg.trim_ratio > 0.f ? 1 : 0;
  hP->useAngle = cfg.max_angle >= 0.f ? 1 : 0;
  hP->force2D = cfg.force2D; hP->force4DOF = cfg.force4DOF;
  hP->maxIter = cfg.max_iter; hP->smoothLength = cfg.smooth_length;
  hP->minDiffRot = cfg.min_diff_rot; hP->minDiffTrans = cfg.min_diff_trans;
  hP->boundRot = cfg.bound_rot; hP->boundTrans = cfg.bound_trans;
  for (int j = 0; j < 3; ++j) hP->checkerOrder[j] = cfg.checker_order[j];
  if (cfg.trim_ratio > 0.f && !ctx->select3) {
    TRY(ensureBuf(ctx, ctx->cand, sizeof(unsigned) * (size_t)n));
    hP->fhist0 = (unsigned*)((char*)ctx->dstate + DSTATE_FH0_OFF);
    hP->fhist1 = (unsigned*)((char*)ctx->dstate + DSTATE_FH1_OFF);
    hP->fcnt = (unsigned*)((char*)ctx->dstate + DSTATE_FCNT_OFF);
    hP->cand = (unsigned*)ctx->cand.p;
  }
  // Blocks of the resident match grid that take work.  A registration on its own uses them all.  The lanes of a batch (loopGridPerSm = 2)
  // take HALF the machine each when the reference does not fit the L2 (32 B per point): those match kernels wait on DRAM, and two of them
  // side by side — plus room for the other lanes' small kernels — beat one after the other (config #2, 5 M points: 3450 -> 3555
  // registrations/s); an L2-resident reference is searched faster by the full grid (config #4, 200 k points: 1139 vs 1109 pairs/s).
  int perSm = KNN1_LOOP_MINB;
  if (ctx->loopGridPerSm > 0 && ctx->loopGridPerSm < perSm && (size_t)icp->index->m * 32 > ctx->l2Bytes) perSm = ctx->loopGridPerSm;
  hP->matchBlocks = perSm * ctx->sm_count;
  hP->lanesForced = (ctx->knnLanes == 4 || ctx->knnLanes == 8) ? ctx->knnLanes : 0;
  hP->seedOn = ctx->knnSeed;
  hP->trace = ctx->trace ? (unsigned long long*)((char*)ctx->dstate + DSTATE_TRACE_OFF) : nullptr;
  if (ctx->trace) CUDA_TRY(ctx, cudaMemsetAsync((char*)ctx->dstate + DSTATE_TRACE_OFF, 0, sizeof(unsigned long long), ctx->stream));
  memset(hS, 0, sizeof(*hS));
  memcpy(hS->T, T_iter, sizeof(T_iter));
  hS->iterate = 1;
  hS->lanes = (ctx->knnLanes == 4) ? 4 : 8;
  checkersInit(hS->chk, T_iter);
  // counters + parameters + the part of the state the loop reads before writing it (head, checker counters, first ring slot)
  CUDA_TRY(ctx, cudaMemcpyAsync(icp->dBlock, hBlock, LOOP_STATE_OFF + sizeof(IcpDevState), cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  CUDA_TRY(ctx, cudaGraphLaunch(icp->exec, ctx->stream));
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(hR, icp->dS, ICP_STATE_HEAD_BYTES, cudaMemcpyDeviceToHost, ctx->stream));
  memcpy(icp->pendTinit, T_init, sizeof(icp->pendTinit));
  icp->pendN = n;
  icp->pending = true;
  return B200REG_OK;
}

// Second half: wait for the stream, read the 144-byte head of the final state, compose the result.
int icpFinish(b200reg_icp* icp, float T_out[16], b200reg_icp_stats* stats) {
  b200reg_ctx* ctx = icp->ctx;
  b200reg_icp_stats& st = icp->pendStats;
  auto done = [&](int code) { if (stats) *stats = st; return code; };
  if (!icp->pending) return done(setErr(ctx, B200REG_E_INVALID_PARAM, "icp_register: nothing in flight"));
  icp->pending = false;
  const cudaError_t e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) return done(setErr(ctx, B200REG_E_CUDA, "ICP loop failed: %s", cudaGetErrorString(e)));
  const IcpDevState* hR = (const IcpDevState*)((char*)ctx->pinned + 16384);
  const double total = (double)icp->pendN * icp->cfg.knn;
  float gms = 0;
  cudaEventElapsedTime(&gms, ctx->ev0, ctx->ev1);
  st.device_ms = gms;
  st.iterations = hR->iterations;
  st.max_iterations_reached = hR->maxIterReached;
  st.points_touched = hR->visits;
  st.kept_last = (int64_t)hR->kept;
  st.point_used_ratio = (float)hR->kept / (float)total;
  st.overlap_ratio = st.point_used_ratio;
  st.last_tau = hR->tau;
  st.convergence_s = nowSec() - icp->pendT1;
  // kernels of the graph: (iterations that ran the full body) x nodes; an iteration that failed early still ran them all
  ctx->launches += (uint64_t)icp->graphNodes * (uint64_t)(hR->iterations + ((hR->errKind == ICPERR_NO_QUANTILE || hR->errKind == ICPERR_NO_POINTS) ? 1 : 0));
  if (!indexHeader(ctx, icp->index)) return done(B200REG_E_INVALID_PARAM);   // non-finite reference: the loop ran on a dummy grid
  const float errv[4] = {hR->errA, hR->errB, hR->errC, hR->errD};
  const int gstatus = icpErrToStatus(ctx, hR->errKind, errv);
  if (gstatus != B200REG_OK) return done(gstatus);
  return done(composeResult(icp, hR->T, icp->pendTinit, T_out));
}

}  // namespace

extern "C" {

int32_t b200reg_icp_register(b200reg_icp* icp, const b200reg_cloud* reading, const float T_init[16], float T_out[16],
                             b200reg_icp_stats* stats) {
  if (!icp || !reading || !T_init || !T_out) return B200REG_E_INVALID_PARAM;
  b200reg_ctx* ctx = icp->ctx;
  const b200reg_icp_config& cfg = icp->cfg;
  // ---- device-resident loop: the whole iterate loop is ONE graph launch; the host comes back for the final state only
  if (wantGraphLoop(icp) && icp->graphState == 0) {
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    icp->graphState = (buildLoopGraph(icp) == 0) ? 1 : -1;
  }
  if (wantGraphLoop(icp) && icp->graphState == 1) {
    const int code = icpEnqueue(icp, reading, T_init);
    if (code != B200REG_OK) { if (stats) *stats = icp->pendStats; return code; }
    return icpFinish(icp, T_out, stats);
  }
  // ---- host-driven loop (k > 1 chains, profiling, A/B): one synchronisation per iteration
  b200reg_icp_stats st;
  memset(&st, 0, sizeof(st));
  auto done = [&](int code) { if (stats) *stats = st; return code; };
  if (icp->pending) return done(setErr(ctx, B200REG_E_INVALID_PARAM, "icp_register: a registration is still in flight on this object"));
  if (!icp->index) return done(setErr(ctx, B200REG_E_NO_POINTS, "ICP: matcher not initialised (empty reference)"));
  if (!icp->index->nrm) return done(setErr(ctx, B200REG_E_INVALID_FIELD, "PointToPlaneErrorMinimizer: reference cloud has no 'normals' descriptor"));
  const int64_t n = reading->n;
  st.reference_point_count = icp->index->m;
  st.reading_point_count = n;
  st.reference_preprocessing_s = icp->refPrepS;
  if (n == 0) return done(setErr(ctx, B200REG_E_NO_POINTS, "The reading point cloud is empty."));   // ICP.cpp:357-359
#define REG_TRY(expr)                                                                                                      \
  do {                                                                                                                     \
    const cudaError_t _e = (expr);                                                                                         \
    if (_e != cudaSuccess) return done(setErr(ctx, B200REG_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__)); \
  } while (0)
  REG_TRY(cudaSetDevice(ctx->device));   // (every failure leaves through done(): *stats is always written)
  const double t0 = nowSec();
  if (!indexHeader(ctx, icp->index)) return done(B200REG_E_INVALID_PARAM);
  { const int c = checkPreTransform(icp, T_init); if (c != B200REG_OK) return done(c); }
  const bool needNrm = reading->nrm && cfg.max_angle >= 0.f;
  { const int c = enqueuePrepReading(icp, reading, T_init, needNrm); if (c != B200REG_OK) return done(c); }
  st.reading_preprocessing_s = nowSec() - t0;

  const double t1 = nowSec();
  float T_iter[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  bool iterate = true, maxIter = false;
  const double total = (double)n * cfg.knn;
  const CheckerCfg ccfg{cfg.max_iter, cfg.smooth_length, cfg.min_diff_rot, cfg.min_diff_trans, cfg.bound_rot, cfg.bound_trans,
                        {cfg.checker_order[0], cfg.checker_order[1], cfg.checker_order[2]}};
  CheckerState chk;
  checkersInit(chk, T_iter);
  REG_TRY(cudaMemsetAsync(ctx->dstate, 0, 2 * sizeof(unsigned long long), ctx->stream));
  REG_TRY(cudaEventRecord(ctx->ev0, ctx->stream));
#undef REG_TRY
  int status = B200REG_OK;
  const HostStep* h = (const HostStep*)ctx->pinned;
  unsigned long long farPrev = 0;
  const int lanesCfg = ctx->knnLanes;
  int lanesAuto = 8;   // the tail kernel's rule (icpTailKernel): 4 lanes per query once < 1/16 of the queries needed rings >= 2
  while (iterate) {
    if (!rigidOk(T_iter)) { status = setErr(ctx, B200REG_E_TRANSFORM, "RigidTransformation: Error, rotation matrix is not orthogonal."); break; }
    if (lanesCfg == 0 && cfg.knn == 1) ctx->knnLanes = lanesAuto;
    status = enqueueStep(ctx, icp->index, (const float4*)icp->reading.p, needNrm ? (const float4*)icp->readingNrm.p : nullptr, n, T_iter,
                         cfg.knn, cfg.max_dist, cfg.trim_ratio, cfg.max_angle, cfg.force2D, st.iterations > 0 ? 1 : 0);
    ctx->knnLanes = lanesCfg;
    if (status != B200REG_OK) break;
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { status = setErr(ctx, B200REG_E_CUDA, "ICP iteration failed: %s", cudaGetErrorString(e)); break; }
    if (cfg.trim_ratio > 0.f && h->sel.error) { status = setErr(ctx, B200REG_E_NOT_CONVERGED, "No matches available for computing distance quantiles"); break; }
    if (h->mom[M_CNT] < 1.0) { status = setErr(ctx, B200REG_E_NOT_CONVERGED, "ErrorMnimizer: no point to minimize"); break; }
    st.kept_last = (int64_t)h->mom[M_CNT];
    lanesAuto = ((h->far - farPrev) * 16ull < (unsigned long long)n) ? 4 : 8;
    farPrev = h->far;
    st.point_used_ratio = (float)h->mom[M_CNT] / (float)total;
    st.overlap_ratio = st.point_used_ratio;
    st.last_tau = (cfg.trim_ratio > 0.f) ? *(const float*)&h->sel.tauBits : INFINITY;
    double A[36], b[6], mr[3], ms[3];
    momentsToSystem(h->mom, A, b, mr, ms);
    float Tstep[16];
    solveP2P(A, b, mr, ms, cfg.force2D, cfg.force4DOF, Tstep, nullptr);
    mul4(Tstep, T_iter, T_iter);   // T_iter = T_step * T_iter (ICP.cpp:433)
    float errv[4] = {0, 0, 0, 0};
    status = icpErrToStatus(ctx, checkersCheck(ccfg, chk, T_iter, iterate, maxIter, errv), errv);
    ++st.iterations;
    if (status != B200REG_OK) break;
    if (iterate && st.iterations >= ICP_HARD_ITERATION_CAP) { status = icpErrToStatus(ctx, ICPERR_ITER_CAP, errv); break; }
  }
  cudaEventRecord(ctx->ev1, ctx->stream);
  cudaEventSynchronize(ctx->ev1);
  float ms = 0;
  cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
  st.device_ms = ms;
  st.points_touched = h->visits;
  st.max_iterations_reached = maxIter ? 1 : 0;
  st.convergence_s = nowSec() - t1;
  if (status != B200REG_OK) return done(status);
  return done(composeResult(icp, T_iter, T_init, T_out));
}

}  // extern "C"

// =================================================================================================================
// b200reg_batch — independent registrations kept in flight on ONE device (loop-closure candidate refinement,
// PlaceRecognition.cpp:71-174: independent iterations of a `for` loop).  A single registration leaves the machine
// under half full during the match and nearly idle during select / reduce / solve (profiles/r1_summary.md), so the batch owns
// `lanes` independent lanes — each a ctx (stream, scratch, pinned staging), an ICP object (loop block + instantiated while-graph)
// and two reusable device clouds — and deals pair i to lane i mod lanes.  Everything a pair needs (uploads, index build, reading
// preparation, the device-resident loop, the read-back of the 144-byte result) is ENQUEUED on the lane's stream without a host
// wait; the host collects a lane's result only when it needs that lane for the next pair.  The kernels of different lanes
// interleave on the device: while one lane solves its 6x6 system another lane's match kernel has the SMs.
// Results are those of the serial path, bit for bit: a lane runs exactly the code b200reg_icp_register runs.
// =================================================================================================================
#define BATCH_MAX_LANES 16
#define BATCH_DEFAULT_LANES 8
#define BATCH_LOOP_GRID_PER_SM 2   /* resident match-kernel blocks per SM of a lane (buildLoopGraph) */
struct b200reg_batch {
  b200reg_ctx* parent = nullptr;
  b200reg_icp_config cfg{};
  int lanes = 0;
  cudaEvent_t parentReady = nullptr;   // the parent stream's position when a batch call starts (its clouds may still be in flight)
  b200reg_index* shared = nullptr;     // b200reg_batch_set_reference: ONE index (built on the parent stream) every lane registers against
  double sharedPrepS = 0;
  struct Lane {
    b200reg_ctx* ctx = nullptr;
    b200reg_icp* icp = nullptr;
    int32_t inFlight = -1;              // pair index
    // the host-buffer entry point stages a pair's clouds one lane-cycle AHEAD: pair i + lanes is uploaded on the lane's copy context
    // (a stream and staging scratch of its own) into the other of two buffer sets while pair i iterates
    b200reg_ctx* copyCtx = nullptr;
    b200reg_cloud* ref[2] = {nullptr, nullptr};       // reusable device clouds, ping-pong
    b200reg_cloud* reading[2] = {nullptr, nullptr};
    cudaEvent_t staged[2] = {nullptr, nullptr};       // the uploads into set b are complete (copy stream)
    int stageCode[2] = {0, 0};
    int32_t stagedPair[2] = {-1, -1};
  } lane[BATCH_MAX_LANES];
};

namespace {

int batchCollect(b200reg_batch* B, b200reg_batch::Lane& L, const float* T_init, float* T_out, b200reg_icp_stats* stats, int32_t* status, int* first) {
  const int32_t i = L.inFlight;
  if (i < 0) return B200REG_OK;
  L.inFlight = -1;
  b200reg_icp_stats st;
  const int code = icpFinish(L.icp, T_out + 16 * (size_t)i, &st);
  if (code != B200REG_OK) {
    memcpy(T_out + 16 * (size_t)i, T_init + 16 * (size_t)i, 16 * sizeof(float));   // what Mapper.cpp:420-428 falls back to when ICP throws
    if (*first == B200REG_OK) { *first = code; setErr(B->parent, code, "pair %d: %s", (int)i, L.ctx->err); }
  }
  B->parent->launches += L.ctx->launches;
  L.ctx->launches = 0;
  if (stats) stats[i] = st;
  if (status) status[i] = code;
  return code;
}

void batchFail(b200reg_batch* B, b200reg_batch::Lane& L, int32_t i, int code, const float* T_init, float* T_out, b200reg_icp_stats* stats,
               int32_t* status, int* first) {
  memcpy(T_out + 16 * (size_t)i, T_init + 16 * (size_t)i, 16 * sizeof(float));
  if (*first == B200REG_OK) { *first = code; setErr(B->parent, code, "pair %d: %s", (int)i, L.ctx->err); }
  if (stats) { stats[i] = L.icp->pendStats; }
  if (status) status[i] = code;
}

// the loop of both batch entry points; `fetch` brings pair i's clouds to the lane (device pointers or uploads)
// `stage(L, j)` (host-buffer entry point; a no-op otherwise) starts the uploads of pair j on the lane's copy context: it is called one
// lane-cycle ahead — pair i + lanes right after pair i - lanes was collected, i.e. when the buffer set it goes into is free — so the
// copies of the next pair of a lane overlap the loop of the current one.
template <class Stage, class Fetch>
int batchRun(b200reg_batch* B, int32_t n, const float* T_init, float* T_out, b200reg_icp_stats* stats, int32_t* status, Stage stage, Fetch fetch) {
  b200reg_ctx* P = B->parent;
  CUDA_TRY(P, cudaSetDevice(P->device));
  CUDA_TRY(P, cudaEventRecord(B->parentReady, P->stream));
  for (int l = 0; l < B->lanes; ++l) CUDA_TRY(P, cudaStreamWaitEvent(B->lane[l].ctx->stream, B->parentReady, 0));
  int first = B200REG_OK;
  for (int32_t i = 0; i < n && i < B->lanes; ++i) stage(B->lane[i], i);   // the first cycle's uploads go first, in pair order: lane 0 starts after ONE copy
  for (int32_t i = 0; i < n; ++i) {
    b200reg_batch::Lane& L = B->lane[i % B->lanes];
    batchCollect(B, L, T_init, T_out, stats, status, &first);
    if (i + B->lanes < n) stage(L, i + B->lanes);
    const b200reg_cloud *ref = nullptr, *reading = nullptr;
    memset(&L.icp->pendStats, 0, sizeof(L.icp->pendStats));
    int code = fetch(L, i, &ref, &reading);
    if (code == B200REG_OK && ref) code = b200reg_icp_set_reference(L.icp, ref);
    else if (code == B200REG_OK) {   // the batch's shared reference (read-only for the lanes)
      if (L.icp->index && L.icp->ownsIndex) b200reg_index_free(L.ctx, L.icp->index);
      L.icp->index = B->shared;
      L.icp->ownsIndex = false;
      L.icp->refPrepS = B->sharedPrepS;
      if (!B->shared) code = setErr(L.ctx, B200REG_E_NO_POINTS, "ICP: matcher not initialised (b200reg_batch_set_reference was not called)");
    }
    if (code == B200REG_OK) code = icpEnqueue(L.icp, reading, T_init + 16 * (size_t)i);
    if (code == B200REG_OK) L.inFlight = i;
    else batchFail(B, L, i, code, T_init, T_out, stats, status, &first);
  }
  for (int l = 0; l < B->lanes; ++l) batchCollect(B, B->lane[l], T_init, T_out, stats, status, &first);
  return status ? B200REG_OK : first;   // with a status array the per-pair codes are the result
}

}  // namespace

extern "C" {

int32_t b200reg_batch_create(b200reg_ctx* ctx, const b200reg_icp_config* cfg, int32_t lanes, b200reg_batch** out) {
  if (!ctx || !cfg || !out) return setErr(ctx, B200REG_E_INVALID_PARAM, "batch_create: bad arguments");
  *out = nullptr;
  if (cfg->knn != 1) return setErr(ctx, B200REG_E_INVALID_PARAM, "batch_create: the batched lanes run the device-resident loop (knn = 1)");
  if (lanes <= 0) {
    const char* env = getenv("B200REG_BATCH_LANES");
    lanes = env ? atoi(env) : 0;
    if (lanes <= 0) lanes = BATCH_DEFAULT_LANES;
  }
  if (lanes > BATCH_MAX_LANES) lanes = BATCH_MAX_LANES;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  b200reg_batch* B = new b200reg_batch();
  B->parent = ctx;
  B->cfg = *cfg;
  B->lanes = lanes;
  int code = B200REG_OK;
  if (cudaEventCreateWithFlags(&B->parentReady, cudaEventDisableTiming) != cudaSuccess) code = setErr(ctx, B200REG_E_CUDA, "batch_create: event");
  for (int l = 0; l < lanes && code == B200REG_OK; ++l) {
    b200reg_batch::Lane& L = B->lane[l];
    code = b200reg_ctx_create(ctx->device, &L.ctx);
    if (code != B200REG_OK) { setErr(ctx, code, "batch_create: lane context"); break; }
    L.ctx->profile = 0; L.ctx->hostLoop = 0;   // lanes always run the graph loop
    L.ctx->loopGridPerSm = BATCH_LOOP_GRID_PER_SM;
    if (const char* e = getenv("B200REG_BATCH_GRID_PER_SM")) { const int v = atoi(e); if (v >= 1 && v <= KNN1_LOOP_MINB) L.ctx->loopGridPerSm = v; }   // A/B
    code = b200reg_icp_create(L.ctx, cfg, &L.icp);
    if (code != B200REG_OK) { setErr(ctx, code, "batch_create: %s", L.ctx->err); break; }
    L.icp->graphState = (buildLoopGraph(L.icp) == 0) ? 1 : -1;
    if (L.icp->graphState != 1) code = setErr(ctx, B200REG_E_CUDA, "batch_create: the device-resident loop (conditional graph nodes) is unavailable");
    if (code == B200REG_OK && (cudaEventCreateWithFlags(&L.staged[0], cudaEventDisableTiming) != cudaSuccess ||
                               cudaEventCreateWithFlags(&L.staged[1], cudaEventDisableTiming) != cudaSuccess))
      code = setErr(ctx, B200REG_E_CUDA, "batch_create: event");
  }
  if (code != B200REG_OK) { b200reg_batch_destroy(B); return code; }
  *out = B;
  return B200REG_OK;
}

void b200reg_batch_destroy(b200reg_batch* B) {
  if (!B) return;
  cudaSetDevice(B->parent->device);
  for (int l = 0; l < B->lanes; ++l) {
    b200reg_batch::Lane& L = B->lane[l];
    if (!L.ctx) continue;
    cudaStreamSynchronize(L.ctx->stream);
    if (L.icp) { L.icp->pending = false; b200reg_icp_destroy(L.icp); }   // a shared index is not the lane's to free
    if (L.copyCtx) cudaStreamSynchronize(L.copyCtx->stream);
    for (int b = 0; b < 2; ++b) {
      if (L.ref[b]) b200reg_cloud_free(L.ctx, L.ref[b]);
      if (L.reading[b]) b200reg_cloud_free(L.ctx, L.reading[b]);
      if (L.staged[b]) cudaEventDestroy(L.staged[b]);
    }
    if (L.copyCtx) b200reg_ctx_destroy(L.copyCtx);
    b200reg_ctx_destroy(L.ctx);
  }
  if (B->shared) b200reg_index_free(B->parent, B->shared);
  if (B->parentReady) cudaEventDestroy(B->parentReady);
  delete B;
}

int32_t b200reg_batch_lanes(const b200reg_batch* B) { return B ? B->lanes : 0; }

// ICP::initReference once for the whole batch: the index is built on the parent stream and every lane registers its readings
// against it (many scans / many initial guesses against one map: re-localisation, multi-sequence mapping).
int32_t b200reg_batch_set_reference(b200reg_batch* B, const b200reg_cloud* reference) {
  if (!B || !reference) return B200REG_E_INVALID_PARAM;
  b200reg_ctx* P = B->parent;
  const double t0 = nowSec();
  CUDA_TRY(P, cudaSetDevice(P->device));
  for (int l = 0; l < B->lanes; ++l) {   // no lane may still hold the old index
    b200reg_batch::Lane& L = B->lane[l];
    CUDA_TRY(P, cudaStreamSynchronize(L.ctx->stream));
    if (L.icp->index && !L.icp->ownsIndex) L.icp->index = nullptr;
  }
  if (B->shared) { b200reg_index_free(P, B->shared); B->shared = nullptr; }
  if (reference->n == 0) return setErr(P, B200REG_E_NO_POINTS, "initReference: reference cloud is empty");
  TRY(buildIndex(P, reference, 1, B->cfg.cell, &B->shared));
  B->sharedPrepS = nowSec() - t0;
  return B200REG_OK;
}

int32_t b200reg_batch_register(b200reg_batch* B, int32_t n, const b200reg_cloud* const* refs, const b200reg_cloud* const* readings,
                               const float* T_init, float* T_out, b200reg_icp_stats* stats, int32_t* status) {
  if (!B) return B200REG_E_INVALID_PARAM;
  if (n < 0 || (n > 0 && (!readings || !T_init || !T_out))) return setErr(B->parent, B200REG_E_INVALID_PARAM, "batch_register: bad arguments");
  return batchRun(B, n, T_init, T_out, stats, status, [](b200reg_batch::Lane&, int32_t) {},
                  [&](b200reg_batch::Lane& L, int32_t i, const b200reg_cloud** ref, const b200reg_cloud** reading) {
    if ((refs && !refs[i]) || !readings[i]) return setErr(L.ctx, B200REG_E_INVALID_PARAM, "batch_register: null cloud");
    *ref = refs ? refs[i] : nullptr;
    *reading = readings[i];
    return (int)B200REG_OK;
  });
}

int32_t b200reg_batch_register_host(b200reg_batch* B, int32_t n, const b200reg_host_cloud* refs, const b200reg_host_cloud* readings,
                                    const float* T_init, float* T_out, b200reg_icp_stats* stats, int32_t* status) {
  if (!B) return B200REG_E_INVALID_PARAM;
  if (n < 0 || (n > 0 && (!readings || !T_init || !T_out))) return setErr(B->parent, B200REG_E_INVALID_PARAM, "batch_register_host: bad arguments");
  const int lanes = B->lanes;
  for (int l = 0; l < lanes && n > 0; ++l) {   // the copy contexts are created the first time host clouds come in
    b200reg_batch::Lane& L = B->lane[l];
    if (L.copyCtx) continue;
    const int code = b200reg_ctx_create(B->parent->device, &L.copyCtx);
    if (code != B200REG_OK) { L.copyCtx = nullptr; return setErr(B->parent, code, "batch_register_host: lane copy context"); }
  }
  // uploads of pair j on the lane's copy context into buffer set (j / lanes) & 1: asynchronous for pinned host memory (pageable memory is
  // staged by the runtime and only the staging copy blocks this thread)
  auto stage = [&](b200reg_batch::Lane& L, int32_t j) {
    const int b = (j / lanes) & 1;
    const b200reg_host_cloud& hq = readings[j];
    int code = B200REG_OK;
    if (refs) {
      const b200reg_host_cloud& hr = refs[j];
      code = L.ref[b] ? b200reg_cloud_update_f32(L.copyCtx, L.ref[b], hr.feat, hr.feat_stride, hr.normals, hr.n, 1)
                      : b200reg_cloud_upload_f32(L.copyCtx, hr.feat, hr.feat_stride, hr.normals, hr.n, &L.ref[b]);
    }
    if (code == B200REG_OK)
      code = L.reading[b] ? b200reg_cloud_update_f32(L.copyCtx, L.reading[b], hq.feat, hq.feat_stride, hq.normals, hq.n, 1)
                          : b200reg_cloud_upload_f32(L.copyCtx, hq.feat, hq.feat_stride, hq.normals, hq.n, &L.reading[b]);
    if (code != B200REG_OK) setErr(L.ctx, code, "%s", L.copyCtx->err);
    else if (cudaEventRecord(L.staged[b], L.copyCtx->stream) != cudaSuccess) code = setErr(L.ctx, B200REG_E_CUDA, "batch_register_host: event record failed");
    L.stageCode[b] = code;
    L.stagedPair[b] = j;
  };
  return batchRun(B, n, T_init, T_out, stats, status, stage,
                  [&](b200reg_batch::Lane& L, int32_t i, const b200reg_cloud** ref, const b200reg_cloud** reading) {
    const int b = (i / lanes) & 1;
    *ref = nullptr;
    if (L.stagedPair[b] != i) return setErr(L.ctx, B200REG_E_INVALID_PARAM, "batch_register_host: pair %d was not staged", (int)i);
    L.stagedPair[b] = -1;
    if (L.stageCode[b] != B200REG_OK) return L.stageCode[b];
    if (cudaStreamWaitEvent(L.ctx->stream, L.staged[b], 0) != cudaSuccess) return setErr(L.ctx, B200REG_E_CUDA, "batch_register_host: stream wait failed");
    if (refs) *ref = L.ref[b];
    *reading = L.reading[b];
    return (int)B200REG_OK;
  });
}

// One-shot form: a batch object with the default number of lanes for the duration of the call.
int32_t b200reg_icp_register_batch(b200reg_ctx* ctx, const b200reg_icp_config* cfg, int32_t n, const b200reg_cloud* const* refs,
                                   const b200reg_cloud* const* readings, const float* T_init, float* T_out, b200reg_icp_stats* stats,
                                   int32_t* status) {
  if (!ctx || !cfg || n < 0 || (n > 0 && (!refs || !readings || !T_init || !T_out)))
    return setErr(ctx, B200REG_E_INVALID_PARAM, "icp_register_batch: bad arguments");
  b200reg_batch* B = nullptr;
  TRY(b200reg_batch_create(ctx, cfg, n < BATCH_DEFAULT_LANES ? (n > 0 ? n : 1) : 0, &B));
  const int code = b200reg_batch_register(B, n, refs, readings, T_init, T_out, stats, status);
  b200reg_batch_destroy(B);
  return code;
}

}  // extern "C"

// =================================================================================================================
// Standalone-module lane: the same kernels behind the Eigen-shaped interfaces of the individual libpointmatcher
// modules (host matrices in, host matrices out) — what the C++ shim classes call when a chain mixes CPU and CUDA modules.
// =================================================================================================================
namespace {

__global__ void trimWeightsKernel(const float* __restrict__ d2, int64_t count, const SelState* __restrict__ st, float* __restrict__ w) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  w[i] = (d2[i] <= __uint_as_float(st->tauBits)) ? 1.f : 0.f;   // OutlierFiltersImpl.cpp:146
}

// MaxDist / MinDist / MedianDist weights: w = (d2 <= limit) or (d2 >= limit) (OutlierFiltersImpl.cpp:74-81,95-102,116-125);
// scaleByTau: limit = fl(limit * tau) with tau the quantile the select kernels left in `st` (MedianDist: factor * median)
__global__ void distWeightsKernel(const float* __restrict__ d2, int64_t count, float limit, int greater, const SelState* __restrict__ st,
                                  int scaleByTau, float* __restrict__ w, float* __restrict__ limitOut) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (scaleByTau) limit = __fmul_rn(limit, __uint_as_float(st->tauBits));
  if (i == 0 && limitOut) *limitOut = limit;
  if (i >= count) return;
  const float d = d2[i];
  w[i] = (greater ? (d >= limit) : (d <= limit)) ? 1.f : 0.f;
}

// SurfaceNormalOutlierFilter::compute (OutlierFiltersImpl.cpp:236-281) on explicit arrays (normals: 3 floats per point)
__global__ void normalGateKernel(const float* __restrict__ nRead, const float* __restrict__ nRef, const int* __restrict__ ids, int k,
                                 int64_t total, float eps, float* __restrict__ w) {
  const int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= total) return;
  const int id = ids[o];
  if (id == -1) { w[o] = 0.f; return; }
  const float* a = nRead + 3 * (o / k);
  const float* b = nRef + 3 * (int64_t)id;
  const float v = __fadd_rn(__fadd_rn(__fmul_rn(a[0], b[0]), __fmul_rn(a[1], b[1])), __fmul_rn(a[2], b[2]));
  w[o] = (v < eps) ? 0.f : 1.f;
}

}  // namespace

extern "C" {

int32_t b200reg_trimmed_dist(b200reg_ctx* ctx, const float* d2, int64_t count, float ratio, float* tau_out, float* w_out) {
  if (!ctx || count < 0 || (count > 0 && !d2)) return setErr(ctx, B200REG_E_INVALID_PARAM, "trimmed_dist: bad arguments");
  if (!(ratio >= 0.f) || ratio > 1.f) return setErr(ctx, B200REG_E_NOT_CONVERGED, "Distance quantile of matches must lie in the range [0,1]");
  if (count == 0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "No matches available for computing distance quantiles");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  TRY(ensureBuf(ctx, ctx->matchD2, sizeof(float) * (size_t)count));
  TRY(ensureBuf(ctx, ctx->misc, sizeof(float) * (size_t)count));
  float* dd = (float*)ctx->matchD2.p;
  float* dw = (float*)ctx->misc.p;
  CUDA_TRY(ctx, cudaMemcpyAsync(dd, d2, sizeof(float) * (size_t)count, cudaMemcpyHostToDevice, ctx->stream));
  SelState* st = selState(ctx);
  TRY(launchSelect(ctx, dd, count, ratio, nullptr));
  trimWeightsKernel<<<divUp(count, 256), 256, 0, ctx->stream>>>(dd, count, st, dw);
  LAUNCH_CHECK(ctx);
  SelState* h = (SelState*)ctx->pinned;
  CUDA_TRY(ctx, cudaMemcpyAsync(h, st, sizeof(SelState), cudaMemcpyDeviceToHost, ctx->stream));
  if (w_out) CUDA_TRY(ctx, cudaMemcpyAsync(w_out, dw, sizeof(float) * (size_t)count, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (h->error) return setErr(ctx, B200REG_E_NOT_CONVERGED, "No matches available for computing distance quantiles");
  if (tau_out) memcpy(tau_out, &h->tauBits, sizeof(float));
  return B200REG_OK;
}

int32_t b200reg_dist_outlier(b200reg_ctx* ctx, int32_t kind, const float* d2, int64_t count, float param, float* limit_out, float* w_out) {
  if (!ctx || count < 0 || (count > 0 && (!d2 || !w_out)) || kind < B200REG_OUTLIER_MAX_DIST || kind > B200REG_OUTLIER_MEDIAN_DIST)
    return setErr(ctx, B200REG_E_INVALID_PARAM, "dist_outlier: bad arguments");
  // parameter bounds of the CPU twins (OutlierFiltersImpl.h:46-49,64-67,83-86)
  if (kind == B200REG_OUTLIER_MEDIAN_DIST && !(param >= 0.0000001f)) return setErr(ctx, B200REG_E_INVALID_PARAM, "MedianDistOutlierFilter: factor must be >= 1e-7");
  if (kind != B200REG_OUTLIER_MEDIAN_DIST && !(param >= 0.0000001f)) return setErr(ctx, B200REG_E_INVALID_PARAM, "%s: distance must be >= 1e-7", kind == B200REG_OUTLIER_MAX_DIST ? "MaxDistOutlierFilter" : "MinDistOutlierFilter");
  if (kind == B200REG_OUTLIER_MEDIAN_DIST && count == 0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "No matches available for computing distance quantiles");
  if (limit_out) *limit_out = 0.f;
  if (count == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  TRY(ensureBuf(ctx, ctx->matchD2, sizeof(float) * (size_t)count));
  TRY(ensureBuf(ctx, ctx->misc, sizeof(float) * (size_t)count + 64));
  float* dd = (float*)ctx->matchD2.p;
  float* dw = (float*)ctx->misc.p;
  float* dLimit = dw + count;
  CUDA_TRY(ctx, cudaMemcpyAsync(dd, d2, sizeof(float) * (size_t)count, cudaMemcpyHostToDevice, ctx->stream));
  SelState* st = selState(ctx);
  const bool median = kind == B200REG_OUTLIER_MEDIAN_DIST;
  if (median) TRY(launchSelect(ctx, dd, count, 0.5f, nullptr));   // input.getDistsQuantile(0.5)
  // pow(maxDist, 2) is evaluated in double and rounded once to T = float: the same value as the fp32 product
  const float limit = median ? param : param * param;
  distWeightsKernel<<<divUp(count, 256), 256, 0, ctx->stream>>>(dd, count, limit, kind == B200REG_OUTLIER_MIN_DIST ? 1 : 0, st, median ? 1 : 0, dw, dLimit);
  LAUNCH_CHECK(ctx);
  SelState* h = (SelState*)ctx->pinned;
  float* hLimit = (float*)((char*)ctx->pinned + 1024);
  CUDA_TRY(ctx, cudaMemcpyAsync(h, st, sizeof(SelState), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(hLimit, dLimit, sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(w_out, dw, sizeof(float) * (size_t)count, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (median && h->error) return setErr(ctx, B200REG_E_NOT_CONVERGED, "No matches available for computing distance quantiles");
  if (limit_out) *limit_out = *hLimit;
  return B200REG_OK;
}

int32_t b200reg_surface_normal_outlier(b200reg_ctx* ctx, const float* normals_reading, const float* normals_reference, int64_t n_reference,
                                       const int32_t* ids, int32_t k, int64_t n, float max_angle, float* w_out) {
  if (!ctx || !ids || !w_out || k < 1 || n < 0) return setErr(ctx, B200REG_E_INVALID_PARAM, "surface_normal_outlier: bad arguments");
  const int64_t total = n * k;
  if (!normals_reading || !normals_reference) {   // "surface normals not available. Skipping filtering" (:268-277)
    for (int64_t i = 0; i < total; ++i) w_out[i] = 1.f;
    return B200REG_OK;
  }
  if (total == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const size_t bR = sizeof(float) * 3 * (size_t)n, bF = sizeof(float) * 3 * (size_t)n_reference, bI = sizeof(int) * (size_t)total;
  TRY(ensureBuf(ctx, ctx->misc, bR + bF + bI + sizeof(float) * (size_t)total + 64));
  float* dR = (float*)ctx->misc.p;
  float* dF = dR + 3 * (size_t)n;
  int* dI = (int*)(dF + 3 * (size_t)n_reference);
  float* dW = (float*)(dI + total);
  CUDA_TRY(ctx, cudaMemcpyAsync(dR, normals_reading, bR, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(dF, normals_reference, bF, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(dI, ids, bI, cudaMemcpyHostToDevice, ctx->stream));
  normalGateKernel<<<divUp(total, 256), 256, 0, ctx->stream>>>(dR, dF, dI, k, total, cosf(max_angle), dW);
  LAUNCH_CHECK(ctx);
  CUDA_TRY(ctx, cudaMemcpyAsync(w_out, dW, sizeof(float) * (size_t)total, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return B200REG_OK;
}

// ErrorMinimizer::compute(reading, reference, weights, matches) (ErrorMinimizer.cpp:218-232) without materialising
// ErrorElements: explicit matches + weights in, normal equations + ErrorElements statistics out.
int32_t b200reg_point_to_plane(b200reg_ctx* ctx, const b200reg_cloud* reading, const b200reg_cloud* reference, const int32_t* ids,
                               const float* d2, const float* w, int32_t k, int32_t force2D, double A_out[36], double b_out[6],
                               b200reg_step_stats* stats) {
  if (!ctx || !reading || !reference || !ids || !d2 || !w || k < 1) return setErr(ctx, B200REG_E_INVALID_PARAM, "point_to_plane: bad arguments");
  if (!reference->nrm) return setErr(ctx, B200REG_E_INVALID_FIELD, "PointToPlaneErrorMinimizer: reference cloud has no 'normals' descriptor");
  const int64_t n = reading->n, total = n * k;
  if (total == 0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "ErrorMnimizer: no point to minimize");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  TRY(ensureBuf(ctx, ctx->matchIds, sizeof(int) * (size_t)total));
  TRY(ensureBuf(ctx, ctx->matchD2, sizeof(float) * (size_t)total));
  TRY(ensureBuf(ctx, ctx->misc, sizeof(float) * (size_t)total));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->matchIds.p, ids, sizeof(int) * (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->matchD2.p, d2, sizeof(float) * (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->misc.p, w, sizeof(float) * (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  ReduceParams P;
  memset(&P, 0, sizeof(P));
  P.reading = reading->xyz;
  P.refPts = reference->xyz;
  P.refNrm = reference->nrm;
  P.ids = (const int*)ctx->matchIds.p;
  P.d2 = (const float*)ctx->matchD2.p;
  P.pos = (const int*)ctx->matchIds.p;   // reference in original order: position == id
  P.n = n;
  P.k = k;
  P.force2D = force2D;
  P.w = (const float*)ctx->misc.p;
  SelState* st = selState(ctx);
  const int rblocks = (int)fmax(1.0, fmin((double)divUp(total, 2 * RED_THREADS), (double)((512 / RED_THREADS) * ctx->sm_count)));
  TRY(ensureBuf(ctx, ctx->partials, sizeof(double) * NUM_MOM * (size_t)rblocks));
  reduceKernel<<<rblocks, RED_THREADS, 0, ctx->stream>>>(P, st, (double*)ctx->partials.p, &st->ticket[3], resultPtr(ctx), nullptr, nullptr);
  LAUNCH_CHECK(ctx);
  HostStep* h = (HostStep*)ctx->pinned;
  CUDA_TRY(ctx, cudaMemcpyAsync(h->mom, resultPtr(ctx), sizeof(double) * NUM_MOM, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (stats) {
    memset(stats, 0, sizeof(*stats));
    stats->kept = (int64_t)h->mom[M_CNT];
    stats->n_finite = (int64_t)h->mom[M_VALID];
    stats->weight_sum = h->mom[M_SPARE];
    stats->tau = INFINITY;
    stats->point_used_ratio = (float)h->mom[M_CNT] / (float)total;
    stats->weighted_point_used_ratio = (float)h->mom[M_SPARE] / (float)total;
  }
  if (h->mom[M_CNT] < 1.0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "ErrorMnimizer: no point to minimize");
  double A[36], b[6], mr[3], ms[3];
  momentsToSystem(h->mom, A, b, mr, ms);
  if (A_out) memcpy(A_out, A, sizeof(A));
  if (b_out) memcpy(b_out, b, sizeof(b));
  if (stats) for (int d = 0; d < 3; ++d) { stats->mean_reading[d] = mr[d]; stats->mean_reference[d] = ms[d]; }
  return B200REG_OK;
}

}  // extern "C"
