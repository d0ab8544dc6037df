// csrc/icp.cu — host driver of the ICP iterate loop (ICP.cpp:258-468) over the device kernels, the 6x6 solve
// (PointToPlane.cpp:185-368) and the TransformationCheckers (TransformationCheckersImpl.cpp:58-225).
#include <chrono>
#include <cmath>
#include <vector>

#include "icp_kernels.cuh"

int runKnn(b200reg_ctx* ctx, const b200reg_index* ix, const float4* queries, int64_t n, const float* T, int k, float maxDist, int allowSelf);
int cloudStats(b200reg_ctx* ctx, const b200reg_cloud* c, double sum[3], float mn[3], float mx[3]);
int buildIndex(b200reg_ctx* ctx, const b200reg_cloud* ref, const float mean[3], const float bbMin[3], const float bbMax[3], float cell,
               b200reg_index** out);

namespace {

inline SelState* selState(b200reg_ctx* ctx) { return (SelState*)((char*)ctx->dstate + DSTATE_SEL_OFF); }
inline unsigned* histPtr(b200reg_ctx* ctx) { return (unsigned*)((char*)ctx->dstate + DSTATE_HIST_OFF); }
inline double* resultPtr(b200reg_ctx* ctx) { return (double*)((char*)ctx->dstate + DSTATE_RESULT_OFF); }

struct HostStep {   // what one iteration brings back over PCIe: NUM_MOM doubles + the select state + the visit counter
  double mom[NUM_MOM];
  SelState sel;
  unsigned long long visits;
};

// device part of one iteration: match -> select -> reduce -> async D2H into pinned memory (no sync here)
int enqueueStep(b200reg_ctx* ctx, const b200reg_index* ix, const float4* reading, const float4* readingNrm, int64_t n, const float* T,
                int k, float maxDist, float trimRatio, float maxAngle, int force2D) {
  if (!ix->nrm) return setErr(ctx, B200REG_E_INVALID_FIELD, "PointToPlaneErrorMinimizer: reference cloud has no 'normals' descriptor");
  TRY(runKnn(ctx, ix, reading, n, T, k, maxDist, 1));
  const int64_t total = n * k;
  SelState* st = selState(ctx);
  const bool useTrim = trimRatio > 0.f;
  if (useTrim) {
    const int blocks = (int)fmin((double)divUp(total, 256 * 4), (double)(2 * ctx->sm_count));
    selectPassKernel<0><<<blocks, 256, 0, ctx->stream>>>((const float*)ctx->matchD2.p, total, trimRatio, st, histPtr(ctx));
    LAUNCH_CHECK(ctx);
    selectPassKernel<1><<<blocks, 256, 0, ctx->stream>>>((const float*)ctx->matchD2.p, total, trimRatio, st, histPtr(ctx));
    LAUNCH_CHECK(ctx);
    selectPassKernel<2><<<blocks, 256, 0, ctx->stream>>>((const float*)ctx->matchD2.p, total, trimRatio, st, histPtr(ctx));
    LAUNCH_CHECK(ctx);
  }
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
  const int rblocks = (int)fmax(1.0, fmin((double)divUp(total, 256), (double)(4 * ctx->sm_count)));
  TRY(ensureBuf(ctx, ctx->partials, sizeof(double) * NUM_MOM * (size_t)rblocks));
  reduceKernel<<<rblocks, 256, 0, ctx->stream>>>(P, st, (double*)ctx->partials.p, &st->ticket[3], resultPtr(ctx));
  LAUNCH_CHECK(ctx);
  HostStep* h = (HostStep*)ctx->pinned;
  CUDA_TRY(ctx, cudaMemcpyAsync(h->mom, resultPtr(ctx), sizeof(double) * NUM_MOM, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(&h->sel, st, sizeof(SelState), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(&h->visits, ctx->dstate, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
  return B200REG_OK;
}

inline void cross3(const double a[3], const double b[3], double o[3]) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}

// un-centred moments -> A = G G^T, b = -G h^T on the matched-mean-centred pairs (PointToPlane.cpp:263-268,107-156)
void momentsToSystem(const double* m, double A[36], double b[6], double meanR[3], double meanS[3]) {
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
bool cholSolve(const double* A, const double* b, int n, double* x) {
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
void jacobiEig(double* a, double* V, int n) {
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
void solveSym(const double* A, const double* b, int n, double* x) {
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

// ---- fp32 4x4 helpers (row-major), depth accumulated k = 0..3 ----------------------------------------------------
void mul4(const float* a, const float* b, float* r) {
  float t[16];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      float acc = a[4 * i] * b[j];
      acc = acc + a[4 * i + 1] * b[4 + j];
      acc = acc + a[4 * i + 2] * b[8 + j];
      acc = acc + a[4 * i + 3] * b[12 + j];
      t[4 * i + j] = acc;
    }
  memcpy(r, t, sizeof(t));
}
bool rigidOk(const float* T) {
  const float det = T[0] * (T[5] * T[10] - T[6] * T[9]) - T[1] * (T[4] * T[10] - T[6] * T[8]) + T[2] * (T[4] * T[9] - T[5] * T[8]);
  return !(fabsf(1.f - det) > 0.001f);
}
void quatOf(const float* T, float q[4]) {   // Eigen::Quaternion(Matrix3): (w, x, y, z)
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
float quatDist(const float a[4], const float b[4]) {   // QuaternionBase::angularDistance
  const float bw = b[0], bx = -b[1], by = -b[2], bz = -b[3];
  const float w = a[0] * bw - a[1] * bx - a[2] * by - a[3] * bz;
  const float x = a[0] * bx + a[1] * bw + a[2] * bz - a[3] * by;
  const float y = a[0] * by + a[2] * bw + a[3] * bx - a[1] * bz;
  const float z = a[0] * bz + a[3] * bw + a[1] * by - a[2] * bx;
  return 2.f * atan2f(sqrtf(x * x + y * y + z * z), fabsf(w));
}

struct Checkers {   // Counter / Differential / Bound (TransformationCheckersImpl.cpp:58-225)
  const b200reg_icp_config& c;
  int counter = 0;
  std::vector<float> rot, tr;   // 4 / 3 floats per pushed transform
  float q0[4], t0[3];
  explicit Checkers(const b200reg_icp_config& cfg) : c(cfg) {}
  void init(const float* T) {
    counter = 0;
    rot.clear(); tr.clear();
    push(T);
    quatOf(T, q0);
    t0[0] = T[3]; t0[1] = T[7]; t0[2] = T[11];
  }
  void push(const float* T) {
    float q[4];
    quatOf(T, q);
    rot.insert(rot.end(), q, q + 4);
    tr.push_back(T[3]); tr.push_back(T[7]); tr.push_back(T[11]);
  }
  // 0 ok, <0 ConvergenceError; may clear `iterate` / set `maxIter`
  int check(b200reg_ctx* ctx, const float* T, bool& iterate, bool& maxIter) {
    if (c.smooth_length >= 0) {
      push(T);
      const size_t sl = (size_t)c.smooth_length, cnt = rot.size() / 4;
      float cv0 = 0.f, cv1 = 0.f;
      if (cnt > sl) {
        for (size_t i = cnt - 1; sl > 0 && i >= cnt - sl; --i) {
          cv0 += fabsf(quatDist(&rot[4 * i], &rot[4 * (i - 1)]));
          const float dx = tr[3 * i] - tr[3 * (i - 1)], dy = tr[3 * i + 1] - tr[3 * (i - 1) + 1], dz = tr[3 * i + 2] - tr[3 * (i - 1) + 2];
          cv1 += fabsf(sqrtf(dx * dx + dy * dy + dz * dz));
        }
        cv0 /= (float)sl; cv1 /= (float)sl;
        if (cv0 < c.min_diff_rot && cv1 < c.min_diff_trans) iterate = false;
      }
      if (cv0 != cv0) return setErr(ctx, B200REG_E_NOT_CONVERGED, "abs rotation norm not a number");
      if (cv1 != cv1) return setErr(ctx, B200REG_E_NOT_CONVERGED, "abs translation norm not a number");
    }
    if (c.bound_rot >= 0.f) {
      float q[4];
      quatOf(T, q);
      const float r = quatDist(q, q0);
      const float dx = T[3] - t0[0], dy = T[7] - t0[1], dz = T[11] - t0[2];
      const float t = sqrtf(dx * dx + dy * dy + dz * dz);
      if (r > c.bound_rot || t > c.bound_trans)
        return setErr(ctx, B200REG_E_NOT_CONVERGED, "limit out of bounds: rot: %g/%g tr: %g/%g", r, c.bound_rot, t, c.bound_trans);
    }
    if (c.max_iter > 0) {
      ++counter;
      if (counter >= c.max_iter) { iterate = false; maxIter = true; }   // MaxNumIterationsReached, caught at ICP.cpp:441-445
    }
    return B200REG_OK;
  }
};

double nowSec() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

}  // namespace

struct b200reg_icp {
  b200reg_ctx* ctx = nullptr;
  b200reg_icp_config cfg{};
  b200reg_index* index = nullptr;   // over the centred, (un)filtered reference — `referenceFiltered` + the matcher
  float mean[3] = {0, 0, 0};        // T_refIn_refMean translation
  double refPrepS = 0;
  DevBuf reading, readingNrm;       // `reading` of ICP.cpp:352-375 (in the refMean frame)
};

extern "C" {

int32_t b200reg_solve_point_to_plane(const double A[36], const double b[6], const double mean_reading[3],
                                     const double mean_reference[3], int32_t force2D, int32_t force4DOF, float T_out[16],
                                     double x_out[6]) {
  if (!A || !b || !mean_reading || !mean_reference || !T_out) return B200REG_E_INVALID_PARAM;
  double x[6] = {0, 0, 0, 0, 0, 0};
  double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, t[3] = {0, 0, 0};
  if (force2D) {   // unknowns [gamma, tx, ty] = rows/cols {2,3,4} (PointToPlane.cpp:247-254, 334-366)
    const int sel[3] = {2, 3, 4};
    double As[9], bs[3], xs[3];
    for (int r = 0; r < 3; ++r) { bs[r] = b[sel[r]]; for (int c = 0; c < 3; ++c) As[3 * r + c] = A[6 * sel[r] + sel[c]]; }
    solveSym(As, bs, 3, xs);
    x[0] = xs[0]; x[1] = xs[1]; x[2] = xs[2];
    R[0] = cos(xs[0]); R[1] = -sin(xs[0]); R[3] = sin(xs[0]); R[4] = cos(xs[0]);
    t[0] = xs[1]; t[1] = xs[2];
  } else {
    double axis[3] = {0, 0, 1}, angle;
    if (force4DOF) {   // unknowns [gamma, tx, ty, tz] = rows/cols {2,3,4,5} (PointToPlane.cpp:128-134,297-313)
      const int sel[4] = {2, 3, 4, 5};
      double As[16], bs[4], xs[4];
      for (int r = 0; r < 4; ++r) { bs[r] = b[sel[r]]; for (int c = 0; c < 4; ++c) As[4 * r + c] = A[6 * sel[r] + sel[c]]; }
      solveSym(As, bs, 4, xs);
      for (int i = 0; i < 4; ++i) x[i] = xs[i];
      angle = xs[0];
      t[0] = xs[1]; t[1] = xs[2]; t[2] = xs[3];
    } else {
      solveSym(A, b, 6, x);
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
  return B200REG_OK;
}

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

__global__ void prepReadingKernel(const float4* __restrict__ xyz, const float4* __restrict__ nrm, int64_t n, Rigid12 T, float t30, float t31,
                                  float t32, float t33, float4* __restrict__ oxyz, float4* __restrict__ onrm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = xyz[i];
  const float3 r = xformPoint(T, p.x, p.y, p.z, p.w);
  const float w = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(t30, p.x), __fmul_rn(t31, p.y)), __fmul_rn(t32, p.z)), __fmul_rn(t33, p.w));
  oxyz[i] = make_float4(r.x, r.y, r.z, w);
  if (nrm) {
    const float4 q = nrm[i];
    const float3 s = xformNormal(T, q.x, q.y, q.z);
    onrm[i] = make_float4(s.x, s.y, s.z, 0.f);
  }
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
  b200reg_icp* icp = new b200reg_icp();
  icp->ctx = ctx;
  icp->cfg = *cfg;
  *out = icp;
  return B200REG_OK;
}

void b200reg_icp_destroy(b200reg_icp* icp) {
  if (!icp) return;
  b200reg_ctx* ctx = icp->ctx;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (icp->index) b200reg_index_free(ctx, icp->index);
  if (icp->reading.p) cudaFree(icp->reading.p);
  if (icp->readingNrm.p) cudaFree(icp->readingNrm.p);
  delete icp;
}

// ICP::initReference (ICP.cpp:293-328): mean of the reference, centring, matcher->init.
int32_t b200reg_icp_set_reference(b200reg_icp* icp, const b200reg_cloud* reference) {
  if (!icp || !reference) return B200REG_E_INVALID_PARAM;
  b200reg_ctx* ctx = icp->ctx;
  const double t0 = nowSec();
  if (icp->index) { b200reg_index_free(ctx, icp->index); icp->index = nullptr; }   // matcherIsInitialized = false
  if (reference->n == 0) return setErr(ctx, B200REG_E_NO_POINTS, "initReference: reference cloud is empty");   // :295-298
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  double sum[3];
  float mn[3], mx[3];
  TRY(cloudStats(ctx, reference, sum, mn, mx));
  // meanReference = features.rowwise().mean() (:313): fp64 sum rounded once to fp32 (contract)
  for (int d = 0; d < 3; ++d) icp->mean[d] = (float)(sum[d] / (double)reference->n);
  TRY(buildIndex(ctx, reference, icp->mean, mn, mx, icp->cfg.cell, &icp->index));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  icp->refPrepS = nowSec() - t0;
  return B200REG_OK;
}

const b200reg_index* b200reg_icp_index(const b200reg_icp* icp) { return icp ? icp->index : nullptr; }

int32_t b200reg_icp_reference_mean(const b200reg_icp* icp, float mean_out[3]) {
  if (!icp || !mean_out) return B200REG_E_INVALID_PARAM;
  for (int d = 0; d < 3; ++d) mean_out[d] = icp->mean[d];
  return B200REG_OK;
}

int32_t b200reg_icp_register(b200reg_icp* icp, const b200reg_cloud* reading, const float T_init[16], float T_out[16],
                             b200reg_icp_stats* stats) {
  if (!icp || !reading || !T_init || !T_out) return B200REG_E_INVALID_PARAM;
  b200reg_ctx* ctx = icp->ctx;
  const b200reg_icp_config& cfg = icp->cfg;
  b200reg_icp_stats st;
  memset(&st, 0, sizeof(st));
  auto done = [&](int code) { if (stats) *stats = st; return code; };
  if (!icp->index) return done(setErr(ctx, B200REG_E_NO_POINTS, "ICP: matcher not initialised (empty reference)"));
  if (!icp->index->nrm) return done(setErr(ctx, B200REG_E_INVALID_FIELD, "PointToPlaneErrorMinimizer: reference cloud has no 'normals' descriptor"));
  const int64_t n = reading->n;
  st.reference_point_count = icp->index->m;
  st.reading_point_count = n;
  st.reference_preprocessing_s = icp->refPrepS;
  if (n == 0) return done(setErr(ctx, B200REG_E_NO_POINTS, "The reading point cloud is empty."));   // ICP.cpp:357-359
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const double t0 = nowSec();
  // T_refMean_readMean = T_refIn_refMean^-1 * T_init (meanReading forced to 0 in this fork, ICP.cpp:363-373)
  float TrefInv[16] = {1, 0, 0, -icp->mean[0], 0, 1, 0, -icp->mean[1], 0, 0, 1, -icp->mean[2], 0, 0, 0, 1};
  float Tpre[16];
  mul4(TrefInv, T_init, Tpre);
  if (!rigidOk(Tpre)) return done(setErr(ctx, B200REG_E_TRANSFORM, "RigidTransformation: Error, rotation matrix is not orthogonal."));
  const bool needNrm = reading->nrm && cfg.max_angle >= 0.f;
  TRY(ensureBuf(ctx, icp->reading, sizeof(float4) * (size_t)n));
  if (needNrm) TRY(ensureBuf(ctx, icp->readingNrm, sizeof(float4) * (size_t)n));
  Rigid12 Rp;
  memcpy(Rp.m, Tpre, sizeof(Rp.m));
  prepReadingKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(reading->xyz, needNrm ? reading->nrm : nullptr, n, Rp, Tpre[12], Tpre[13], Tpre[14],
                                                            Tpre[15], (float4*)icp->reading.p, (float4*)icp->readingNrm.p);
  LAUNCH_CHECK(ctx);
  st.reading_preprocessing_s = nowSec() - t0;

  const double t1 = nowSec();
  float T_iter[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  bool iterate = true, maxIter = false;
  Checkers checkers(cfg);
  checkers.init(T_iter);
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->dstate, 0, sizeof(unsigned long long), ctx->stream));
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  int status = B200REG_OK;
  const HostStep* h = (const HostStep*)ctx->pinned;
  const double total = (double)n * cfg.knn;
  while (iterate) {
    if (!rigidOk(T_iter)) { status = setErr(ctx, B200REG_E_TRANSFORM, "RigidTransformation: Error, rotation matrix is not orthogonal."); break; }
    status = enqueueStep(ctx, icp->index, (const float4*)icp->reading.p, needNrm ? (const float4*)icp->readingNrm.p : nullptr, n, T_iter,
                         cfg.knn, cfg.max_dist, cfg.trim_ratio, cfg.max_angle, cfg.force2D);
    if (status != B200REG_OK) break;
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { status = setErr(ctx, B200REG_E_CUDA, "ICP iteration failed: %s", cudaGetErrorString(e)); break; }
    if (cfg.trim_ratio > 0.f && h->sel.error) { status = setErr(ctx, B200REG_E_NOT_CONVERGED, "No matches available for computing distance quantiles"); break; }
    if (h->mom[M_CNT] < 1.0) { status = setErr(ctx, B200REG_E_NOT_CONVERGED, "ErrorMnimizer: no point to minimize"); break; }
    st.kept_last = (int64_t)h->mom[M_CNT];
    st.point_used_ratio = (float)h->mom[M_CNT] / (float)total;
    st.overlap_ratio = st.point_used_ratio;
    st.last_tau = (cfg.trim_ratio > 0.f) ? *(const float*)&h->sel.tauBits : INFINITY;
    double A[36], b[6], mr[3], ms[3];
    momentsToSystem(h->mom, A, b, mr, ms);
    float Tstep[16];
    b200reg_solve_point_to_plane(A, b, mr, ms, cfg.force2D, cfg.force4DOF, Tstep, nullptr);
    mul4(Tstep, T_iter, T_iter);   // T_iter = T_step * T_iter (ICP.cpp:433)
    status = checkers.check(ctx, T_iter, iterate, maxIter);
    ++st.iterations;
    if (status != B200REG_OK) break;
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
  // icpCorrected_T_refIn_readIn = T_refIn_refMean * (T_iter * T_refMean_readMean) * I   (ICP.cpp:459-465)
  float Tref[16] = {1, 0, 0, icp->mean[0], 0, 1, 0, icp->mean[1], 0, 0, 1, icp->mean[2], 0, 0, 0, 1};
  float tmp[16];
  mul4(T_iter, Tpre, tmp);
  mul4(Tref, tmp, T_out);
  return done(B200REG_OK);
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
  const int blocks = (int)fmax(1.0, fmin((double)divUp(count, 256 * 4), (double)(2 * ctx->sm_count)));
  selectPassKernel<0><<<blocks, 256, 0, ctx->stream>>>(dd, count, ratio, st, histPtr(ctx));
  LAUNCH_CHECK(ctx);
  selectPassKernel<1><<<blocks, 256, 0, ctx->stream>>>(dd, count, ratio, st, histPtr(ctx));
  LAUNCH_CHECK(ctx);
  selectPassKernel<2><<<blocks, 256, 0, ctx->stream>>>(dd, count, ratio, st, histPtr(ctx));
  LAUNCH_CHECK(ctx);
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
  const int rblocks = (int)fmax(1.0, fmin((double)divUp(total, 256), (double)(4 * ctx->sm_count)));
  TRY(ensureBuf(ctx, ctx->partials, sizeof(double) * NUM_MOM * (size_t)rblocks));
  reduceKernel<<<rblocks, 256, 0, ctx->stream>>>(P, st, (double*)ctx->partials.p, &st->ticket[3], resultPtr(ctx));
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
