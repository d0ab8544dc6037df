// csrc/icp_kernels.cuh — device kernels of one ICP iteration after the match:
//   (1) radix select of the trimmed-distance limit  (Matches::getDistsQuantile, Matches.cpp:61-87)
//   (2) fused outlier gates + ErrorElements statistics + point-to-plane moment reduction (fp64)
//       (OutlierFiltersImpl.cpp:140-147,226-281; ErrorMinimizer.cpp:59-140; PointToPlane.cpp:107-156,263-268)
#pragma once
#include <cstddef>

#include "common.cuh"

// ---- device state block layout (ctx->dstate, 64 KB) -------------------------------------------------------------
struct SelState {
  unsigned prefix;     // selected high digits so far
  unsigned rank;       // remaining rank inside the selected bucket
  unsigned nFinite;    // finite matches
  unsigned tauBits;    // result: bit pattern of the limit
  unsigned ticket[4];  // last-block tickets (3 select passes + reduce)
  int error;           // 1: no finite match
  int pad[7];
};
#define DSTATE_VISITS_OFF 0
#define DSTATE_SEL_OFF 64
#define DSTATE_HIST_OFF 256              /* 2048 x uint32 */
#define DSTATE_RESULT_OFF (256 + 8192)   /* NUM_MOM doubles */

#define NUM_MOM 36
// threads per block of reduceKernel.  128-thread blocks (16 K registers: the footprint of one block of the match kernel, so a lane's
// reduction could start in the first slot another lane frees) were measured: 3635 vs 3730 registrations/s on the lanes, 1975 vs 2200 alone
#ifndef RED_THREADS
#define RED_THREADS 256
#endif
// moment slots
enum {
  M_CNT = 0, M_VALID = 1,
  M_R = 2,      // sum r (3)
  M_S = 5,      // sum s (3)
  M_UU = 8,     // sum u u^T: xx xy xz yy yz zz (6)
  M_UN = 14,    // sum u n^T row-major (9)
  M_NN = 23,    // sum n n^T: xx xy xz yy yz zz (6)
  M_UE = 29,    // sum u e (3)
  M_NE = 32,    // sum n e (3)
  M_SPARE = 35
};

// pass 0: digit = bits[31:21]; pass 1: bits[20:10]; pass 2: bits[9:0].  Values are non-negative floats, so the
// uint32 order of the bit patterns is the float order.  The last block to finish picks the bucket.
template <int PASS>
__global__ void __launch_bounds__(256) selectPassKernel(const float* __restrict__ d2, int64_t count, float ratio, SelState* st,
                                                        unsigned* __restrict__ hist, const IcpDevParams* __restrict__ DP) {
  if (DP) { d2 = DP->d2; count = DP->n; ratio = DP->trimRatio; traceStamp(DP->trace, 2 + PASS); }   // graph-driven loop (k = 1)
  __shared__ unsigned sh[2048];
  __shared__ unsigned scan[256];
  __shared__ bool last;
  constexpr int NB = (PASS == 2) ? 1024 : 2048;
  for (int b = threadIdx.x; b < NB; b += blockDim.x) sh[b] = 0;
  __syncthreads();
  const unsigned prefix = (PASS == 0) ? 0u : st->prefix;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned v = __float_as_uint(d2[i]);
    if (v == 0x7f800000u) continue;   // dists != inf (Matches.cpp:70)
    if (PASS == 0) atomicAdd(&sh[v >> 21], 1u);
    else if (PASS == 1) { if ((v >> 21) == prefix) atomicAdd(&sh[(v >> 10) & 0x7ffu], 1u); }
    else { if ((v >> 10) == prefix) atomicAdd(&sh[v & 0x3ffu], 1u); }
  }
  __syncthreads();
  for (int b = threadIdx.x; b < NB; b += blockDim.x)
    if (sh[b]) atomicAdd(&hist[b], sh[b]);
  __threadfence();
  if (threadIdx.x == 0) last = (atomicAdd(&st->ticket[PASS], 1u) == gridDim.x - 1);
  __syncthreads();
  if (!last) return;
  __threadfence();
  // ---- finalisation by the last block: find the bucket holding the wanted rank
  constexpr int PER = NB / 256;
  unsigned local[PER];
  unsigned sum = 0;
  for (int j = 0; j < PER; ++j) { local[j] = __ldcg(&hist[threadIdx.x * PER + j]); sum += local[j]; }
  scan[threadIdx.x] = sum;
  __syncthreads();
  for (int o = 1; o < 256; o <<= 1) {   // inclusive Hillis-Steele scan
    unsigned v = (threadIdx.x >= o) ? scan[threadIdx.x - o] : 0u;
    __syncthreads();
    scan[threadIdx.x] += v;
    __syncthreads();
  }
  __shared__ unsigned sRank;
  if (threadIdx.x == 0) {
    unsigned rank = st->rank;
    if (PASS == 0) {
      const unsigned nF = scan[255];
      st->nFinite = nF;
      st->error = (nF == 0) ? 1 : 0;
      if (nF > 0) {
        // rank = size_t(float(n) * quantile) — float multiply + truncation (Matches.cpp:85); quantile 1 -> max (:83)
        rank = (ratio == 1.0f) ? nF - 1 : (unsigned)__fmul_rn((float)nF, ratio);
        if (rank > nF - 1) rank = nF - 1;
      } else rank = 0;
    }
    sRank = rank;
  }
  __syncthreads();
  const unsigned rank = sRank;
  const unsigned incl = scan[threadIdx.x], excl = incl - sum;
  if (rank >= excl && rank < incl) {   // exactly one thread (when nFinite > 0)
    unsigned acc = excl;
    for (int j = 0; j < PER; ++j) {
      if (rank < acc + local[j]) {
        const unsigned digit = threadIdx.x * PER + j;
        if (PASS == 0) st->prefix = digit;
        else if (PASS == 1) st->prefix = (prefix << 11) | digit;
        else st->tauBits = (prefix << 10) | digit;
        st->rank = rank - acc;
        break;
      }
      acc += local[j];
    }
  }
  for (int b = threadIdx.x; b < 2048; b += blockDim.x) hist[b] = 0;   // ready for the next pass / next iteration
  if (threadIdx.x == 0) st->ticket[PASS] = 0;
}

// ---- the selection of the device-resident loop: ONE launch after the match --------------------------------------------
// The match kernel has already histogrammed the top 11 bits of every finite d2 (pass 0) into fhist0.  Here every block
// re-derives the pass-0 bucket from that histogram (2048 loads, redundant but off the critical path of nothing), runs
// pass 1 over its share of d2 and, in the same sweep, compacts the distances of the pass-0 bucket into `cand`; the last
// block to finish picks the pass-1 digit and runs pass 2 over the compacted candidates only.  Same tau bits as the
// three-launch version (selectPassKernel), which remains for the standalone module and the host-driven loop.
#define SELF_THREADS 1024

__device__ __forceinline__ unsigned blockExclusiveScan1024(unsigned v, unsigned* warpSums, unsigned& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  __syncthreads();   // warpSums may still be read from a previous call
  if (lane == 31) warpSums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const unsigned w = warpSums[lane];
    unsigned wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    warpSums[lane] = wi - w;             // exclusive prefix of the warp sums
    if (lane == 31) warpSums[32] = wi;   // grand total
  }
  __syncthreads();
  total = warpSums[32];
  return warpSums[warp] + incl - v;
}

// picks the bin holding `rank` out of per-thread counts (c0, c1) of bins (2 tid, 2 tid + 1); result through bcast[0..1]
__device__ __forceinline__ void pickBin2(unsigned c0, unsigned c1, unsigned rank, unsigned* warpSums, unsigned* bcast, unsigned& total) {
  const unsigned excl = blockExclusiveScan1024(c0 + c1, warpSums, total);
  if (rank >= excl && rank < excl + c0) { bcast[0] = 2 * threadIdx.x; bcast[1] = rank - excl; }
  else if (rank >= excl + c0 && rank < excl + c0 + c1) { bcast[0] = 2 * threadIdx.x + 1; bcast[1] = rank - excl - c0; }
  __syncthreads();
}

__global__ void __launch_bounds__(SELF_THREADS, 2) selectFusedKernel(const IcpDevParams* __restrict__ DP, SelState* st) {
  __shared__ unsigned hist[2048];
  __shared__ unsigned warpSums[40];
  __shared__ unsigned bcast[2];
  __shared__ bool last;
  traceStamp(DP->trace, 2);
  const float* __restrict__ d2 = DP->d2;
  const int64_t count = DP->n;
  const float ratio = DP->trimRatio;
  unsigned* g0 = DP->fhist0;
  unsigned* g1 = DP->fhist1;
  unsigned* cnt = DP->fcnt;
  unsigned* cand = DP->cand;
  const int tid = threadIdx.x;
  // ---- pass 0 bucket, from the histogram the match kernel accumulated
  unsigned nF;
  pickBin2(__ldcg(&g0[2 * tid]), __ldcg(&g0[2 * tid + 1]), 0xffffffffu, warpSums, bcast, nF);   // rank unknown yet: only the total
  if (nF == 0) {   // no finite match at all (uniform over the grid): "No matches available for computing distance quantiles"
    if (blockIdx.x == 0 && tid == 0) { st->nFinite = 0; st->error = 1; }
    return;        // fhist0 is all zero already, nothing was compacted
  }
  // rank = size_t(float(n) * quantile) — float multiply + truncation (Matches.cpp:85); quantile 1 -> max (:83)
  unsigned rank = (ratio == 1.0f) ? nF - 1 : (unsigned)__fmul_rn((float)nF, ratio);
  if (rank > nF - 1) rank = nF - 1;
  unsigned dummy;
  pickBin2(__ldcg(&g0[2 * tid]), __ldcg(&g0[2 * tid + 1]), rank, warpSums, bcast, dummy);
  const unsigned digit0 = bcast[0], rank0 = bcast[1];
  // ---- pass 1 over this block's share + compaction of the pass-0 bucket
  hist[2 * tid] = 0;
  hist[2 * tid + 1] = 0;
  __syncthreads();
  for (int64_t base = (int64_t)blockIdx.x * SELF_THREADS; base < count; base += (int64_t)gridDim.x * SELF_THREADS) {
    const int64_t i = base + tid;
    const unsigned v = (i < count) ? __float_as_uint(d2[i]) : 0x7f800000u;
    const bool match = (v != 0x7f800000u) && (v >> 21) == digit0;   // dists != inf (Matches.cpp:70)
    if (match) atomicAdd(&hist[(v >> 10) & 0x7ffu], 1u);
    const unsigned m = __ballot_sync(0xffffffffu, match);
    if (m) {
      const int lane = tid & 31, leader = __ffs(m) - 1;
      unsigned slot = 0;
      if (lane == leader) slot = atomicAdd(&cnt[0], (unsigned)__popc(m));
      slot = __shfl_sync(0xffffffffu, slot, leader) + __popc(m & ((1u << lane) - 1u));
      if (match) cand[slot] = v;
    }
  }
  __syncthreads();
  if (hist[2 * tid]) atomicAdd(&g1[2 * tid], hist[2 * tid]);
  if (hist[2 * tid + 1]) atomicAdd(&g1[2 * tid + 1], hist[2 * tid + 1]);
  __threadfence();
  __syncthreads();
  if (tid == 0) last = (atomicAdd(&cnt[1], 1u) == gridDim.x - 1);
  __syncthreads();
  if (!last) return;
  __threadfence();
  // ---- the last block: pass-1 digit, then pass 2 over the candidates
  pickBin2(__ldcg(&g1[2 * tid]), __ldcg(&g1[2 * tid + 1]), rank0, warpSums, bcast, dummy);
  const unsigned digit1 = bcast[0], rank1 = bcast[1];
  hist[2 * tid] = 0;
  hist[2 * tid + 1] = 0;
  __syncthreads();
  const unsigned nc = __ldcg(&cnt[0]);
  for (unsigned i = tid; i < nc; i += SELF_THREADS) {
    const unsigned v = __ldcg(&cand[i]);
    if (((v >> 10) & 0x7ffu) == digit1) atomicAdd(&hist[v & 0x3ffu], 1u);
  }
  __syncthreads();
  pickBin2(hist[2 * tid], hist[2 * tid + 1], rank1, warpSums, bcast, dummy);   // bins >= 1024 are empty
  const unsigned digit2 = bcast[0];
  if (tid == 0) {
    st->nFinite = nF;
    st->error = 0;
    st->tauBits = (digit0 << 21) | (digit1 << 10) | digit2;
    cnt[0] = 0;
    cnt[1] = 0;
  }
  g0[2 * tid] = 0; g0[2 * tid + 1] = 0;   // ready for the next iteration
  g1[2 * tid] = 0; g1[2 * tid + 1] = 0;
}

struct ReduceParams {
  const float4* reading;       // pre-iteration reading (already in the refMean frame)
  const float4* readingNrm;    // nullable
  const float4* refPts;        // sorted reference points
  const float4* refNrm;        // sorted reference normals
  const int* ids;
  const float* d2;
  const int* pos;
  int64_t n;
  int k;
  int hasT;
  Rigid12 T;
  int useTrim;
  int useAngle;
  float cosAngle;              // eps = cos(maxAngle) in float (OutlierFiltersImpl.cpp:228)
  int force2D;
  const float* w;              // explicit OutlierWeights (standalone-module lane); nullable. When set, the gates are skipped.
};

__device__ __forceinline__ double warpSum(double v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Warp reduction of N values per lane by recursive halving: at each butterfly step a lane keeps one half of its values and
// trades the other half with its partner, so 36 moments cost 18 + 9 + 5 + 3 + 2 = 37 shuffles instead of 36 x 5.  After the
// last step lane L holds the warp totals of the index range [lo, lo + len) (len <= 2); the ranges of the 32 lanes tile [0, N).
template <int N>
__device__ __forceinline__ void halveStep(const double (&a)[N], double (&b)[(N + 1) / 2], int off, int& lo, int& len) {
  constexpr int H = (N + 1) / 2;
  const bool upper = (threadIdx.x & off) != 0;
#pragma unroll
  for (int j = 0; j < H; ++j) {
    const double lowV = a[j];
    const double highV = (H + j < N) ? a[H + j] : 0.0;
    const double recv = __shfl_xor_sync(0xffffffffu, upper ? lowV : highV, off);
    b[j] = (upper ? highV : lowV) + recv;
  }
  if (upper) { lo += H; len -= H; } else len = (len < H) ? len : H;
  if (len < 0) len = 0;
}

// one kept pair -> the 36 un-centred moments
__device__ __forceinline__ void accumulatePair(double* m, const ReduceParams& P, float3 r, const float4 s4, const float4 nn4, double w) {
  double rx = r.x, ry = r.y, rz = r.z, sx = s4.x, sy = s4.y, sz = s4.z, nx = nn4.x, ny = nn4.y, nz = nn4.z;
  if (P.force2D) { rz = 0.0; sz = 0.0; nz = 0.0; }
  const double ux = ry * nz - rz * ny, uy = rz * nx - rx * nz, uz = rx * ny - ry * nx;   // u = r x n
  const double e = (rx - sx) * nx + (ry - sy) * ny + (rz - sz) * nz;
  m[M_CNT] += 1.0;
  m[M_SPARE] += w;   // sum of weights (ErrorMinimizer.cpp:123)
  m[M_R] += rx; m[M_R + 1] += ry; m[M_R + 2] += rz;
  m[M_S] += sx; m[M_S + 1] += sy; m[M_S + 2] += sz;
  m[M_UU] += ux * ux; m[M_UU + 1] += ux * uy; m[M_UU + 2] += ux * uz; m[M_UU + 3] += uy * uy; m[M_UU + 4] += uy * uz; m[M_UU + 5] += uz * uz;
  m[M_UN] += ux * nx; m[M_UN + 1] += ux * ny; m[M_UN + 2] += ux * nz;
  m[M_UN + 3] += uy * nx; m[M_UN + 4] += uy * ny; m[M_UN + 5] += uy * nz;
  m[M_UN + 6] += uz * nx; m[M_UN + 7] += uz * ny; m[M_UN + 8] += uz * nz;
  m[M_NN] += nx * nx; m[M_NN + 1] += nx * ny; m[M_NN + 2] += nx * nz; m[M_NN + 3] += ny * ny; m[M_NN + 4] += ny * nz; m[M_NN + 5] += nz * nz;
  m[M_UE] += ux * e; m[M_UE + 1] += uy * e; m[M_UE + 2] += uz * e;
  m[M_NE] += nx * e; m[M_NE + 1] += ny * e; m[M_NE + 2] += nz * e;
}

// gates of one (reading point, neighbour slot) pair: true = the pair enters the normal equations.  On true, pos is valid.
__device__ __forceinline__ bool pairKept(const ReduceParams& P, int64_t o, float d, float tau, double* m) {
  if (d == CUDART_INF_F) return false;          // ErrorMinimizer.cpp:103-106
  m[M_VALID] += 1.0;
  if (P.w) return P.w[o] != 0.f;                 // ErrorMinimizer.cpp:108
  return !(P.useTrim && !(d <= tau));            // OutlierFiltersImpl.cpp:146
}

// Grid-stride over (reading point, neighbour slot) pairs, two pairs in flight per thread so the dependent gathers
// (d2 -> sorted position -> reference point + normal) of one overlap those of the other.  Weights are binary and act
// only through compaction — exactly the fork's A = G G^T, b = -G h^T (PointToPlane.cpp:153-155).  Moments are
// un-centred; the matched-mean centring of PointToPlane.cpp:263-268 is applied algebraically afterwards.
// Reduction: 36 fp64 moments per thread -> warp shuffle -> block tree -> per-block partial -> the last block adds the
// partials in a fixed order (deterministic for a given grid).
__global__ void __launch_bounds__(RED_THREADS, 512 / RED_THREADS) reduceKernel(ReduceParams P, const SelState* __restrict__ st, double* __restrict__ partials,
                                                    unsigned* ticket, double* __restrict__ result, const IcpDevParams* __restrict__ DP,
                                                    const IcpDevState* __restrict__ DS) {
  if (DP) {   // graph-driven loop: arguments from the device-resident blocks, T_iter from the loop state
    P.reading = DP->reading; P.readingNrm = DP->readingNrm; P.refPts = DP->pts; P.refNrm = DP->nrm;
    P.ids = DP->ids; P.d2 = DP->d2; P.pos = DP->pos; P.n = DP->n; P.k = 1; P.hasT = 1;
#pragma unroll
    for (int j = 0; j < 12; ++j) P.T.m[j] = DS->T[j];
    P.useTrim = DP->useTrim; P.useAngle = DP->useAngle; P.cosAngle = DP->cosAngle; P.force2D = DP->force2D; P.w = nullptr;
    partials = DP->partials;
    traceStamp(DP->trace, 5);
  }
  double m[NUM_MOM];
#pragma unroll
  for (int j = 0; j < NUM_MOM; ++j) m[j] = 0.0;
  const float tau = P.useTrim ? __uint_as_float(st->tauBits) : CUDART_INF_F;
  const int64_t total = P.n * P.k;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const bool angleGate = !P.w && P.useAngle && P.readingNrm;
  for (int64_t o0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o0 < total; o0 += 2 * stride) {
    const int64_t o1 = o0 + stride;
    const bool has1 = o1 < total;
    const float d0 = P.d2[o0];
    const float d1 = has1 ? P.d2[o1] : CUDART_INF_F;
    bool k0 = pairKept(P, o0, d0, tau, m);
    bool k1 = has1 && pairKept(P, o1, d1, tau, m);
    const int p0 = k0 ? P.pos[o0] : -1, p1 = k1 ? P.pos[o1] : -1;
    k0 = k0 && p0 >= 0;
    k1 = k1 && p1 >= 0;
    const int64_t i0 = o0 / P.k, i1 = has1 ? o1 / P.k : 0;
    float4 q0 = make_float4(0.f, 0.f, 0.f, 1.f), q1 = q0, n0 = q0, n1 = q0, s0 = q0, s1 = q0, a0 = q0, a1 = q0;
    if (k0) { q0 = P.reading[i0]; n0 = P.refNrm[p0]; s0 = P.refPts[p0]; if (angleGate) a0 = P.readingNrm[i0]; }
    if (k1) { q1 = P.reading[i1]; n1 = P.refNrm[p1]; s1 = P.refPts[p1]; if (angleGate) a1 = P.readingNrm[i1]; }
    if (k0) {
      float3 r = make_float3(q0.x, q0.y, q0.z);
      if (P.hasT) r = xformPoint(P.T, q0.x, q0.y, q0.z, q0.w);
      bool keep = true;
      if (angleGate) {          // OutlierFiltersImpl.cpp:259-264
        float3 a = make_float3(a0.x, a0.y, a0.z);
        if (P.hasT) a = xformNormal(P.T, a0.x, a0.y, a0.z);
        keep = !(__fadd_rn(__fadd_rn(__fmul_rn(a.x, n0.x), __fmul_rn(a.y, n0.y)), __fmul_rn(a.z, n0.z)) < P.cosAngle);
      }
      if (keep) accumulatePair(m, P, r, s0, n0, P.w ? (double)P.w[o0] : 1.0);
    }
    if (k1) {
      float3 r = make_float3(q1.x, q1.y, q1.z);
      if (P.hasT) r = xformPoint(P.T, q1.x, q1.y, q1.z, q1.w);
      bool keep = true;
      if (angleGate) {
        float3 a = make_float3(a1.x, a1.y, a1.z);
        if (P.hasT) a = xformNormal(P.T, a1.x, a1.y, a1.z);
        keep = !(__fadd_rn(__fadd_rn(__fmul_rn(a.x, n1.x), __fmul_rn(a.y, n1.y)), __fmul_rn(a.z, n1.z)) < P.cosAngle);
      }
      if (keep) accumulatePair(m, P, r, s1, n1, P.w ? (double)P.w[o1] : 1.0);
    }
  }
  // warp shuffle -> block tree (shared) -> per-block partial
  __shared__ double sm[RED_THREADS / 32][NUM_MOM];
  __shared__ bool last;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  {
    double a18[18], a9[9], a5[5], a3[3], a2[2];
    int lo = 0, len = NUM_MOM;
    halveStep<36>(m, a18, 16, lo, len);
    halveStep<18>(a18, a9, 8, lo, len);
    halveStep<9>(a9, a5, 4, lo, len);
    halveStep<5>(a5, a3, 2, lo, len);
    halveStep<3>(a3, a2, 1, lo, len);
    if (len > 0) sm[warp][lo] = a2[0];
    if (len > 1) sm[warp][lo + 1] = a2[1];
    (void)lane;
  }
  __syncthreads();
  if (threadIdx.x < NUM_MOM) {
    double v = 0.0;
    for (int w = 0; w < RED_THREADS / 32; ++w) v += sm[w][threadIdx.x];
    partials[(size_t)blockIdx.x * NUM_MOM + threadIdx.x] = v;
  }
  __threadfence();
  if (threadIdx.x == 0) last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!last) return;
  __threadfence();
  // the last block: 7 slices x 36 moments; slice s adds partials s, s+7, ... in order, then the slices are added in order
  constexpr int SL = RED_THREADS / NUM_MOM;   // slices of NUM_MOM threads
  const int j = threadIdx.x % NUM_MOM, sl = threadIdx.x / NUM_MOM;
  if (sl < SL) {
    double v = 0.0;
#pragma unroll 4
    for (unsigned b = sl; b < gridDim.x; b += SL) v += __ldcg(&partials[(size_t)b * NUM_MOM + j]);
    sm[sl][j] = v;
  }
  __syncthreads();
  if (threadIdx.x < NUM_MOM) {
    double v = 0.0;
    for (int s = 0; s < SL; ++s) v += sm[s][threadIdx.x];
    result[threadIdx.x] = v;
  }
  if (threadIdx.x == 0) *ticket = 0;
}
