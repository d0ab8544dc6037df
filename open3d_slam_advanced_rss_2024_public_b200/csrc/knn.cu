// csrc/knn.cu — KDTreeMatcher::findClosests replacement (MatchersImpl.cpp:117-132): exact k-NN by neighbour-cell scan
// over the radix-sorted voxel grid, fused with the per-iteration RigidTransformation (ICP.cpp:409-413).
//
// Contract: d2 = ((dx*dx + dy*dy) + dz*dz) with every fp32 op rounded; candidates need d2 < maxDist^2;
// neighbours ascending by (d2, original index) — ties -> smallest index; padding (-1, +inf); epsilon treated as 0.
//
// Exactness of the pruning: a cell's lower bound is evaluated with the SAME rounded operations on per-axis gaps that
// are (after subtracting GridDesc::slack) never larger than the true per-axis separation, so by monotonicity of
// round-to-nearest the bound never exceeds the computed d2 of any point in the cell.  Cells whose bound equals the
// current k-th distance are still visited (a tie with a smaller index may hide there).
#include <cfloat>
#include <cmath>

#include "common.cuh"

namespace {

template <int K>
struct TopK {
  float d[K];
  int id[K];
  int pos[K];
  int k;       // requested k (<= K)
  int count;
  __device__ __forceinline__ void init(int k_) {
    k = k_;
    count = 0;
#pragma unroll
    for (int j = 0; j < K; ++j) { d[j] = CUDART_INF_F; id[j] = -1; pos[j] = -1; }
  }
  __device__ __forceinline__ bool full() const { return count == k; }
  __device__ __forceinline__ float worst() const {
    if (K == 1) return d[0];
    float w = d[0];
#pragma unroll
    for (int j = 1; j < K; ++j) if (j == k - 1) w = d[j];
    return (k == 1) ? d[0] : w;
  }
  __device__ __forceinline__ void offer(float dd, int ii, int pp) {
    if (K == 1) {
      if (dd < d[0] || (dd == d[0] && (unsigned)ii < (unsigned)id[0])) { d[0] = dd; id[0] = ii; pos[0] = pp; count = 1; }
      return;
    }
    // position of insertion = number of entries strictly better than (dd, ii)
    int ins = 0;
#pragma unroll
    for (int j = 0; j < K; ++j) ins += (j < count && (d[j] < dd || (d[j] == dd && id[j] < ii))) ? 1 : 0;
    if (ins >= k) return;
#pragma unroll
    for (int j = K - 1; j > 0; --j)
      if (j > ins && j < k) { d[j] = d[j - 1]; id[j] = id[j - 1]; pos[j] = pos[j - 1]; }
#pragma unroll
    for (int j = 0; j < K; ++j)
      if (j == ins) { d[j] = dd; id[j] = ii; pos[j] = pp; }
    if (count < k) ++count;
  }
};

struct AxisGeom {
  float q, o;   // query coordinate, grid origin
  int c;        // query's (unclamped) cell coordinate
};

// conservative gap between the query and cell `i` along one axis (0 for the query's own cell)
__device__ __forceinline__ float axisGap(const AxisGeom& a, int i, float cell, float slack) {
  if (i == a.c) return 0.f;
  float g;
  if (i > a.c) g = __fsub_rn(__fadd_rn(a.o, __fmul_rn((float)i, cell)), a.q);           // lo(i) - q
  else         g = __fsub_rn(a.q, __fadd_rn(a.o, __fmul_rn((float)(i + 1), cell)));     // q - hi(i)
  return fmaxf(__fsub_rn(g, slack), 0.f);
}
__device__ __forceinline__ float lb3(float gx, float gy, float gz) {
  return __fadd_rn(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gy, gy)), __fmul_rn(gz, gz));
}

template <int K>
__global__ void __launch_bounds__(128) knnKernel(const float4* __restrict__ pts, const uint32_t* __restrict__ cellStart, GridDesc g,
                                                 const float4* __restrict__ queries, int64_t n, int hasT, Rigid12 T, int k, float maxd2,
                                                 int allowSelf, int* __restrict__ outIds, float* __restrict__ outD2,
                                                 int* __restrict__ outPos, unsigned long long* __restrict__ visitCounter) {
  const int64_t qi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned visits = 0;
  if (qi < n) {
    const float4 q4 = queries[qi];
    float3 q = make_float3(q4.x, q4.y, q4.z);
    if (hasT) q = xformPoint(T, q4.x, q4.y, q4.z, q4.w);
    TopK<K> best;
    best.init(k);
    const bool finite = (fabsf(q.x) <= FLT_MAX) && (fabsf(q.y) <= FLT_MAX) && (fabsf(q.z) <= FLT_MAX);
    if (finite) {
      const float lim = 1.0e9f;
      AxisGeom ax{q.x, g.ox, (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.x, g.ox), g.inv_cell), -lim), lim))};
      AxisGeom ay{q.y, g.oy, (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.y, g.oy), g.inv_cell), -lim), lim))};
      AxisGeom az{q.z, g.oz, (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.z, g.oz), g.inv_cell), -lim), lim))};
      // first ring that touches the grid, last ring that still does
      const int r0 = max(max(max(-ax.c, ax.c - (g.nx - 1)), max(-ay.c, ay.c - (g.ny - 1))), max(max(-az.c, az.c - (g.nz - 1)), 0));
      const int rEnd = max(max(max(ax.c, g.nx - 1 - ax.c), max(ay.c, g.ny - 1 - ay.c)), max(az.c, g.nz - 1 - az.c));
      for (int r = r0; r <= rEnd; ++r) {
        if (r > 0) {
          // every cell of ring r has |d| == r on some axis => its gap there >= the matching face gap
          const float f = fminf(fminf(fminf(axisGap(ax, ax.c - r, g.cell, g.slack), axisGap(ax, ax.c + r, g.cell, g.slack)),
                                      fminf(axisGap(ay, ay.c - r, g.cell, g.slack), axisGap(ay, ay.c + r, g.cell, g.slack))),
                                fminf(axisGap(az, az.c - r, g.cell, g.slack), axisGap(az, az.c + r, g.cell, g.slack)));
          const float ringLb = lb3(0.f, 0.f, f);
          if (ringLb >= maxd2) break;
          if (best.full() && ringLb > best.worst()) break;
        }
        const int z0 = max(az.c - r, 0), z1 = min(az.c + r, g.nz - 1);
        const int y0 = max(ay.c - r, 0), y1 = min(ay.c + r, g.ny - 1);
        for (int iz = z0; iz <= z1; ++iz) {
          const float gz = axisGap(az, iz, g.cell, g.slack);
          const int adz = abs(iz - az.c);
          for (int iy = y0; iy <= y1; ++iy) {
            const float gy = axisGap(ay, iy, g.cell, g.slack);
            const float rowLb = lb3(0.f, gy, gz);
            if (rowLb >= maxd2) continue;
            if (best.full() && rowLb > best.worst()) continue;
            const bool shellRow = (adz == r) || (abs(iy - ay.c) == r);
            // x reach implied by the current bound (conservative: +1 cell on both sides)
            const float bound = best.full() ? best.worst() : maxd2;
            int xr = r;
            if (bound < CUDART_INF_F) {
              const float reach = __fadd_rn(sqrtf(fmaxf(__fsub_rn(bound, rowLb), 0.f)), g.slack);
              const float cellsF = fminf(__fmul_rn(reach, g.inv_cell), 1.0e9f);
              xr = min(r, (int)cellsF + 2);
            }
            const size_t rowBase = (size_t)g.nx * ((size_t)iy + (size_t)g.ny * (size_t)iz);
#pragma unroll 1
            for (int seg = 0; seg < 2; ++seg) {
              int xa, xb;
              if (shellRow) {
                if (seg == 1) break;
                xa = ax.c - xr; xb = ax.c + xr;
              } else {   // interior row of the shell: only the two end cells (they are beyond reach if xr < r)
                if (xr < r) break;
                if (r == 0 && seg == 1) break;
                xa = xb = (seg == 0) ? ax.c - r : ax.c + r;
              }
              xa = max(xa, 0);
              xb = min(xb, g.nx - 1);
              if (xa > xb) continue;
              const uint32_t s = cellStart[rowBase + xa], e = cellStart[rowBase + xb + 1];
              for (uint32_t j = s; j < e; ++j) {
                const float4 p = pts[j];
                const int id = __float_as_int(p.w);
                if (!allowSelf && id == (int)qi) continue;
                const float d = dist2_rn(q.x, q.y, q.z, p.x, p.y, p.z);
                ++visits;
                if (d < maxd2) best.offer(d, id, (int)j);
              }
            }
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < K; ++j)
      if (j < k) {
        outIds[(int64_t)k * qi + j] = best.id[j];
        outD2[(int64_t)k * qi + j] = best.d[j];
        outPos[(int64_t)k * qi + j] = best.pos[j];
      }
  }
  // touched-point statistic (Matcher::visitCounter)
  for (int o = 16; o > 0; o >>= 1) visits += __shfl_xor_sync(0xffffffffu, visits, o);
  if ((threadIdx.x & 31) == 0 && visits) atomicAdd(visitCounter, (unsigned long long)visits);
}

// ---- k = 1 (the ICP matcher): eight lanes per query ---------------------------------------------------------------
// One thread per query leaves a B200 two-thirds empty at 131 072 queries and serialises ~20 dependent loads per
// query (profiles/r1a_summary.md).  Here a query is owned by an aligned group of 8 lanes:
//   * the 9 (y,z) rows of the 3x3x3 neighbourhood are 9 contiguous point ranges; their 18 directory entries are
//     fetched in ONE round of independent loads (centre row by every lane, the other 8 rows one per lane);
//   * the centre row is scanned cooperatively (lane-strided float4 loads), the group minimum prunes the other rows
//     by their exact fp32 lower bounds, survivors are scanned cooperatively as well;
//   * rings >= 2 (sparse neighbourhoods, large initial misalignment) distribute the (y,z) rows of the shell over
//     the 8 lanes, each lane walking its rows like the general kernel.
// The result is the lexicographic minimum of (d2, index) over every candidate with d2 < maxDist^2 that pruning cannot
// exclude, so it is identical to the one-thread kernel and to the oracle (ties -> smallest index).
struct Best1 {
  float d;
  int id, pos;
};
__device__ __forceinline__ void offer1(Best1& b, float d, int id, int pos) {
  if (d < b.d || (d == b.d && (unsigned)id < (unsigned)b.id)) { b.d = d; b.id = id; b.pos = pos; }
}
__device__ __forceinline__ void groupMin8(Best1& b) {
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) {
    const float d = __shfl_xor_sync(0xffffffffu, b.d, o);
    const int id = __shfl_xor_sync(0xffffffffu, b.id, o);
    const int pos = __shfl_xor_sync(0xffffffffu, b.pos, o);
    offer1(b, d, id, pos);
  }
}
__device__ __forceinline__ void test1(const float4 p, uint32_t j, const float3 q, float maxd2, int selfId, Best1& b, unsigned& visits) {
  const int id = __float_as_int(p.w);
  if (id == selfId) return;
  const float d = dist2_rn(q.x, q.y, q.z, p.x, p.y, p.z);
  ++visits;
  if (d < maxd2) offer1(b, d, id, (int)j);
}
// lanes of a group stride through [s, e); two independent loads in flight per lane
__device__ __forceinline__ void scanCoop8(const float4* __restrict__ pts, uint32_t s, uint32_t e, int l, const float3 q, float maxd2,
                                          int selfId, Best1& b, unsigned& visits) {
  for (uint32_t j = s + l; j < e; j += 16) {
    const bool h1 = (j + 8 < e);
    const float4 p0 = pts[j];
    const float4 p1 = pts[h1 ? j + 8 : j];
    test1(p0, j, q, maxd2, selfId, b, visits);
    if (h1) test1(p1, j + 8, q, maxd2, selfId, b, visits);
  }
}
__device__ __forceinline__ float ringLowerBound(const AxisGeom& ax, const AxisGeom& ay, const AxisGeom& az, int r, float cell, float slack) {
  // every cell of ring r has |d| == r on some axis => its gap there >= the matching face gap
  const float f = fminf(fminf(fminf(axisGap(ax, ax.c - r, cell, slack), axisGap(ax, ax.c + r, cell, slack)),
                              fminf(axisGap(ay, ay.c - r, cell, slack), axisGap(ay, ay.c + r, cell, slack))),
                        fminf(axisGap(az, az.c - r, cell, slack), axisGap(az, az.c + r, cell, slack)));
  return lb3(0.f, 0.f, f);
}

__global__ void __launch_bounds__(256) knn1GroupKernel(const float4* __restrict__ pts, const uint32_t* __restrict__ cellStart, GridDesc g,
                                                       const float4* __restrict__ queries, int64_t n, int hasT, Rigid12 T,
                                                       const float* __restrict__ Tdev, float maxd2, int allowSelf,
                                                       int* __restrict__ outIds, float* __restrict__ outD2, int* __restrict__ outPos,
                                                       unsigned long long* __restrict__ visitCounter) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31, l = lane & 7, gbase = lane & 24;
  const int64_t qi = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  const bool live = qi < n;
  if (Tdev) {   // device-resident iteration transform (graph-driven loop): rows 0..2 of the row-major 4x4
#pragma unroll
    for (int j = 0; j < 12; ++j) T.m[j] = Tdev[j];
    hasT = 1;
  }
  float4 q4 = make_float4(0.f, 0.f, 0.f, 1.f);
  if (live) q4 = queries[qi];
  float3 q = make_float3(q4.x, q4.y, q4.z);
  if (hasT) q = xformPoint(T, q4.x, q4.y, q4.z, q4.w);
  const bool finite = live && (fabsf(q.x) <= FLT_MAX) && (fabsf(q.y) <= FLT_MAX) && (fabsf(q.z) <= FLT_MAX);
  const int selfId = allowSelf ? -2 : (int)qi;
  const float lim = 1.0e9f;
  const AxisGeom ax{q.x, g.ox, finite ? (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.x, g.ox), g.inv_cell), -lim), lim)) : 0};
  const AxisGeom ay{q.y, g.oy, finite ? (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.y, g.oy), g.inv_cell), -lim), lim)) : 0};
  const AxisGeom az{q.z, g.oz, finite ? (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.z, g.oz), g.inv_cell), -lim), lim)) : 0};
  Best1 best{CUDART_INF_F, -1, -1};
  unsigned visits = 0;

  // ---- rings 0 and 1: nine (y,z) rows, each the contiguous x range [cx-1, cx+1]
  uint32_t sC = 0, eC = 0, sL = 0, eL = 0;
  float lbL = CUDART_INF_F;
  {
    const int xa = max(ax.c - 1, 0), xb = min(ax.c + 1, g.nx - 1);
    if (finite && xa <= xb) {
      if (ay.c >= 0 && ay.c < g.ny && az.c >= 0 && az.c < g.nz) {
        const size_t base = (size_t)g.nx * ((size_t)ay.c + (size_t)g.ny * (size_t)az.c);
        sC = cellStart[base + xa];
        eC = cellStart[base + xb + 1];
      }
      const int ri = l + (l >= 4 ? 1 : 0);
      const int iy = ay.c + (ri % 3) - 1, iz = az.c + (ri / 3) - 1;
      if (iy >= 0 && iy < g.ny && iz >= 0 && iz < g.nz) {
        const size_t base = (size_t)g.nx * ((size_t)iy + (size_t)g.ny * (size_t)iz);
        sL = cellStart[base + xa];
        eL = cellStart[base + xb + 1];
        lbL = lb3(0.f, axisGap(ay, iy, g.cell, g.slack), axisGap(az, iz, g.cell, g.slack));
      }
    }
  }
  scanCoop8(pts, sC, eC, l, q, maxd2, selfId, best, visits);
  groupMin8(best);
  {
    const bool alive = (eL > sL) && (lbL < maxd2) && !(lbL > best.d);
    unsigned m = (__ballot_sync(FULL, alive) >> gbase) & 0xffu;
    while (__any_sync(FULL, m != 0u)) {
      const int r = m ? (__ffs(m) - 1) : 0;
      const uint32_t s = __shfl_sync(FULL, sL, gbase + r);
      uint32_t e = __shfl_sync(FULL, eL, gbase + r);
      const float lb = __shfl_sync(FULL, lbL, gbase + r);
      if (!m || lb > best.d) e = s;   // lane-local re-prune (its own best only improves)
      m &= m - 1u;
      scanCoop8(pts, s, e, l, q, maxd2, selfId, best, visits);
    }
    groupMin8(best);
  }

  // ---- rings >= 2: the (y,z) rows of each shell are dealt to the 8 lanes
  {
    const int r0 = max(max(max(-ax.c, ax.c - (g.nx - 1)), max(-ay.c, ay.c - (g.ny - 1))), max(max(-az.c, az.c - (g.nz - 1)), 0));
    const int rEnd = max(max(max(ax.c, g.nx - 1 - ax.c), max(ay.c, g.ny - 1 - ay.c)), max(az.c, g.nz - 1 - az.c));
    int r = max(2, r0);
    bool need = finite && r <= rEnd;
    if (need) {
      const float rl = ringLowerBound(ax, ay, az, r, g.cell, g.slack);
      need = (rl < maxd2) && !(rl > best.d);
    }
    while (__any_sync(FULL, need)) {
      if (need) {
        const int w = 2 * r + 1;
        for (int t = l; t < w * w; t += 8) {
          const int dy = t % w - r, dz = t / w - r;
          const int iy = ay.c + dy, iz = az.c + dz;
          if (iy < 0 || iy >= g.ny || iz < 0 || iz >= g.nz) continue;
          const float rowLb = lb3(0.f, axisGap(ay, iy, g.cell, g.slack), axisGap(az, iz, g.cell, g.slack));
          if (rowLb >= maxd2 || rowLb > best.d) continue;
          const bool shellRow = (abs(dz) == r) || (abs(dy) == r);
          const float bound = fminf(best.d, maxd2);
          int xr = r;
          if (bound < CUDART_INF_F) {   // x reach implied by the current bound (conservative: +1 cell on both sides)
            const float reach = __fadd_rn(sqrtf(fmaxf(__fsub_rn(bound, rowLb), 0.f)), g.slack);
            xr = min(r, (int)fminf(__fmul_rn(reach, g.inv_cell), 1.0e9f) + 2);
          }
          const size_t rowBase = (size_t)g.nx * ((size_t)iy + (size_t)g.ny * (size_t)iz);
#pragma unroll 1
          for (int seg = 0; seg < 2; ++seg) {
            int xa, xb;
            if (shellRow) {
              if (seg == 1) break;
              xa = ax.c - xr; xb = ax.c + xr;
            } else {   // interior row of the shell: only its two end cells (beyond reach if xr < r)
              if (xr < r) break;
              xa = xb = (seg == 0) ? ax.c - r : ax.c + r;
            }
            xa = max(xa, 0);
            xb = min(xb, g.nx - 1);
            if (xa > xb) continue;
            const uint32_t s = cellStart[rowBase + xa], e = cellStart[rowBase + xb + 1];
            for (uint32_t j = s; j < e; ++j) test1(pts[j], j, q, maxd2, selfId, best, visits);
          }
        }
      }
      groupMin8(best);
      if (need) {
        ++r;
        need = r <= rEnd;
        if (need) {
          const float rl = ringLowerBound(ax, ay, az, r, g.cell, g.slack);
          need = (rl < maxd2) && !(rl > best.d);
        }
      }
    }
  }
  if (live && l == 0) {
    outIds[qi] = best.id;
    outD2[qi] = best.d;
    outPos[qi] = best.pos;
  }
  // touched-point statistic (Matcher::visitCounter): warp shuffle -> shared -> one atomic per block
  __shared__ unsigned blockVisits;
  if (threadIdx.x == 0) blockVisits = 0;
  __syncthreads();
  for (int o = 16; o > 0; o >>= 1) visits += __shfl_xor_sync(FULL, visits, o);
  if (lane == 0 && visits) atomicAdd(&blockVisits, visits);
  __syncthreads();
  if (threadIdx.x == 0 && blockVisits) atomicAdd(visitCounter, (unsigned long long)blockVisits);
}

template <int K>
void launchKnn(b200reg_ctx* ctx, const b200reg_index* ix, const float4* q, int64_t n, int hasT, const Rigid12& T, int k, float maxd2,
               int allowSelf, int* ids, float* d2, int* pos, unsigned long long* visits) {
  knnKernel<K><<<divUp(n, 128), 128, 0, ctx->stream>>>(ix->pts, ix->cellStart, ix->g, q, n, hasT, T, k, maxd2, allowSelf, ids, d2, pos, visits);
}

}  // namespace

// Device-side entry shared with icp.cu / normals.cu.  Results land in ctx->matchIds / matchD2 / matchPos (k*n entries);
// the touched-point counter is dstate[0] (unsigned long long), accumulated.
int runKnn(b200reg_ctx* ctx, const b200reg_index* ix, const float4* queries, int64_t n, const float* T, int k, float maxDist, int allowSelf) {
  if (k < 1 || k > 64) return setErr(ctx, B200REG_E_INVALID_PARAM, "knn: k = %d outside the supported range [1, 64]", k);
  if (!(maxDist >= 0.f)) return setErr(ctx, B200REG_E_INVALID_PARAM, "knn: maxDist must be >= 0");
  TRY(ensureBuf(ctx, ctx->matchIds, sizeof(int) * (size_t)k * n));
  TRY(ensureBuf(ctx, ctx->matchD2, sizeof(float) * (size_t)k * n));
  TRY(ensureBuf(ctx, ctx->matchPos, sizeof(int) * (size_t)k * n));
  if (n == 0) return B200REG_OK;
  Rigid12 R;
  memset(&R, 0, sizeof(R));
  if (T) memcpy(R.m, T, sizeof(R.m));
  const float maxd2 = maxDist * maxDist;   // fp32 product (inf stays inf)
  int* ids = (int*)ctx->matchIds.p;
  float* d2 = (float*)ctx->matchD2.p;
  int* pos = (int*)ctx->matchPos.p;
  unsigned long long* visits = (unsigned long long*)ctx->dstate;
  const int hasT = T ? 1 : 0;
  if (k == 1 && !ctx->knnOneThread)
    knn1GroupKernel<<<divUp(n * 8, 256), 256, 0, ctx->stream>>>(ix->pts, ix->cellStart, ix->g, queries, n, hasT, R, nullptr, maxd2, allowSelf,
                                                                ids, d2, pos, visits);
  else if (k == 1) launchKnn<1>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else if (k <= 4) launchKnn<4>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else if (k <= 8) launchKnn<8>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else if (k <= 16) launchKnn<16>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else if (k <= 32) launchKnn<32>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else launchKnn<64>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  LAUNCH_CHECK(ctx);
  return B200REG_OK;
}

extern "C" int32_t b200reg_knn(b200reg_ctx* ctx, const b200reg_index* index, const b200reg_cloud* query, const float* T, int32_t k,
                               float max_dist, float epsilon, int32_t allow_self, int32_t* ids_out, float* d2_out,
                               uint64_t* visit_count) {
  if (!ctx || !index || !query) return setErr(ctx, B200REG_E_INVALID_PARAM, "knn: bad arguments");
  if (!(epsilon >= 0.f)) return setErr(ctx, B200REG_E_INVALID_PARAM, "knn: epsilon must be >= 0");   // MatchersImpl.h:98
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const int64_t n = query->n;
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->dstate, 0, sizeof(unsigned long long), ctx->stream));
  TRY(runKnn(ctx, index, query->xyz, n, T, k, max_dist, allow_self));
  if (ids_out && n) CUDA_TRY(ctx, cudaMemcpyAsync(ids_out, ctx->matchIds.p, sizeof(int) * (size_t)k * n, cudaMemcpyDeviceToHost, ctx->stream));
  if (d2_out && n) CUDA_TRY(ctx, cudaMemcpyAsync(d2_out, ctx->matchD2.p, sizeof(float) * (size_t)k * n, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pinned, ctx->dstate, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (visit_count) *visit_count = *(const unsigned long long*)ctx->pinned;
  return B200REG_OK;
}
