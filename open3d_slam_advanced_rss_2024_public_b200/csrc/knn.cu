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
#include <algorithm>
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
// the index build's cell function (index.cu cellCoord): monotone non-decreasing in p, clamped into the grid
__device__ __forceinline__ int cellOfClamped(float p, float o, float inv, int nmax) {
  const float c = floorf(__fmul_rn(__fsub_rn(p, o), inv));
  return (int)fminf(fmaxf(c, 0.f), (float)(nmax - 1));
}
__device__ __forceinline__ float lb3(float gx, float gy, float gz) {
  return __fadd_rn(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gy, gy)), __fmul_rn(gz, gz));
}

template <int K>
__global__ void __launch_bounds__(128) knnKernel(const float4* __restrict__ pts, const uint32_t* __restrict__ cellStart, const GridDesc* __restrict__ gp,
                                                 const float4* __restrict__ queries, int64_t n, int hasT, Rigid12 T, int k, float maxd2,
                                                 int allowSelf, int* __restrict__ outIds, float* __restrict__ outD2,
                                                 int* __restrict__ outPos, unsigned long long* __restrict__ visitCounter) {
  const int64_t qi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const GridDesc g = *gp;   // device-derived grid (index.cu)
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
  unsigned long long key;   // (bit pattern of d2) << 32 | original index: unsigned order == (d2, index) lexicographic order (d2 >= 0)
  int pos;                  // sorted position of the holder (feeds the gather of the reduction)
  __device__ __forceinline__ float d() const { return __uint_as_float((unsigned)(key >> 32)); }
};
__device__ __forceinline__ unsigned long long packKey(float d, int id) {
  return ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)id;
}
template <int G>
__device__ __forceinline__ void groupMin(Best1& b) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) {
    const unsigned long long k = __shfl_xor_sync(0xffffffffu, b.key, o);
    const int pos = __shfl_xor_sync(0xffffffffu, b.pos, o);
    if (k < b.key) { b.key = k; b.pos = pos; }
  }
}
__device__ __forceinline__ float4 ldPoint(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
template <bool SELFCHK>
__device__ __forceinline__ void test1(const float4 p, uint32_t j, const float3 q, int selfId, Best1& b, unsigned& visits) {
  const int id = __float_as_int(p.w);
  if (SELFCHK && id == selfId) return;
  const unsigned long long k = packKey(dist2_rn(q.x, q.y, q.z, p.x, p.y, p.z), id);
  ++visits;
  if (k < b.key) { b.key = k; b.pos = (int)j; }
}
// lanes of a group stride through [s, e); two independent loads in flight per lane
template <int G, bool SELFCHK>
__device__ __forceinline__ void scanCoop(const float4* __restrict__ pts, uint32_t s, uint32_t e, int l, const float3 q, int selfId, Best1& b,
                                         unsigned& visits) {
  for (uint32_t j = s + l; j < e; j += 2 * G) {
    const bool h1 = (j + G < e);
    // both loads are issued before either is used (volatile asm keeps the compiler from sinking the second one
    // behind the first candidate's compare)
    const float4 p0 = ldPoint(pts + j);
    const float4 p1 = ldPoint(pts + (h1 ? j + G : j));
    test1<SELFCHK>(p0, j, q, selfId, b, visits);
    if (h1) test1<SELFCHK>(p1, j + G, q, selfId, b, visits);
  }
}
__device__ __forceinline__ float ringLowerBound(const AxisGeom& ax, const AxisGeom& ay, const AxisGeom& az, int r, float cell, float slack) {
  // every cell of ring r has |d| == r on some axis => its gap there >= the matching face gap
  const float f = fminf(fminf(fminf(axisGap(ax, ax.c - r, cell, slack), axisGap(ax, ax.c + r, cell, slack)),
                              fminf(axisGap(ay, ay.c - r, cell, slack), axisGap(ay, ay.c + r, cell, slack))),
                        fminf(axisGap(az, az.c - r, cell, slack), axisGap(az, az.c + r, cell, slack)));
  return lb3(0.f, 0.f, f);
}

// One query of an aligned G-lane group (G = 4 or 8); SELFCHK: exclude the query's own index (allowSelf == 0).  Every lane
// of the warp must call this together (the group loops are closed with warp-wide votes).
template <int G, bool SELFCHK>
__device__ __forceinline__ void knn1Query(const float4* __restrict__ pts, const uint32_t* __restrict__ cellStart, const GridDesc& g,
                                          const float4* __restrict__ queries, int64_t n, int64_t qi, int hasT, const Rigid12& T, float maxd2,
                                          int* __restrict__ outIds, float* __restrict__ outD2, int* outPos, unsigned& visits, unsigned& farQ,
                                          unsigned* __restrict__ sHist, const int* seedPos) {
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int NR = 8 / G;   // non-centre rows owned by each lane
  const int lane = threadIdx.x & 31, l = lane & (G - 1), gbase = lane & (32 - G);
  const bool live = qi < n;
  float4 q4 = make_float4(0.f, 0.f, 0.f, 1.f);
  if (live) q4 = queries[qi];
  float3 q = make_float3(q4.x, q4.y, q4.z);
  if (hasT) q = xformPoint(T, q4.x, q4.y, q4.z, q4.w);
  const bool finite = live && (fabsf(q.x) <= FLT_MAX) && (fabsf(q.y) <= FLT_MAX) && (fabsf(q.z) <= FLT_MAX);
  const int selfId = (int)qi;
  const float lim = 1.0e9f;
  const AxisGeom ax{q.x, g.ox, finite ? (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.x, g.ox), g.inv_cell), -lim), lim)) : 0};
  const AxisGeom ay{q.y, g.oy, finite ? (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.y, g.oy), g.inv_cell), -lim), lim)) : 0};
  const AxisGeom az{q.z, g.oz, finite ? (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.z, g.oz), g.inv_cell), -lim), lim)) : 0};
  // a candidate must beat (maxDist^2, index 0): d2 < maxDist^2 strictly, any index (NaN patterns sort above +inf and never win)
  const unsigned long long initKey = packKey(maxd2, 0);
  Best1 best{initKey, -1};
  // Seed (iterations >= 2 of a registration): the point this query matched in the previous iteration, re-measured under the new
  // transform.  It is an ordinary candidate (same key, same tie rule), so the result is unchanged — but the search starts with a
  // bound that is usually the answer itself: rows beyond it are never fetched and the x range of the others shrinks to 1-2 cells.
  if (seedPos && finite) {
    const int sp = seedPos[qi];
    if (sp >= 0) test1<SELFCHK>(ldPoint(pts + sp), (uint32_t)sp, q, selfId, best, visits);
  }
  const float bound0 = best.d();   // group-uniform; maxDist^2 without a seed
  const bool tight = bound0 < g.ring2Safe;   // the bound lies inside the 3x3x3 neighbourhood: clip before fetching

  // ---- rings 0 and 1: nine (y,z) rows, each the contiguous x range [cx-1, cx+1] (clipped to the bound's reach when tight)
  uint32_t sC = 0, eC = 0, sL[NR], eL[NR];
  float lbL[NR];
#pragma unroll
  for (int t = 0; t < NR; ++t) { sL[t] = 0; eL[t] = 0; lbL[t] = CUDART_INF_F; }
  {
    int xa = max(ax.c - 1, 0), xb = min(ax.c + 1, g.nx - 1);
    if (tight) {   // no point with d2 <= bound0 lies outside the cells of q.x -+ (sqrt(bound0) + slack): same argument as the shell walk below
      const float reach = __fadd_rn(sqrtf(bound0), g.slack);
      xa = max(xa, cellOfClamped(__fsub_rn(__fsub_rn(q.x, reach), g.slack), g.ox, g.inv_cell, g.nx));
      xb = min(xb, cellOfClamped(__fadd_rn(__fadd_rn(q.x, reach), g.slack), g.ox, g.inv_cell, g.nx));
    }
    if (finite && xa <= xb) {
      if ((unsigned)ay.c < (unsigned)g.ny && (unsigned)az.c < (unsigned)g.nz) {
        const uint32_t base = (uint32_t)g.nx * ((uint32_t)ay.c + (uint32_t)g.ny * (uint32_t)az.c);
        sC = cellStart[base + xa];
        eC = cellStart[base + xb + 1];
      }
#pragma unroll
      for (int t = 0; t < NR; ++t) {
        const int rr = l + t * G;
        const int ri = rr + (rr >= 4 ? 1 : 0);
        const int iy = ay.c + (ri % 3) - 1, iz = az.c + (ri / 3) - 1;
        if ((unsigned)iy < (unsigned)g.ny && (unsigned)iz < (unsigned)g.nz) {
          const float lb = lb3(0.f, axisGap(ay, iy, g.cell, g.slack), axisGap(az, iz, g.cell, g.slack));
          if (!(lb > bound0)) {
            const uint32_t base = (uint32_t)g.nx * ((uint32_t)iy + (uint32_t)g.ny * (uint32_t)iz);
            sL[t] = cellStart[base + xa];
            eL[t] = cellStart[base + xb + 1];
            lbL[t] = lb;
          }
        }
      }
    }
  }
  scanCoop<G, SELFCHK>(pts, sC, eC, l, q, selfId, best, visits);
  {
    // pruning needs only the group's best DISTANCE so far (3 shuffles); the lane-local (key, pos) pairs are merged once, below
    float bound = best.d();
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) bound = fminf(bound, __shfl_xor_sync(FULL, bound, o));
    unsigned m = 0;
#pragma unroll
    for (int t = 0; t < NR; ++t) {
      const bool alive = (eL[t] > sL[t]) && !(lbL[t] > bound);
      m |= ((__ballot_sync(FULL, alive) >> gbase) & ((1u << G) - 1u)) << (t * G);
    }
    while (__any_sync(FULL, m != 0u)) {
      const int r = m ? (__ffs(m) - 1) : 0;
      const int src = gbase + (r & (G - 1));
      uint32_t s = __shfl_sync(FULL, sL[0], src), e = __shfl_sync(FULL, eL[0], src);
      float lb = __shfl_sync(FULL, lbL[0], src);
      if (NR == 2) {
        const uint32_t s1 = __shfl_sync(FULL, sL[NR - 1], src), e1 = __shfl_sync(FULL, eL[NR - 1], src);
        const float lb1 = __shfl_sync(FULL, lbL[NR - 1], src);
        if (r >= G) { s = s1; e = e1; lb = lb1; }
      }
      if (!m || lb > fminf(bound, best.d())) e = s;   // lane-local re-prune (a lane's own best only improves)
      m &= m - 1u;
      scanCoop<G, SELFCHK>(pts, s, e, l, q, selfId, best, visits);
    }
    groupMin<G>(best);
  }

  // ---- beyond the 3x3x3 neighbourhood: rings in the (y,z) plane only.  Ring r = the rows with max(|dy|, |dz|) == r; along x a row is
  // ONE contiguous range of the directory, [xlo, xhi] = the cells a point with d2 <= bound can lie in (two loads, whatever its length),
  // so a shell costs 8 r row look-ups instead of the 8 r rows + 2 (2r - 1)^2 end cells of a cubic shell.  Rings 0 and 1 are the nine
  // rows of the first stage again, now beyond x = cx -+ 1 (two ranges per row).  A point found anywhere tightens bound, which prunes
  // later rows (rowLb > bound) and shortens their x range; the walk ends at the first ring whose nearest face is farther than the best.
  {
    const bool need0 = finite && !(best.d() < g.ring2Safe);   // any point outside the 3x3x3 cells is farther than sqrt(ring2Safe)
    int r = 0, rEnd = -1;
    if (need0) {
      r = max(max(max(-ay.c, ay.c - (g.ny - 1)), max(-az.c, az.c - (g.nz - 1))), 0);   // first ring with a row inside the grid
      rEnd = max(max(ay.c, g.ny - 1 - ay.c), max(az.c, g.nz - 1 - az.c));
    }
    auto ringLb2 = [&](int rr) {
      const float f = fminf(fminf(axisGap(ay, ay.c - rr, g.cell, g.slack), axisGap(ay, ay.c + rr, g.cell, g.slack)),
                            fminf(axisGap(az, az.c - rr, g.cell, g.slack), axisGap(az, az.c + rr, g.cell, g.slack)));
      return lb3(0.f, 0.f, f);
    };
    bool need = need0 && r <= rEnd && !(ringLb2(r) > best.d());
    if (need && l == 0) ++farQ;   // statistic: queries the 3x3x3 neighbourhood could not settle (drives the lanes-per-query choice)
    const unsigned gmask = ((1u << G) - 1u) << gbase;
    float bound = best.d();   // group-wide pruning bound (starts at maxDist^2); lane-local bests are merged after each ring
    while (__any_sync(FULL, need)) {
      if (need) {
        // the ring's rows inside the grid: sides z = cz -+ r over y in [ya, yb], sides y = cy -+ r over z in [za, zb]
        const int ya = max(ay.c - r, 0), yb = min(ay.c + r, g.ny - 1), za = max(az.c - r + 1, 0), zb = min(az.c + r - 1, g.nz - 1);
        const int wy = max(yb - ya + 1, 0), wz = max(zb - za + 1, 0);
        const int nA = ((unsigned)(az.c - r) < (unsigned)g.nz) ? wy : 0;
        const int nB = (r > 0 && (unsigned)(az.c + r) < (unsigned)g.nz) ? wy : 0;
        const int nC = (r > 0 && (unsigned)(ay.c - r) < (unsigned)g.ny) ? wz : 0;
        const int nD = (r > 0 && (unsigned)(ay.c + r) < (unsigned)g.ny) ? wz : 0;
        const int nrows = nA + nB + nC + nD;
        for (int t0 = 0; t0 < nrows; t0 += G) {   // one round = G rows, one per lane; the group stays converged
          int t = t0 + l;
          if (t < nrows) {
            int iy, iz;
            if (t < nA) { iy = ya + t; iz = az.c - r; }
            else if ((t -= nA) < nB) { iy = ya + t; iz = az.c + r; }
            else if ((t -= nB) < nC) { iy = ay.c - r; iz = za + t; }
            else { iy = ay.c + r; iz = za + (t - nC); }
            const float rowLb = lb3(0.f, axisGap(ay, iy, g.cell, g.slack), axisGap(az, iz, g.cell, g.slack));
            if (!(rowLb > bound)) {
              // x cells a better point of this row can lie in: the index build's own (monotone) cell function applied to
              // q.x -+ reach, reach = sqrt(bound - rowLb) + slack  =>  no point with d2 <= bound falls outside [xlo, xhi]
              int xlo = 0, xhi = g.nx - 1;
              if (bound < CUDART_INF_F) {
                const float reach = __fadd_rn(sqrtf(fmaxf(__fsub_rn(bound, rowLb), 0.f)), g.slack);
                xlo = cellOfClamped(__fsub_rn(__fsub_rn(q.x, reach), g.slack), g.ox, g.inv_cell, g.nx);
                xhi = cellOfClamped(__fadd_rn(__fadd_rn(q.x, reach), g.slack), g.ox, g.inv_cell, g.nx);
              }
              const uint32_t rowBase = (uint32_t)g.nx * ((uint32_t)iy + (uint32_t)g.ny * (uint32_t)iz);
#pragma unroll 1
              for (int seg = 0; seg < 2; ++seg) {
                int xa = xlo, xb = xhi;
                if (r <= 1) {   // cells cx - 1 .. cx + 1 of these rows were the first stage's
                  if (seg == 0) xb = min(xb, ax.c - 2); else xa = max(xa, ax.c + 2);
                } else if (seg == 1) break;
                if (xa > xb) continue;
                const uint32_t s = cellStart[rowBase + xa], e = cellStart[rowBase + xb + 1];
                for (uint32_t j = s; j < e; j += 4) {   // four independent loads in flight per lane (this loop is latency bound)
                  const uint32_t j1 = min(j + 1, e - 1), j2 = min(j + 2, e - 1), j3 = min(j + 3, e - 1);
                  const float4 p0 = ldPoint(pts + j), p1 = ldPoint(pts + j1), p2 = ldPoint(pts + j2), p3 = ldPoint(pts + j3);
                  test1<SELFCHK>(p0, j, q, selfId, best, visits);
                  if (j + 1 < e) test1<SELFCHK>(p1, j1, q, selfId, best, visits);
                  if (j + 2 < e) test1<SELFCHK>(p2, j2, q, selfId, best, visits);
                  if (j + 3 < e) test1<SELFCHK>(p3, j3, q, selfId, best, visits);
                }
              }
              bound = fminf(bound, best.d());
            }
          }
          // the first hit of any lane tightens everybody's row and x-reach pruning from the next round on
#pragma unroll
          for (int o = G / 2; o > 0; o >>= 1) bound = fminf(bound, __shfl_xor_sync(gmask, bound, o));
        }
      }
      groupMin<G>(best);
      if (need) {
        ++r;
        need = r <= rEnd && !(ringLb2(r) > best.d());
      }
    }
  }
  if (live && l == 0) {
    const bool found = best.key != initKey;
    outIds[qi] = found ? (int)(unsigned)best.key : -1;
    outD2[qi] = found ? best.d() : CUDART_INF_F;
    outPos[qi] = found ? best.pos : -1;
    if (sHist && found && best.d() != CUDART_INF_F) {   // pass 0 of the trimmed-distance radix select, fused (icp_kernels.cuh)
      const unsigned bin = (unsigned)(best.key >> 53);   // top 11 bits of the d2 pattern
      atomicAdd(&sHist[bin], 1u);
      atomicMin(&sHist[2048], bin);
      atomicMax(&sHist[2049], bin);
    }
  }
}

// touched-point statistic (Matcher::visitCounter): warp shuffle -> shared; the last warp out adds the block total (no barrier
// at the end of the kernel: warps retire as they finish)
__device__ __forceinline__ void flushVisits(unsigned visits, unsigned farQ, unsigned* sVisits, unsigned* sFarQ, unsigned* sTicket,
                                            unsigned long long* visitCounter, const unsigned* sHist = nullptr, unsigned* gHist = nullptr) {
  __syncwarp();   // the group leaders' shared-memory atomics are ordered before lane 0's ticket
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    visits += __shfl_xor_sync(0xffffffffu, visits, o);
    farQ += __shfl_xor_sync(0xffffffffu, farQ, o);
  }
  unsigned lastOut = 0;
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(sVisits, visits);
    if (farQ) atomicAdd(sFarQ, farQ);
    __threadfence_block();
    lastOut = (atomicAdd(sTicket, 1u) == (blockDim.x >> 5) - 1u) ? 1u : 0u;
    if (lastOut) {
      const unsigned total = atomicAdd(sVisits, 0u), far = atomicAdd(sFarQ, 0u);
      if (total) atomicAdd(visitCounter, (unsigned long long)total);
      if (far) atomicAdd(visitCounter + 1, (unsigned long long)far);   // counters[1]: far-query count
    }
  }
  if (sHist) {   // the last warp out adds the block's touched bins to the global pass-0 histogram
    lastOut = __shfl_sync(0xffffffffu, lastOut, 0);
    if (lastOut) {
      __threadfence_block();
      const unsigned lo = ((volatile const unsigned*)sHist)[2048], hi = ((volatile const unsigned*)sHist)[2049];
      if (lo <= hi)
        for (unsigned b = lo + (threadIdx.x & 31); b <= hi; b += 32) {
          const unsigned c = ((volatile const unsigned*)sHist)[b];
          if (c) atomicAdd(&gHist[b], c);
        }
    }
  }
}

// Stand-alone match (b200reg_knn, the host-driven loop): like knn1LoopKernel a resident grid (KNN1_MINB blocks per SM) whose warps
// take 32/G queries per grab from `work` (zeroed by the launcher) — query cost varies by an order of magnitude and far queries come
// in runs, so no warp idles behind one.
template <int G, bool SELFCHK>
__global__ void __launch_bounds__(256, KNN1_LOOP_MINB) knn1GroupKernel(const float4* __restrict__ pts, const uint32_t* __restrict__ cellStart, const GridDesc* __restrict__ gp,
                                                                       const float4* __restrict__ queries, int64_t n, int hasT, Rigid12 T, float maxd2,
                                                                       int* __restrict__ outIds, float* __restrict__ outD2, int* outPos,
                                                                       unsigned long long* __restrict__ visitCounter, const int* seedPos,
                                                                       unsigned long long* __restrict__ work) {
  constexpr int LOG2G = (G == 8) ? 3 : 2;
  constexpr int QPW = 32 >> LOG2G;   // queries of one warp-wide call
  __shared__ unsigned sVisits, sFarQ, sTicket;
  const GridDesc g = *gp;   // device-derived grid (index.cu)
  if (threadIdx.x == 0) { sVisits = 0; sFarQ = 0; sTicket = 0; }
  __syncthreads();   // all warps are at the start here: nobody waits on a straggler
  unsigned visits = 0, farQ = 0;
  const int lane = threadIdx.x & 31;
  for (;;) {
    long long base = 0;
    if (lane == 0) base = (long long)atomicAdd(work, (unsigned long long)QPW);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= n) break;
    knn1Query<G, SELFCHK>(pts, cellStart, g, queries, n, base + (lane >> LOG2G), hasT, T, maxd2, outIds, outD2, outPos, visits, farQ, nullptr, seedPos);
  }
  flushVisits(visits, farQ, &sVisits, &sFarQ, &sTicket, visitCounter);
}

// The same search inside the device-resident registration loop (icp.cu): arguments and the iteration transform are read from
// the loop's blocks in HBM, the grid is fixed (a multiple of the SM count) and strides over the queries, so one
// instantiated CUDA graph serves every reading size and reference.
template <int G>
__device__ __forceinline__ void knn1LoopBody(const IcpDevParams* __restrict__ P, const IcpDevState* __restrict__ S, unsigned& visits,
                                             unsigned& farQ, unsigned* __restrict__ sHist, unsigned long long* __restrict__ work) {
  constexpr int LOG2G = (G == 8) ? 3 : 2;
  constexpr int QPW = 32 >> LOG2G;   // queries of one warp-wide call
  constexpr int CHUNK = QPW;         // queries a warp takes from the work counter at a time
  const GridDesc g = *P->gp;
  const float4* pts = P->pts;
  const uint32_t* cellStart = P->cellStart;
  const float4* queries = P->reading;
  const int64_t n = P->n;
  const float maxd2 = P->maxd2;
  int* ids = P->ids;
  float* d2 = P->d2;
  int* pos = P->pos;
  // the previous iteration's matches (same reference) seed the search once far queries are rare (the 4-lane regime): while many
  // queries still move by more than a cell per iteration the seed's extra dependent load costs more than its bound saves
  const int* seed = (P->seedOn && S->iterations > 0 && G == 4) ? P->pos : nullptr;
  Rigid12 T;
#pragma unroll
  for (int j = 0; j < 12; ++j) T.m[j] = S->T[j];
  // Query cost varies by an order of magnitude (a query far from the map walks rings, and such queries come in runs: the scan is in
  // sensor order), so the warps of the resident grid take chunks of 32 consecutive queries from a counter instead of a fixed stride:
  // no warp idles while another still has a run of far queries ahead of it.  The loop's tail resets the counter.
  const int lane = threadIdx.x & 31;
  for (;;) {
    long long base = 0;
    if (lane == 0) base = (long long)atomicAdd(work, (unsigned long long)CHUNK);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= n) break;
#pragma unroll 1
    for (int sub = 0; sub < CHUNK; sub += QPW)
      knn1Query<G, false>(pts, cellStart, g, queries, n, base + sub + (lane >> LOG2G), 1, T, maxd2, ids, d2, pos, visits, farQ, sHist, seed);
  }
}

// S->lanes (written by the loop's tail from the previous iteration's far-query count) picks the decomposition: 8 lanes per
// query while many queries need rings >= 2 (the first iterations of a badly aligned scan), 4 once they are rare.
__global__ void __launch_bounds__(256, KNN1_LOOP_MINB) knn1LoopKernel(const IcpDevParams* __restrict__ P, const IcpDevState* __restrict__ S,
                                                         unsigned long long* __restrict__ visitCounter) {
  if ((int)blockIdx.x >= P->matchBlocks) return;   // the registration asked for a part of the resident grid (icp.cu icpEnqueue)
  __shared__ unsigned sVisits, sFarQ, sTicket;
  __shared__ unsigned sHistBuf[2050];   // pass-0 histogram of the block's d2 + touched bin range
  unsigned* gHist = P->fhist0;
  unsigned* sHist = gHist ? sHistBuf : nullptr;
  if (threadIdx.x == 0) { sVisits = 0; sFarQ = 0; sTicket = 0; }
  if (sHist) {
    for (int b = threadIdx.x; b < 2048; b += blockDim.x) sHistBuf[b] = 0;
    if (threadIdx.x == 0) { sHistBuf[2048] = 0xffffffffu; sHistBuf[2049] = 0; }
  }
  __syncthreads();
  traceStamp(P->trace, 1);
  unsigned visits = 0, farQ = 0;
  if (S->lanes == 4) knn1LoopBody<4>(P, S, visits, farQ, sHist, visitCounter + 2);
  else knn1LoopBody<8>(P, S, visits, farQ, sHist, visitCounter + 2);
  flushVisits(visits, farQ, &sVisits, &sFarQ, &sTicket, visitCounter, sHist, gHist);
}

// ---- 2 <= k <= 32 (SurfaceNormal's self-kNN, k-NN matchers): W lanes per query, W = 8 / 16 / 32 >= k ------------------
// The running result is a sorted list DISTRIBUTED over the group: lane j holds the j-th best packed (d2, index) key and
// its sorted position.  A contiguous point range is tested W candidates at a time (one coalesced float4 load per lane);
// candidates that beat the current k-th key are found with one ballot and inserted one by one with a shuffle-up
// (O(1) per insertion for the whole group, ~k ln(candidates / k) insertions per query).  Cells are visited like the k = 1
// kernel: the 3x3x3 neighbourhood as 9 rows whose directory entries are fetched in one round, then whole shells, every
// row and shell pruned by its exact fp32 lower bound against the k-th distance (maxDist^2 while the list is not full).
// queries == nullptr: the index's own points are the queries, taken in SORTED order (neighbouring groups share cells;
// results are written at the original index) — the self-kNN of SurfaceNormal.cpp:170-188.
// TWO (W = 32 only, 32 < k <= 64): every lane holds TWO entries — entry l of the lower half (the 32 best) and entry 32 + l of the
// upper half; an insertion into the lower half pushes its last entry into the head of the upper half.
template <int W, bool TWO = false>
struct KnnList {
  unsigned long long key, kth;   // this lane's entry; the group's k-th entry
  int pos;
  unsigned gmask;
  int l, kLane;                  // kLane: lane holding the k-th entry (in the upper half when TWO)
  unsigned long long keyB;       // upper half (TWO)
  int posB;
  __device__ __forceinline__ float kthD() const { return __uint_as_float((unsigned)(kth >> 32)); }
  __device__ __forceinline__ static void insertInto(unsigned long long& key, int& pos, unsigned long long ck, int cp, unsigned gmask, int l) {
    const unsigned long long upK = __shfl_up_sync(gmask, key, 1, W);
    const int upP = __shfl_up_sync(gmask, pos, 1, W);
    if (key > ck) {
      const bool take = (l == 0) || (upK < ck);
      key = take ? ck : upK;
      pos = take ? cp : upP;
    }
  }
  __device__ __forceinline__ void insert(unsigned long long ck, int cp) {
    if (!TWO) {
      insertInto(key, pos, ck, cp, gmask, l);
      kth = __shfl_sync(gmask, key, kLane, W);
    } else {
      const unsigned long long lastK = __shfl_sync(gmask, key, W - 1, W);
      const int lastP = __shfl_sync(gmask, pos, W - 1, W);
      if (ck < lastK) {   // uniform over the group: into the lower half, whose last entry moves to the head of the upper half
        insertInto(key, pos, ck, cp, gmask, l);
        const unsigned long long upK = __shfl_up_sync(gmask, keyB, 1, W);
        const int upP = __shfl_up_sync(gmask, posB, 1, W);
        keyB = (l == 0) ? lastK : upK;
        posB = (l == 0) ? lastP : upP;
      } else insertInto(keyB, posB, ck, cp, gmask, l);
      kth = __shfl_sync(gmask, keyB, kLane, W);
    }
  }
};

// compare-exchange of a bitonic network over the W lanes of a group: lane l and lane l ^ j; the lower lane keeps the minimum when `up`
template <int W>
__device__ __forceinline__ void bitonicStep(unsigned long long& k, int& p, int j, bool up, unsigned gmask, int l) {
  const unsigned long long ok = __shfl_xor_sync(gmask, k, j, W);
  const int op = __shfl_xor_sync(gmask, p, j, W);
  const bool takeMin = ((l & j) == 0) == up;
  if (takeMin ? (ok < k) : (ok > k)) { k = ok; p = op; }
}
// A chunk with many accepted candidates is not inserted one by one (a shuffle-up of the whole list per candidate): the chunk is sorted by
// a bitonic network, reversed against the list — the W smallest of the union are then min(list[l], chunk[W - 1 - l]), a bitonic sequence
// — and one bitonic merge restores the order.  ~(log2 W (log2 W + 1) / 2 + log2 W + 1) exchange steps whatever the number of candidates;
// the same list results (the W best keys are the W best keys however they got there).
#define KNN_BULK_MIN 12
template <int W>
__device__ __forceinline__ void bulkMerge(unsigned long long& key, int& pos, unsigned long long ck, int cp, unsigned gmask, int l) {
#pragma unroll
  for (int size = 2; size <= W; size <<= 1)
#pragma unroll
    for (int j = size >> 1; j > 0; j >>= 1) bitonicStep<W>(ck, cp, j, (l & size) == 0 || size == W, gmask, l);
  const unsigned long long rk = __shfl_sync(gmask, ck, W - 1 - l, W);
  const int rp = __shfl_sync(gmask, cp, W - 1 - l, W);
  if (rk < key) { key = rk; pos = rp; }
#pragma unroll
  for (int j = W >> 1; j > 0; j >>= 1) bitonicStep<W>(key, pos, j, true, gmask, l);
}

template <int W, bool SELFCHK, bool TWO>
__device__ __forceinline__ void scanList(const float4* __restrict__ pts, uint32_t s, uint32_t e, const float3 q, int selfId, KnnList<W, TWO>& L,
                                         unsigned& visits) {
  if (s >= e) return;
  const int gbase = (threadIdx.x & 31) & (32 - W);
  float4 p = ldPoint(pts + min(s + (uint32_t)L.l, e - 1));
  for (uint32_t j0 = s; j0 < e; j0 += W) {
    const uint32_t j = j0 + L.l;
    const float4 pn = ldPoint(pts + min(j + W, e - 1));   // next chunk in flight while this one is inserted
    const int id = __float_as_int(p.w);
    const bool ok = (j < e) && !(SELFCHK && id == selfId);
    const unsigned long long key = packKey(dist2_rn(q.x, q.y, q.z, p.x, p.y, p.z), id);
    visits += ok ? 1u : 0u;
    const bool accept = ok && key < L.kth;
    unsigned m = __ballot_sync(L.gmask, accept);
    if (W < 32) m = (m >> gbase) & ((1u << (W & 31)) - 1u);
    if (!TWO && W == 32 && __popc(m) >= KNN_BULK_MIN) {   // uniform over the group
      bulkMerge<W>(L.key, L.pos, accept ? key : ~0ull, (int)j, L.gmask, L.l);
      L.kth = __shfl_sync(L.gmask, L.key, L.kLane, W);
      m = 0;
    }
    while (m) {
      const int b = __ffs(m) - 1;
      m &= m - 1u;
      const unsigned long long ck = __shfl_sync(L.gmask, key, b, W);
      const int cp = __shfl_sync(L.gmask, (int)j, b, W);
      if (ck < L.kth) L.insert(ck, cp);
    }
    p = pn;
  }
}

template <int W, bool SELFCHK, bool TWO>
__global__ void __launch_bounds__(256, 4) knnListKernel(const float4* __restrict__ pts, const uint32_t* __restrict__ cellStart, const GridDesc* __restrict__ gp,
                                                     const float4* __restrict__ queries, int64_t n, int hasT, Rigid12 T, int k, float maxd2,
                                                     int* __restrict__ outIds, float* __restrict__ outD2, int* __restrict__ outPos,
                                                     unsigned long long* __restrict__ visitCounter) {
  constexpr int NT = (9 + W - 1) / W;   // near rows per lane
  const int lane = threadIdx.x & 31, l = lane & (W - 1), gbase = lane & (32 - W);
  const unsigned gmask = (W == 32) ? 0xffffffffu : (((1u << (W & 31)) - 1u) << gbase);
  const int64_t slot = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / W;
  const GridDesc g = *gp;   // device-derived grid (index.cu)
  unsigned visits = 0;
  if (slot < n) {
    float4 q4;
    int64_t qi = slot;
    if (queries) q4 = queries[slot];
    else { q4 = pts[slot]; qi = (int64_t)__float_as_int(q4.w); q4.w = 1.f; }
    float3 q = make_float3(q4.x, q4.y, q4.z);
    if (hasT) q = xformPoint(T, q4.x, q4.y, q4.z, q4.w);
    const bool finite = (fabsf(q.x) <= FLT_MAX) && (fabsf(q.y) <= FLT_MAX) && (fabsf(q.z) <= FLT_MAX);
    const int selfId = (int)qi;
    const float lim = 1.0e9f;
    const AxisGeom ax{q.x, g.ox, finite ? (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.x, g.ox), g.inv_cell), -lim), lim)) : 0};
    const AxisGeom ay{q.y, g.oy, finite ? (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.y, g.oy), g.inv_cell), -lim), lim)) : 0};
    const AxisGeom az{q.z, g.oz, finite ? (int)floorf(fminf(fmaxf(__fmul_rn(__fsub_rn(q.z, g.oz), g.inv_cell), -lim), lim)) : 0};
    const unsigned long long initKey = packKey(maxd2, 0);   // an entry must beat (maxDist^2, 0): d2 < maxDist^2 strictly
    KnnList<W, TWO> L{initKey, initKey, -1, gmask, l, TWO ? k - 1 - W : k - 1, initKey, -1};

    // ---- rings 0 and 1: nine (y,z) rows [cx-1, cx+1], directory entries fetched in one round
    uint32_t sR[NT], eR[NT];
    float lbR[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      sR[t] = 0; eR[t] = 0; lbR[t] = CUDART_INF_F;
      const int ri = l + t * W;
      const int xa = max(ax.c - 1, 0), xb = min(ax.c + 1, g.nx - 1);
      const int iy = ay.c + (ri % 3) - 1, iz = az.c + (ri / 3) - 1;
      if (finite && ri < 9 && xa <= xb && (unsigned)iy < (unsigned)g.ny && (unsigned)iz < (unsigned)g.nz) {
        const uint32_t base = (uint32_t)g.nx * ((uint32_t)iy + (uint32_t)g.ny * (uint32_t)iz);
        sR[t] = cellStart[base + xa];
        eR[t] = cellStart[base + xb + 1];
        lbR[t] = lb3(0.f, axisGap(ay, iy, g.cell, g.slack), axisGap(az, iz, g.cell, g.slack));
      }
    }
    scanList<W, SELFCHK, TWO>(pts, __shfl_sync(gmask, sR[0], 4, W), __shfl_sync(gmask, eR[0], 4, W), q, selfId, L, visits);   // centre row first
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      unsigned m = __ballot_sync(gmask, eR[t] > sR[t] && (l + t * W) != 4 && !(lbR[t] > L.kthD()));
      if (W < 32) m = (m >> gbase) & ((1u << (W & 31)) - 1u);
      while (m) {
        const int b = __ffs(m) - 1;
        m &= m - 1u;
        const uint32_t s = __shfl_sync(gmask, sR[t], b, W), e = __shfl_sync(gmask, eR[t], b, W);
        const float lb = __shfl_sync(gmask, lbR[t], b, W);
        if (!(lb > L.kthD())) scanList<W, SELFCHK, TWO>(pts, s, e, q, selfId, L, visits);
      }
    }

    // ---- rings >= 2: one round = W rows of the shell, one per lane (bounds, x reach, directory), then scanned by the group
    bool need = finite && !(L.kthD() < g.ring2Safe);
    int r = 2, rEnd = 0;
    if (need) {
      const int r0 = max(max(max(-ax.c, ax.c - (g.nx - 1)), max(-ay.c, ay.c - (g.ny - 1))), max(max(-az.c, az.c - (g.nz - 1)), 0));
      rEnd = max(max(max(ax.c, g.nx - 1 - ax.c), max(ay.c, g.ny - 1 - ay.c)), max(az.c, g.nz - 1 - az.c));
      r = max(2, r0);
      need = r <= rEnd && !(ringLowerBound(ax, ay, az, r, g.cell, g.slack) > L.kthD());
    }
    while (need) {
      const int y0 = max(ay.c - r, 0), z0 = max(az.c - r, 0);   // shell rows clipped to the grid
      const int wy = max(min(ay.c + r, g.ny - 1) - y0 + 1, 0), wz = max(min(az.c + r, g.nz - 1) - z0 + 1, 0);
      const int nrows = wy * wz;
      for (int t0 = 0; t0 < nrows; t0 += W) {
        const int t = t0 + l;
        const int iy = y0 + t % max(wy, 1), iz = z0 + t / max(wy, 1);
        const int dy = iy - ay.c, dz = iz - az.c;
        uint32_t s0 = 0, e0 = 0, s1 = 0, e1 = 0;
        float rowLb = CUDART_INF_F;
        if (t < nrows) {
          rowLb = lb3(0.f, axisGap(ay, iy, g.cell, g.slack), axisGap(az, iz, g.cell, g.slack));
          const float bound = L.kthD();
          if (!(rowLb > bound)) {
            int xlo = 0, xhi = g.nx - 1;
            if (bound < CUDART_INF_F) {
              const float reach = __fadd_rn(sqrtf(fmaxf(__fsub_rn(bound, rowLb), 0.f)), g.slack);
              xlo = cellOfClamped(__fsub_rn(__fsub_rn(q.x, reach), g.slack), g.ox, g.inv_cell, g.nx);
              xhi = cellOfClamped(__fadd_rn(__fadd_rn(q.x, reach), g.slack), g.ox, g.inv_cell, g.nx);
            }
            const uint32_t rowBase = (uint32_t)g.nx * ((uint32_t)iy + (uint32_t)g.ny * (uint32_t)iz);
            if ((abs(dz) == r) || (abs(dy) == r)) {   // shell row: the whole x range
              const int xa = max(ax.c - r, xlo), xb = min(ax.c + r, xhi);
              if (xa <= xb) { s0 = cellStart[rowBase + xa]; e0 = cellStart[rowBase + xb + 1]; }
            } else {                                   // interior row: its two end cells
              const int xl = ax.c - r, xr = ax.c + r;
              if (xl >= xlo && xl <= xhi) { s0 = cellStart[rowBase + xl]; e0 = cellStart[rowBase + xl + 1]; }
              if (xr >= xlo && xr <= xhi) { s1 = cellStart[rowBase + xr]; e1 = cellStart[rowBase + xr + 1]; }
            }
          }
        }
        unsigned m = __ballot_sync(gmask, e0 > s0 || e1 > s1);
        if (W < 32) m = (m >> gbase) & ((1u << (W & 31)) - 1u);
        while (m) {
          const int b = __ffs(m) - 1;
          m &= m - 1u;
          const float lb = __shfl_sync(gmask, rowLb, b, W);
          const uint32_t a0 = __shfl_sync(gmask, s0, b, W), b0 = __shfl_sync(gmask, e0, b, W);
          const uint32_t a1 = __shfl_sync(gmask, s1, b, W), b1 = __shfl_sync(gmask, e1, b, W);
          if (!(lb > L.kthD())) {
            scanList<W, SELFCHK, TWO>(pts, a0, b0, q, selfId, L, visits);
            scanList<W, SELFCHK, TWO>(pts, a1, b1, q, selfId, L, visits);
          }
        }
      }
      ++r;
      need = r <= rEnd && !(ringLowerBound(ax, ay, az, r, g.cell, g.slack) > L.kthD());
    }
    if (l < k) {
      const bool found = L.key != initKey;
      const int64_t o = (int64_t)k * qi + l;
      outIds[o] = found ? (int)(unsigned)L.key : -1;
      outD2[o] = found ? __uint_as_float((unsigned)(L.key >> 32)) : CUDART_INF_F;
      outPos[o] = found ? L.pos : -1;
    }
    if (TWO && W + l < k) {
      const bool found = L.keyB != initKey;
      const int64_t o = (int64_t)k * qi + W + l;
      outIds[o] = found ? (int)(unsigned)L.keyB : -1;
      outD2[o] = found ? __uint_as_float((unsigned)(L.keyB >> 32)) : CUDART_INF_F;
      outPos[o] = found ? L.posB : -1;
    }
  }
  for (int o = 16; o > 0; o >>= 1) visits += __shfl_xor_sync(0xffffffffu, visits, o);
  if ((threadIdx.x & 31) == 0 && visits) atomicAdd(visitCounter, (unsigned long long)visits);
}

template <int W, bool TWO = false>
void launchKnnList(b200reg_ctx* ctx, const b200reg_index* ix, const float4* q, int64_t n, int hasT, const Rigid12& T, int k, float maxd2,
                   int allowSelf, int* ids, float* d2, int* pos, unsigned long long* visits) {
  const int blocks = divUp(n * W, 256);
  if (allowSelf)
    knnListKernel<W, false, TWO><<<blocks, 256, 0, ctx->stream>>>(ix->pts, ix->cellStart, &ix->dHdr->g, q, n, hasT, T, k, maxd2, ids, d2, pos, visits);
  else
    knnListKernel<W, true, TWO><<<blocks, 256, 0, ctx->stream>>>(ix->pts, ix->cellStart, &ix->dHdr->g, q, n, hasT, T, k, maxd2, ids, d2, pos, visits);
}

template <int G>
void launchKnn1(b200reg_ctx* ctx, const b200reg_index* ix, const float4* q, int64_t n, int hasT, const Rigid12& T, float maxd2, int allowSelf, int* ids, float* d2, int* pos, unsigned long long* visits, const int* seed) {
  const int blocks = (int)std::min<int64_t>(divUp(n * G, 256), (int64_t)KNN1_LOOP_MINB * ctx->sm_count);   // resident grid
  unsigned long long* work = visits + 2;   // dstate[2], zeroed by runKnn
  if (allowSelf)
    knn1GroupKernel<G, false><<<blocks, 256, 0, ctx->stream>>>(ix->pts, ix->cellStart, &ix->dHdr->g, q, n, hasT, T, maxd2, ids, d2, pos, visits, seed, work);
  else
    knn1GroupKernel<G, true><<<blocks, 256, 0, ctx->stream>>>(ix->pts, ix->cellStart, &ix->dHdr->g, q, n, hasT, T, maxd2, ids, d2, pos, visits, seed, work);
}

template <int K>
void launchKnn(b200reg_ctx* ctx, const b200reg_index* ix, const float4* q, int64_t n, int hasT, const Rigid12& T, int k, float maxd2,
               int allowSelf, int* ids, float* d2, int* pos, unsigned long long* visits) {
  knnKernel<K><<<divUp(n, 128), 128, 0, ctx->stream>>>(ix->pts, ix->cellStart, &ix->dHdr->g, q, n, hasT, T, k, maxd2, allowSelf, ids, d2, pos, visits);
}

}  // namespace

// k = 1 match of the device-resident loop; `blocks` is fixed by the caller (graph capture)
int launchKnn1Loop(b200reg_ctx* ctx, const IcpDevParams* P, const IcpDevState* S, unsigned long long* counters, int blocks) {
  knn1LoopKernel<<<blocks, 256, 0, ctx->stream>>>(P, S, counters);
  return B200REG_OK;
}

// Device-side entry shared with icp.cu / normals.cu.  Results land in ctx->matchIds / matchD2 / matchPos (k*n entries);
// queries == nullptr (2 <= k <= 32 and the list kernel only): the index's own points, n == index size.
// the touched-point counter is dstate[0] (unsigned long long), accumulated.
// seeded != 0 (k = 1 only): ctx->matchPos still holds the previous iteration's matches of the SAME queries against the SAME index;
// they are re-measured and used as the starting bound (knn1Query).
int runKnn(b200reg_ctx* ctx, const b200reg_index* ix, const float4* queries, int64_t n, const float* T, int k, float maxDist, int allowSelf,
           int seeded) {
  if (k < 1 || k > 64) return setErr(ctx, B200REG_E_INVALID_PARAM, "knn: k = %d outside the supported range [1, 64]", k);
  if (!(maxDist >= 0.f)) return setErr(ctx, B200REG_E_INVALID_PARAM, "knn: maxDist must be >= 0");
  TRY(ensureBuf(ctx, ctx->matchIds, sizeof(int) * (size_t)k * n));
  TRY(ensureBuf(ctx, ctx->matchD2, sizeof(float) * (size_t)k * n));
  TRY(ensureBuf(ctx, ctx->matchPos, sizeof(int) * (size_t)k * n));
  if (n == 0) return B200REG_OK;
  if (!queries && (k < 2 || ctx->knnLanes == 1 || n != ix->m)) return setErr(ctx, B200REG_E_INVALID_PARAM, "knn: self mode needs k >= 2 and the list kernel");
  Rigid12 R;
  memset(&R, 0, sizeof(R));
  if (T) memcpy(R.m, T, sizeof(R.m));
  const float maxd2 = maxDist * maxDist;   // fp32 product (inf stays inf)
  int* ids = (int*)ctx->matchIds.p;
  float* d2 = (float*)ctx->matchD2.p;
  int* pos = (int*)ctx->matchPos.p;
  unsigned long long* visits = (unsigned long long*)ctx->dstate;
  const int hasT = T ? 1 : 0;
  if (k == 1 && ctx->knnLanes != 1) CUDA_TRY(ctx, cudaMemsetAsync(visits + 2, 0, sizeof(unsigned long long), ctx->stream));   // work counter of knn1GroupKernel
  profBegin(ctx);
  const int* seed = (seeded && ctx->knnSeed && ctx->knnLanes == 4) ? pos : nullptr;   // as in knn1LoopBody: only in the 4-lane regime
  if (k == 1 && (ctx->knnLanes == 8 || ctx->knnLanes == 0)) launchKnn1<8>(ctx, ix, queries, n, hasT, R, maxd2, allowSelf, ids, d2, pos, visits, seed);
  else if (k == 1 && ctx->knnLanes == 4) launchKnn1<4>(ctx, ix, queries, n, hasT, R, maxd2, allowSelf, ids, d2, pos, visits, seed);
  else if (k == 1) launchKnn<1>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else if (k <= 8 && ctx->knnLanes != 1) launchKnnList<8>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else if (k <= 16 && ctx->knnLanes != 1) launchKnnList<16>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else if (k <= 32 && ctx->knnLanes != 1) launchKnnList<32>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else if (k > 32 && ctx->knnLanes != 1) launchKnnList<32, true>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);   // two entries per lane
  else if (k <= 4) launchKnn<4>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else if (k <= 8) launchKnn<8>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else if (k <= 16) launchKnn<16>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else if (k <= 32) launchKnn<32>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
  else return setErr(ctx, B200REG_E_INVALID_PARAM, "knn: the one-thread-per-query kernel (B200REG_KNN_LANES=1) holds at most 32 neighbours");
  profEnd(ctx);
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
  TRY(runKnn(ctx, index, query->xyz, n, T, k, max_dist, allow_self, 0));
  if (ids_out && n) CUDA_TRY(ctx, cudaMemcpyAsync(ids_out, ctx->matchIds.p, sizeof(int) * (size_t)k * n, cudaMemcpyDeviceToHost, ctx->stream));
  if (d2_out && n) CUDA_TRY(ctx, cudaMemcpyAsync(d2_out, ctx->matchD2.p, sizeof(float) * (size_t)k * n, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pinned, ctx->dstate, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (visit_count) *visit_count = *(const unsigned long long*)ctx->pinned;
  return B200REG_OK;
}
