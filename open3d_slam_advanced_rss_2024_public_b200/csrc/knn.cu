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
  if (k == 1) launchKnn<1>(ctx, ix, queries, n, hasT, R, k, maxd2, allowSelf, ids, d2, pos, visits);
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
