// csrc/index.cu — KDTreeMatcher::init replacement (MatchersImpl.cpp:109-114): a radix-sorted voxel-grid index.
//
// Layout in HBM: the (centred) reference is re-ordered by linear cell id (x fastest) into
//   pts[M]  float4 = (x, y, z, original index as int bits)   — one 16 B coalesced load yields position AND id
//   nrm[M]  float4 = reference normals in the same order
//   cellStart[numCells + 1] uint32                           — CSR directory; x-adjacent cells are contiguous ranges
// Build = stats (mean + bbox, fp64 sums) -> key/count -> 32-bit radix sort of (cell, index) -> exclusive scan -> gather.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cfloat>
#include <cmath>

#include "common.cuh"

namespace {

struct StatsPartial {
  double sx, sy, sz;
  float mnx, mny, mnz, mxx, mxy, mxz;
};

__global__ void statsKernel(const float4* __restrict__ xyz, int64_t n, StatsPartial* __restrict__ partials) {
  double sx = 0, sy = 0, sz = 0;
  float mnx = FLT_MAX, mny = FLT_MAX, mnz = FLT_MAX, mxx = -FLT_MAX, mxy = -FLT_MAX, mxz = -FLT_MAX;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 p = xyz[i];
    sx += p.x; sy += p.y; sz += p.z;
    mnx = fminf(mnx, p.x); mny = fminf(mny, p.y); mnz = fminf(mnz, p.z);
    mxx = fmaxf(mxx, p.x); mxy = fmaxf(mxy, p.y); mxz = fmaxf(mxz, p.z);
  }
  for (int o = 16; o > 0; o >>= 1) {
    sx += __shfl_xor_sync(0xffffffffu, sx, o);
    sy += __shfl_xor_sync(0xffffffffu, sy, o);
    sz += __shfl_xor_sync(0xffffffffu, sz, o);
    mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o));
    mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o));
    mnz = fminf(mnz, __shfl_xor_sync(0xffffffffu, mnz, o));
    mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
    mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
    mxz = fmaxf(mxz, __shfl_xor_sync(0xffffffffu, mxz, o));
  }
  __shared__ StatsPartial sm[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) sm[warp] = {sx, sy, sz, mnx, mny, mnz, mxx, mxy, mxz};
  __syncthreads();
  if (threadIdx.x == 0) {
    StatsPartial a = sm[0];
    for (int w = 1; w < (blockDim.x >> 5); ++w) {   // fixed order: deterministic
      a.sx += sm[w].sx; a.sy += sm[w].sy; a.sz += sm[w].sz;
      a.mnx = fminf(a.mnx, sm[w].mnx); a.mny = fminf(a.mny, sm[w].mny); a.mnz = fminf(a.mnz, sm[w].mnz);
      a.mxx = fmaxf(a.mxx, sm[w].mxx); a.mxy = fmaxf(a.mxy, sm[w].mxy); a.mxz = fmaxf(a.mxz, sm[w].mxz);
    }
    partials[blockIdx.x] = a;
  }
}

__device__ __forceinline__ int cellCoord(float p, float o, float inv, int nmax) {
  const int c = (int)floorf(__fmul_rn(__fsub_rn(p, o), inv));
  return min(max(c, 0), nmax - 1);
}

// centred point = fl(p - mean) exactly as ICP.cpp:320 does; key = linear cell (x fastest)
__global__ void keyKernel(const float4* __restrict__ xyz, int64_t n, float mx, float my, float mz, GridDesc g,
                          uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* __restrict__ counts) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = xyz[i];
  const float x = __fsub_rn(p.x, mx), y = __fsub_rn(p.y, my), z = __fsub_rn(p.z, mz);
  const int cx = cellCoord(x, g.ox, g.inv_cell, g.nx), cy = cellCoord(y, g.oy, g.inv_cell, g.ny), cz = cellCoord(z, g.oz, g.inv_cell, g.nz);
  const uint32_t key = (uint32_t)cx + (uint32_t)g.nx * ((uint32_t)cy + (uint32_t)g.ny * (uint32_t)cz);
  keys[i] = key;
  vals[i] = (uint32_t)i;
  atomicAdd(&counts[key], 1u);
}

__global__ void gatherKernel(const float4* __restrict__ xyz, const float4* __restrict__ nrm, const uint32_t* __restrict__ vals,
                             int64_t n, float mx, float my, float mz, float4* __restrict__ pts, float4* __restrict__ outN) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t i = vals[j];
  const float4 p = xyz[i];
  pts[j] = make_float4(__fsub_rn(p.x, mx), __fsub_rn(p.y, my), __fsub_rn(p.z, mz), __int_as_float((int)i));
  if (nrm) outN[j] = nrm[i];
}

}  // namespace

// shared with icp.cu: mean (fp64 sum rounded once to fp32) and bbox of a cloud
int cloudStats(b200reg_ctx* ctx, const b200reg_cloud* c, double sum[3], float mn[3], float mx[3]) {
  const int blocks = (int)fmin((double)(4 * ctx->sm_count), (double)divUp(c->n, 256));
  TRY(ensureBuf(ctx, ctx->partials, sizeof(StatsPartial) * (size_t)blocks));
  TRY(ensurePinned(ctx, sizeof(StatsPartial) * (size_t)blocks));
  statsKernel<<<blocks, 256, 0, ctx->stream>>>(c->xyz, c->n, (StatsPartial*)ctx->partials.p);
  LAUNCH_CHECK(ctx);
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pinned, ctx->partials.p, sizeof(StatsPartial) * (size_t)blocks, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  const StatsPartial* h = (const StatsPartial*)ctx->pinned;
  sum[0] = sum[1] = sum[2] = 0;
  mn[0] = mn[1] = mn[2] = FLT_MAX;
  mx[0] = mx[1] = mx[2] = -FLT_MAX;
  for (int b = 0; b < blocks; ++b) {
    sum[0] += h[b].sx; sum[1] += h[b].sy; sum[2] += h[b].sz;
    mn[0] = fminf(mn[0], h[b].mnx); mn[1] = fminf(mn[1], h[b].mny); mn[2] = fminf(mn[2], h[b].mnz);
    mx[0] = fmaxf(mx[0], h[b].mxx); mx[1] = fmaxf(mx[1], h[b].mxy); mx[2] = fmaxf(mx[2], h[b].mxz);
  }
  return B200REG_OK;
}

// Build the index over `ref` after subtracting `mean` (pass zeros for an un-centred index).
int buildIndex(b200reg_ctx* ctx, const b200reg_cloud* ref, const float mean[3], const float bbMin[3], const float bbMax[3], float cell,
               b200reg_index** out) {
  const int64_t m = ref->n;
  if (m <= 0) return setErr(ctx, B200REG_E_NO_POINTS, "index_build: reference cloud is empty");
  if (m >= (int64_t)0x7fffffff) return setErr(ctx, B200REG_E_INVALID_PARAM, "index_build: more than 2^31-1 points");
  b200reg_index* ix = new b200reg_index();
  ix->m = m;
  for (int d = 0; d < 3; ++d) ix->mean[d] = mean[d];
  // bbox of the centred cloud: fl(min - mean) .. fl(max - mean) (fp32 subtraction is monotone)
  float lo[3], hi[3], ext[3];
  float maxAbs = 0.f;
  for (int d = 0; d < 3; ++d) {
    lo[d] = bbMin[d] - mean[d];
    hi[d] = bbMax[d] - mean[d];
    ext[d] = fmaxf(hi[d] - lo[d], 1e-6f);
    maxAbs = fmaxf(maxAbs, fmaxf(fabsf(lo[d]), fabsf(hi[d])));
  }
  if (!(cell > 0.f)) {   // automatic: ~2.5 x the mean spacing of points spread over the bbox surface
    const double area = 2.0 * ((double)ext[0] * ext[1] + (double)ext[1] * ext[2] + (double)ext[0] * ext[2]);
    cell = (float)(2.5 * sqrt(area / (double)m));
    const float emax = fmaxf(ext[0], fmaxf(ext[1], ext[2]));
    cell = fmaxf(cell, emax * 1e-4f);
  }
  for (;;) {   // keep the directory below 2^30 cells
    const double nc = (floor(ext[0] / cell) + 1) * (floor(ext[1] / cell) + 1) * (floor(ext[2] / cell) + 1);
    if (nc <= (double)(1u << 30)) break;
    cell *= 1.26f;
  }
  GridDesc& g = ix->g;
  g.ox = lo[0]; g.oy = lo[1]; g.oz = lo[2];
  g.cell = cell;
  g.inv_cell = 1.0f / cell;
  g.nx = (int)floorf(ext[0] / cell) + 1;
  g.ny = (int)floorf(ext[1] / cell) + 1;
  g.nz = (int)floorf(ext[2] / cell) + 1;
  // slack >= error of fl((p-o)*inv) cell assignment + error of evaluating cell bounds in fp32 (see knn.cu)
  g.slack = 16.f * 1.1920929e-7f * (maxAbs + fmaxf(ext[0], fmaxf(ext[1], ext[2])) + cell);
  {   // (cell - 4 slack)^2 with a 1 % margin, rounded toward zero: a point >= 2 cells away cannot have a smaller fp32 d2
    const double c = (double)cell - 4.0 * (double)g.slack;
    g.ring2Safe = (c > 0.0) ? nextafterf((float)(0.99 * c * c), 0.f) : 0.f;
  }
  const size_t numCells = (size_t)g.nx * g.ny * g.nz;

  cudaError_t e = devAlloc(ctx, &ix->pts, sizeof(float4) * (size_t)m);
  if (e == cudaSuccess && ref->nrm) e = devAlloc(ctx, &ix->nrm, sizeof(float4) * (size_t)m);
  if (e == cudaSuccess) e = devAlloc(ctx, &ix->cellStart, sizeof(uint32_t) * (numCells + 1));
  if (e != cudaSuccess) {
    b200reg_index_free(ctx, ix);
    return setErr(ctx, B200REG_E_OOM, "index_build: allocation failed (%zu cells, %lld points): %s", numCells, (long long)m, cudaGetErrorString(e));
  }
  TRY(ensureBuf(ctx, ctx->keys0, sizeof(uint32_t) * (size_t)m));
  TRY(ensureBuf(ctx, ctx->keys1, sizeof(uint32_t) * (size_t)m));
  TRY(ensureBuf(ctx, ctx->vals0, sizeof(uint32_t) * (size_t)m));
  TRY(ensureBuf(ctx, ctx->vals1, sizeof(uint32_t) * (size_t)m));
  uint32_t* counts = ix->cellStart;   // counted in place, then scanned in place
  CUDA_TRY(ctx, cudaMemsetAsync(counts, 0, sizeof(uint32_t) * (numCells + 1), ctx->stream));
  keyKernel<<<divUp(m, 256), 256, 0, ctx->stream>>>(ref->xyz, m, mean[0], mean[1], mean[2], g, (uint32_t*)ctx->keys0.p,
                                                    (uint32_t*)ctx->vals0.p, counts);
  LAUNCH_CHECK(ctx);
  int endBit = 1;
  while (endBit < 32 && ((size_t)1 << endBit) < numCells) ++endBit;
  cub::DoubleBuffer<uint32_t> dk((uint32_t*)ctx->keys0.p, (uint32_t*)ctx->keys1.p), dv((uint32_t*)ctx->vals0.p, (uint32_t*)ctx->vals1.p);
  size_t tmpSort = 0, tmpScan = 0;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmpSort, dk, dv, (int)m, 0, endBit, ctx->stream));
  CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmpScan, counts, counts, (int64_t)(numCells + 1), ctx->stream));
  TRY(ensureBuf(ctx, ctx->cubtmp, tmpSort > tmpScan ? tmpSort : tmpScan));
  size_t tmpBytes = ctx->cubtmp.cap;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(ctx->cubtmp.p, tmpBytes, dk, dv, (int)m, 0, endBit, ctx->stream));
  ctx->launches += (uint64_t)((endBit + 7) / 8 + 2);
  tmpBytes = ctx->cubtmp.cap;
  CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(ctx->cubtmp.p, tmpBytes, counts, counts, (int64_t)(numCells + 1), ctx->stream));
  ctx->launches += 2;
  gatherKernel<<<divUp(m, 256), 256, 0, ctx->stream>>>(ref->xyz, ref->nrm, dv.Current(), m, mean[0], mean[1], mean[2], ix->pts, ix->nrm);
  LAUNCH_CHECK(ctx);
  *out = ix;
  return B200REG_OK;
}

extern "C" {

int32_t b200reg_index_build(b200reg_ctx* ctx, const b200reg_cloud* ref, float cell, b200reg_index** out) {
  if (!ctx || !ref || !out) return setErr(ctx, B200REG_E_INVALID_PARAM, "index_build: bad arguments");
  *out = nullptr;
  if (ref->n <= 0) return setErr(ctx, B200REG_E_NO_POINTS, "index_build: reference cloud is empty");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  double sum[3];
  float mn[3], mx[3];
  TRY(cloudStats(ctx, ref, sum, mn, mx));
  const float zero[3] = {0.f, 0.f, 0.f};
  TRY(buildIndex(ctx, ref, zero, mn, mx, cell, out));
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev1));
  float ms = 0;
  cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
  (*out)->build_ms = ms;
  return B200REG_OK;
}

void b200reg_index_free(b200reg_ctx* ctx, b200reg_index* index) {
  if (!index) return;
  if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
  if (index->pts) devFree(ctx, index->pts);
  if (index->nrm) devFree(ctx, index->nrm);
  if (index->cellStart) devFree(ctx, index->cellStart);
  delete index;
}

int64_t b200reg_index_size(const b200reg_index* index) { return index ? index->m : 0; }
float b200reg_index_cell(const b200reg_index* index) { return index ? index->g.cell : 0.f; }

}  // extern "C"
