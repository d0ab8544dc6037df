// csrc/index.cu — KDTreeMatcher::init replacement (MatchersImpl.cpp:109-114): a counting-sorted voxel-grid index, built
// without a host round trip.
//
// Layout in HBM: the (centred) reference is re-ordered by linear cell id (x fastest) into
//   pts[M]  float4 = (x, y, z, original index as int bits)   — one 16 B coalesced load yields position AND id
//   nrm[M]  float4 = reference normals in the same order
//   cellStart[numCells + 1] uint32                           — CSR directory; x-adjacent cells are contiguous ranges
//   hdr     IndexHeader (device)                             — grid geometry + mean, derived ON THE DEVICE from the cloud's bbox
//
// Build (all on the ctx stream, nothing waits on the host):
//   statsKernel      fp64 coordinate sums + bbox per block                                   16 B/pt read
//   gridKernel       one block: mean (fp64 sum rounded once to fp32, ICP.cpp:313), bbox of the centred cloud, cell size, grid
//   zeroDirKernel    directory + scan tile states
//   countKernel      cell of fl(p - mean) (ICP.cpp:320); rank of the point inside its cell = the value the atomic returns   16 B/pt read, 8 B/pt written
//   scanDirKernel    single-pass inclusive scan of the counts (decoupled look-back), length read from the header        8 B/cell
//   permKernel       perm[cellStart[cell] + rank] = original index (4-byte scatter into an L2-resident array)           12 B/pt read, 4 B/pt written
//   gatherPtsKernel  pts[j] = (centred point perm[j], index): coalesced streaming writes, 16-byte gathers from an L2-resident array   20 B/pt read, 16 B/pt written
//   gatherNrmKernel  the normals alike, in a pass of their own                                                          20 B/pt read, 16 B/pt written
// The order of the points INSIDE a cell is the arrival order of the atomics; nothing downstream depends on it: every search
// resolves ties by (d2, original index) and the reductions run in reading order (knn.cu, icp_kernels.cuh).
// The host learns the grid (for b200reg_index_cell, error reporting) from an asynchronous copy of the header, waited for only
// when somebody asks.
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

// One block: block partials -> mean, bbox of the centred cloud, cell size, grid dimensions (what the host used to do between two
// synchronisations).  The partials are added in block order by one thread per quantity, so the mean is the same whatever ran
// beside this kernel.  Non-finite coordinates (a LiDAR driver's +-inf "no return") leave status = 1 and a 1-cell grid, so the
// kernels behind this one run harmlessly and the host reports InvalidParameter when it reads the header.
__global__ void __launch_bounds__(288) gridKernel(const StatsPartial* __restrict__ partials, int blocks, long long m, int centre, float userCell,
                                                  unsigned capCells, IndexHeader* __restrict__ hdr) {
  __shared__ double sSum[3];
  __shared__ float sMn[3], sMx[3];
  // nine warps, one quantity each (3 sums, 3 minima, 3 maxima): lane l adds partials l, l + 32, ... in order, then a shuffle tree —
  // a fixed order, so the mean does not depend on what ran beside this kernel
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < 3) {
    double s = 0;
    for (int b = lane; b < blocks; b += 32) s += (warp == 0) ? partials[b].sx : (warp == 1) ? partials[b].sy : partials[b].sz;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) sSum[warp] = s;
  } else if (warp < 6) {
    const int d = warp - 3;
    float a = FLT_MAX;
    for (int b = lane; b < blocks; b += 32) a = fminf(a, d == 0 ? partials[b].mnx : d == 1 ? partials[b].mny : partials[b].mnz);
    for (int o = 16; o > 0; o >>= 1) a = fminf(a, __shfl_xor_sync(0xffffffffu, a, o));
    if (lane == 0) sMn[d] = a;
  } else {
    const int d = warp - 6;
    float z = -FLT_MAX;
    for (int b = lane; b < blocks; b += 32) z = fmaxf(z, d == 0 ? partials[b].mxx : d == 1 ? partials[b].mxy : partials[b].mxz);
    for (int o = 16; o > 0; o >>= 1) z = fmaxf(z, __shfl_xor_sync(0xffffffffu, z, o));
    if (lane == 0) sMx[d] = z;
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  IndexHeader h;
  h.m = m;
  bool finite = true;
  for (int d = 0; d < 3; ++d) {
    h.sum[d] = sSum[d];
    h.bbMin[d] = sMn[d]; h.bbMax[d] = sMx[d];
    // meanReference = features.rowwise().mean() (ICP.cpp:313): fp64 sum rounded once to fp32 (contract)
    h.mean[d] = centre ? (float)(sSum[d] / (double)m) : 0.f;
    finite = finite && (fabs(sSum[d]) <= DBL_MAX) && (fabsf(sMn[d]) <= FLT_MAX) && (fabsf(sMx[d]) <= FLT_MAX);
  }
  h.status = finite ? 0 : 1;
  GridDesc& g = h.g;
  if (!finite) {
    for (int d = 0; d < 3; ++d) h.mean[d] = 0.f;
    g.ox = g.oy = g.oz = 0.f; g.cell = 1.f; g.inv_cell = 1.f; g.slack = 0.f; g.ring2Safe = 0.f; g.nx = g.ny = g.nz = 1;
    h.numCells = 1;
    *hdr = h;
    return;
  }
  // bbox of the centred cloud: fl(min - mean) .. fl(max - mean) (fp32 subtraction is monotone)
  float lo[3], hi[3], ext[3];
  float maxAbs = 0.f;
  for (int d = 0; d < 3; ++d) {
    lo[d] = h.bbMin[d] - h.mean[d];
    hi[d] = h.bbMax[d] - h.mean[d];
    ext[d] = fmaxf(hi[d] - lo[d], 1e-6f);
    maxAbs = fmaxf(maxAbs, fmaxf(fabsf(lo[d]), fabsf(hi[d])));
  }
  float cell = userCell;
  const float emax = fmaxf(ext[0], fmaxf(ext[1], ext[2]));
  if (!(cell > 0.f) || !(cell <= FLT_MAX)) {   // automatic: ~2.5 x the mean spacing of points spread over the bbox surface
    const double area = 2.0 * ((double)ext[0] * ext[1] + (double)ext[1] * ext[2] + (double)ext[0] * ext[2]);
    cell = (float)(2.5 * sqrt(area / (double)m));
    cell = fmaxf(cell, emax * 1e-4f);
  }
  cell = fmaxf(cell, emax * 1e-6f);   // a user cell far below the extent would overflow the int cell coordinates
  for (int it = 0; it < 400; ++it) {   // keep the directory inside its allocation (capCells, chosen by the host from m alone)
    const double nc = (floor((double)(ext[0] / cell)) + 1) * (floor((double)(ext[1] / cell)) + 1) * (floor((double)(ext[2] / cell)) + 1);
    if (nc <= (double)capCells) break;
    cell *= 1.26f;
  }
  g.ox = lo[0]; g.oy = lo[1]; g.oz = lo[2];
  g.cell = cell;
  g.inv_cell = 1.0f / cell;
  g.nx = (int)floorf(ext[0] / cell) + 1;
  g.ny = (int)floorf(ext[1] / cell) + 1;
  g.nz = (int)floorf(ext[2] / cell) + 1;
  // slack >= error of fl((p-o)*inv) cell assignment + error of evaluating cell bounds in fp32 (see knn.cu)
  g.slack = 16.f * 1.1920929e-7f * (maxAbs + emax + cell);
  {   // (cell - 4 slack)^2 with a 1 % margin, rounded toward zero: a point >= 2 cells away cannot have a smaller fp32 d2
    const double c = (double)cell - 4.0 * (double)g.slack;
    g.ring2Safe = (c > 0.0) ? nextafterf((float)(0.99 * c * c), 0.f) : 0.f;
  }
  h.numCells = (unsigned)g.nx * (unsigned)g.ny * (unsigned)g.nz;
  *hdr = h;
}

// ---- directory: counts -> inclusive scan, single pass ------------------------------------------------------------------------
#define SCAN_THREADS 256
#define SCAN_ITEMS 16
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)
// tile state: bits [63:62] = 0 empty, 1 aggregate of this tile only, 2 inclusive prefix; bits [31:0] = value (< 2^31)
#define TS_AGG (1ull << 62)
#define TS_INC (2ull << 62)

// Directory allocation: dirBase = [3 pad words | cellStart[0] | counts of cells 0 .. numCells-1 | padding to a whole scan tile];
// cellStart = dirBase + 3, so the counts (cellStart + 1 = dirBase + 4) are 16-byte aligned for the scan's vector accesses.
// zeroes dirBase[0 .. 4 + numCells) (rounded up to 16 bytes), the tile states of the scan and its tile counter
__global__ void zeroDirKernel(const IndexHeader* __restrict__ hdr, uint32_t* __restrict__ dirBase, unsigned long long* __restrict__ tileState,
                              unsigned* __restrict__ tileCounter) {
  const unsigned numCells = hdr->numCells;
  const size_t quads = ((size_t)numCells + 4 + 3) / 4;   // dirBase is 16-byte aligned (pool allocation) and padded
  const size_t stride = (size_t)gridDim.x * blockDim.x, t0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint4* d4 = reinterpret_cast<uint4*>(dirBase);
  for (size_t i = t0; i < quads; i += stride) d4[i] = make_uint4(0u, 0u, 0u, 0u);
  const size_t tiles = ((size_t)numCells + SCAN_TILE - 1) / SCAN_TILE;
  for (size_t i = t0; i < tiles; i += stride) tileState[i] = 0ull;
  if (t0 == 0) *tileCounter = 0u;
}

__device__ __forceinline__ int cellCoord(float p, float o, float inv, int nmax) {
  const int c = (int)floorf(__fmul_rn(__fsub_rn(p, o), inv));
  return min(max(c, 0), nmax - 1);
}
__device__ __forceinline__ uint32_t cellKeyOf(const float4 p, const float mx, const float my, const float mz, const GridDesc& g, float& x, float& y,
                                              float& z) {
  // centred point = fl(p - mean) exactly as ICP.cpp:320 does; key = linear cell (x fastest)
  x = __fsub_rn(p.x, mx); y = __fsub_rn(p.y, my); z = __fsub_rn(p.z, mz);
  const int cx = cellCoord(x, g.ox, g.inv_cell, g.nx), cy = cellCoord(y, g.oy, g.inv_cell, g.ny), cz = cellCoord(z, g.oz, g.inv_cell, g.nz);
  return (uint32_t)cx + (uint32_t)g.nx * ((uint32_t)cy + (uint32_t)g.ny * (uint32_t)cz);
}

__global__ void __launch_bounds__(256) countKernel(const float4* __restrict__ xyz, int64_t n, const IndexHeader* __restrict__ hdr,
                                                   uint32_t* __restrict__ cellStart, uint2* __restrict__ keyRank) {
  const GridDesc g = hdr->g;
  const float mx = hdr->mean[0], my = hdr->mean[1], mz = hdr->mean[2];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float x, y, z;
    const uint32_t key = cellKeyOf(xyz[i], mx, my, mz, g, x, y, z);
    const uint32_t r = atomicAdd(&cellStart[key + 1], 1u);   // counts live one slot up: the scan turns them into the next cell's start
    keyRank[i] = make_uint2(key, r);
  }
}

// Inclusive scan of the counts, cellStart[1 .. numCells], in place (cellStart[0] stays 0): afterwards cellStart[c] = first sorted
// position of cell c and cellStart[numCells] = m.  One pass over the directory: tiles are handed out by a counter (so a tile's predecessors are always
// running or done), each publishes its aggregate, then looks back over its predecessors' states a warp at a time.
__global__ void __launch_bounds__(SCAN_THREADS) scanDirKernel(const IndexHeader* __restrict__ hdr, uint32_t* __restrict__ cellStart,
                                                              unsigned long long* __restrict__ tileState, unsigned* __restrict__ tileCounter) {
  const unsigned numCells = hdr->numCells;
  const unsigned tiles = (numCells + SCAN_TILE - 1) / SCAN_TILE;
  uint32_t* __restrict__ data = cellStart + 1;   // 16-byte aligned (zeroDirKernel)
  __shared__ unsigned sTile, sWarp[SCAN_ITEMS / 4][SCAN_THREADS / 32], sPrefix;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (;;) {
    __syncthreads();   // sTile / sPrefix / sWarp of the previous round are no longer read
    if (threadIdx.x == 0) sTile = atomicAdd(tileCounter, 1u);
    __syncthreads();
    const unsigned tile = sTile;
    if (tile >= tiles) return;
    // group (j, tid) = the four words at tile base + 4 (j * SCAN_THREADS + tid): consecutive threads load consecutive 16 bytes.
    // The allocation is padded to a whole tile; words beyond numCells are ignored.
    const size_t tbase = (size_t)tile * SCAN_TILE;
    constexpr int NG = SCAN_ITEMS / 4;
    uint4 q[NG];
#pragma unroll
    for (int j = 0; j < NG; ++j) q[j] = __ldcg(reinterpret_cast<const uint4*>(data + tbase) + j * SCAN_THREADS + threadIdx.x);
    unsigned gsum[NG], gincl[NG];
#pragma unroll
    for (int j = 0; j < NG; ++j) {
      const size_t w0 = tbase + 4 * ((size_t)j * SCAN_THREADS + threadIdx.x);
      if (w0 + 0 >= numCells) q[j].x = 0u;
      if (w0 + 1 >= numCells) q[j].y = 0u;
      if (w0 + 2 >= numCells) q[j].z = 0u;
      if (w0 + 3 >= numCells) q[j].w = 0u;
      q[j].y += q[j].x; q[j].z += q[j].y; q[j].w += q[j].z;   // inclusive inside the group
      gsum[j] = q[j].w;
      unsigned incl = gsum[j];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      gincl[j] = incl;
      if (lane == 31) sWarp[j][warp] = incl;
    }
    __syncthreads();
    // offset of group (j, tid) inside the tile = all groups of rows j' < j + the warps before this one in row j + the lanes before
    unsigned goff[NG], tileSum = 0;
#pragma unroll
    for (int j = 0; j < NG; ++j) {
      unsigned before = 0, row = 0;
#pragma unroll
      for (int w = 0; w < SCAN_THREADS / 32; ++w) {
        const unsigned sw = sWarp[j][w];
        if (w < warp) before += sw;
        row += sw;
      }
      goff[j] = tileSum + before + (gincl[j] - gsum[j]);
      tileSum += row;
    }
    if (warp == 0) {   // publish, then look back
      if (lane == 0) {
        __threadfence();
        atomicExch(&tileState[tile], (tile == 0 ? TS_INC : TS_AGG) | (unsigned long long)tileSum);
      }
      unsigned prefix = 0;
      if (tile > 0) {
        int look = (int)tile - 1;
        for (;;) {
          const int idx = look - lane;
          unsigned long long s = TS_INC;   // lanes before tile 0 behave like an inclusive prefix of 0
          if (idx >= 0) {
            do { s = *((volatile unsigned long long*)&tileState[idx]); } while ((s >> 62) == 0ull);
          }
          const unsigned incMask = __ballot_sync(0xffffffffu, (s >> 62) == 2ull);
          const int firstInc = incMask ? (__ffs(incMask) - 1) : 32;   // nearest predecessor holding an inclusive prefix
          unsigned contrib = (lane <= firstInc) ? (unsigned)s : 0u;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
          prefix += contrib;
          if (incMask) break;
          look -= 32;
        }
        if (lane == 0) {
          __threadfence();
          atomicExch(&tileState[tile], TS_INC | (unsigned long long)(prefix + tileSum));
        }
      }
      if (lane == 0) sPrefix = prefix;
    }
    __syncthreads();
    const unsigned prefix = sPrefix;
#pragma unroll
    for (int j = 0; j < NG; ++j) {   // words beyond numCells land in the allocation's padding
      const unsigned off = prefix + goff[j];
      reinterpret_cast<uint4*>(data + tbase)[j * SCAN_THREADS + threadIdx.x] = make_uint4(q[j].x + off, q[j].y + off, q[j].z + off, q[j].w + off);
    }
  }
}

// perm[sorted position] = original index.  The random 4-byte writes land in a 4 M-byte array that stays in L2; scattering the
// 32-byte payload instead was measured at 313 us for 5 M points (partial-sector writes), this + the gather below at ~180 us.
__global__ void __launch_bounds__(256) permKernel(const uint2* __restrict__ keyRank, int64_t n, const uint32_t* __restrict__ cellStart,
                                                  uint32_t* __restrict__ perm) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint2 kr = keyRank[i];
    perm[cellStart[kr.x] + kr.y] = (uint32_t)i;
  }
}

// sorted payload: pts[j] = (fl(p - mean), original index), then the normals in a pass of their own.  One pass over both arrays
// gathers from 32 B/pt of input — more than the L2 holds for a 5 M-point map, so half-used sectors are fetched twice; one array
// at a time (16 B/pt, L2-resident) with the output streamed past the cache fetches every input sector once.
__device__ __forceinline__ float4 ldGather(const float4* p) {   // keep the gathered line (its other half is some other thread's point)
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__global__ void __launch_bounds__(256) gatherPtsKernel(const float4* __restrict__ xyz, int64_t n, const IndexHeader* __restrict__ hdr,
                                                       const uint32_t* __restrict__ perm, float4* __restrict__ pts) {
  const float mx = hdr->mean[0], my = hdr->mean[1], mz = hdr->mean[2];
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t i = __ldcs(perm + j);
    const float4 p = ldGather(xyz + i);
    __stcs(pts + j, make_float4(__fsub_rn(p.x, mx), __fsub_rn(p.y, my), __fsub_rn(p.z, mz), __int_as_float((int)i)));
  }
}
__global__ void __launch_bounds__(256) gatherNrmKernel(const float4* __restrict__ nrm, int64_t n, const uint32_t* __restrict__ perm,
                                                       float4* __restrict__ outN) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x)
    __stcs(outN + j, ldGather(nrm + __ldcs(perm + j)));
}

}  // namespace

// shared with voxelgrid.cu: mean sums and bbox of a cloud on the HOST (one synchronisation; the index build does not use this)
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

static void indexRelease(b200reg_ctx* ctx, b200reg_index* ix) {
  if (!ix) return;
  if (ix->hdrReady) {
    cudaEventSynchronize(ix->hdrReady);   // the header copy targets ix->hHdr: it must have landed before the slot is reused
    cudaEventDestroy(ix->hdrReady);
  }
  if (ix->hHdr) releaseHeaderSlot(ctx, ix->hHdr);
  if (ix->pts) devFree(ctx, ix->pts);
  if (ix->nrmStore) devFree(ctx, ix->nrmStore);
  if (ix->dirBase) devFree(ctx, ix->dirBase);
  if (ix->dHdr) devFree(ctx, ix->dHdr);
  delete ix;
}

// Enqueue the build of an index over `ref` (centre != 0: after subtracting the cloud's mean, ICP.cpp:313-320).  Nothing here
// waits for the device; `ref` must stay alive until the stream has passed this point (stream-ordered frees on the same ctx do).
int buildIndex(b200reg_ctx* ctx, const b200reg_cloud* ref, int centre, float cell, b200reg_index** out, b200reg_index* recycle) {
  const int64_t m = ref->n;
  *out = nullptr;
  if (m <= 0) { indexRelease(ctx, recycle); return setErr(ctx, B200REG_E_NO_POINTS, "index_build: reference cloud is empty"); }
  if (m >= (int64_t)0x7fffffff) { indexRelease(ctx, recycle); return setErr(ctx, B200REG_E_INVALID_PARAM, "index_build: more than 2^31-1 points"); }
  // directory capacity from m alone (the host does not know the bbox): 8 cells per point, at least 2^16, at most 2^30; the
  // device grows the cell until the grid fits (the automatic cell gives ~4 cells per point on surface-like clouds)
  const uint64_t cap = std::min<uint64_t>(std::max<uint64_t>(8ull * (uint64_t)m, 1ull << 16), 1ull << 30);
  b200reg_index* ix = nullptr;
  bool recycled = false;   // (not `ix == recycle`: a freshly allocated index may land at the address of the one just released)
  if (recycle && recycle->mCap >= m && recycle->cellCap >= cap && recycle->cellCap <= 4 * cap && (recycle->nrm || !ref->nrm)) {
    ix = recycle;   // same stream: everything enqueued against the old content runs before the kernels below
    recycled = true;
    if (!ix->hdrFetched) cudaEventSynchronize(ix->hdrReady);   // the old header copy must have landed before the slot is rewritten
    ix->hdrFetched = false;
    ix->hasNrm = ref->nrm != nullptr;
  } else {
    indexRelease(ctx, recycle);
    ix = new b200reg_index();
  }
  ix->m = m;
  const size_t tilesCap = ((recycled ? (uint64_t)ix->cellCap : cap) + SCAN_TILE - 1) / SCAN_TILE;
  const int sblocks = (int)fmin((double)(4 * ctx->sm_count), (double)divUp(m, 256));
  int st = B200REG_OK;
  cudaError_t e = cudaSuccess;
  auto fail = [&](int code) { indexRelease(ctx, ix); return code; };
  if (!recycled) {
    ix->cellCap = (uint32_t)cap;
    ix->mCap = m + m / 8;   // head-room: a periodically rebuilt reference (the Mapper's map patch) changes size a little every time
    ix->hasNrm = ref->nrm != nullptr;
    e = devAlloc(ctx, &ix->pts, sizeof(float4) * (size_t)ix->mCap);
    if (e == cudaSuccess && ref->nrm) e = devAlloc(ctx, &ix->nrmStore, sizeof(float4) * (size_t)ix->mCap);
    if (e == cudaSuccess) e = devAlloc(ctx, &ix->dirBase, sizeof(uint32_t) * ((size_t)tilesCap * SCAN_TILE + 8));
    ix->cellStart = ix->dirBase ? ix->dirBase + 3 : nullptr;
    if (e == cudaSuccess) e = devAlloc(ctx, &ix->dHdr, sizeof(IndexHeader));
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ix->hdrReady, cudaEventDisableTiming);
    if (e != cudaSuccess)
      return fail(setErr(ctx, B200REG_E_OOM, "index_build: allocation failed (%llu directory cells, %lld points): %s", (unsigned long long)cap, (long long)m, cudaGetErrorString(e)));
    ix->hHdr = acquireHeaderSlot(ctx);
    if (!ix->hHdr) return fail(setErr(ctx, B200REG_E_OOM, "index_build: no pinned header slot"));
  }
  ix->nrm = ix->hasNrm ? ix->nrmStore : nullptr;
  if ((st = ensureBuf(ctx, ctx->partials, sizeof(StatsPartial) * (size_t)sblocks)) != B200REG_OK) return fail(st);
  if ((st = ensureBuf(ctx, ctx->keys0, sizeof(uint2) * (size_t)m)) != B200REG_OK) return fail(st);
  if ((st = ensureBuf(ctx, ctx->vals0, sizeof(uint32_t) * (size_t)m)) != B200REG_OK) return fail(st);
  if ((st = ensureBuf(ctx, ctx->cubtmp, sizeof(unsigned long long) * tilesCap + 64)) != B200REG_OK) return fail(st);
  unsigned long long* tileState = (unsigned long long*)ctx->cubtmp.p;
  unsigned* tileCounter = (unsigned*)((char*)ctx->cubtmp.p + sizeof(unsigned long long) * tilesCap);
  uint2* keyRank = (uint2*)ctx->keys0.p;
  uint32_t* perm = (uint32_t*)ctx->vals0.p;
  const int wide = 8 * ctx->sm_count;   // grid of the streaming passes: 8 blocks of 256 threads per SM
  const int pblocks = (int)std::min<int64_t>(divUp(m, 256), wide);
#define IDX_STEP(name)                                                                                                   \
  do {                                                                                                                   \
    ctx->launches++;                                                                                                     \
    if (ctx->debugSync) {                                                                                                \
      const cudaError_t de = cudaStreamSynchronize(ctx->stream);                                                         \
      if (de != cudaSuccess) return fail(setErr(ctx, B200REG_E_CUDA, "index_build: %s: %s", name, cudaGetErrorString(de))); \
    }                                                                                                                    \
  } while (0)
  statsKernel<<<sblocks, 256, 0, ctx->stream>>>(ref->xyz, m, (StatsPartial*)ctx->partials.p);
  IDX_STEP("statsKernel");
  gridKernel<<<1, 288, 0, ctx->stream>>>((const StatsPartial*)ctx->partials.p, sblocks, (long long)m, centre, cell, ix->cellCap, ix->dHdr);
  IDX_STEP("gridKernel");
  zeroDirKernel<<<(int)std::min<uint64_t>(wide, (cap / 4 + 255) / 256), 256, 0, ctx->stream>>>(ix->dHdr, ix->dirBase, tileState, tileCounter);
  IDX_STEP("zeroDirKernel");
  countKernel<<<pblocks, 256, 0, ctx->stream>>>(ref->xyz, m, ix->dHdr, ix->cellStart, keyRank);
  IDX_STEP("countKernel");
  scanDirKernel<<<(int)std::min<uint64_t>(4 * ctx->sm_count, tilesCap), SCAN_THREADS, 0, ctx->stream>>>(ix->dHdr, ix->cellStart, tileState, tileCounter);
  IDX_STEP("scanDirKernel");
  permKernel<<<pblocks, 256, 0, ctx->stream>>>(keyRank, m, ix->cellStart, perm);
  IDX_STEP("permKernel");
  gatherPtsKernel<<<pblocks, 256, 0, ctx->stream>>>(ref->xyz, m, ix->dHdr, perm, ix->pts);
  IDX_STEP("gatherPtsKernel");
  if (ref->nrm) {
    gatherNrmKernel<<<pblocks, 256, 0, ctx->stream>>>(ref->nrm, m, perm, ix->nrm);
    IDX_STEP("gatherNrmKernel");
  }
#undef IDX_STEP
  e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(ix->hHdr, ix->dHdr, sizeof(IndexHeader), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaEventRecord(ix->hdrReady, ctx->stream);
  if (e != cudaSuccess) return fail(setErr(ctx, B200REG_E_CUDA, "index_build: %s", cudaGetErrorString(e)));
  *out = ix;
  return B200REG_OK;
}

// Host view of the header: waits (once) for the asynchronous copy enqueued by the build.  Returns nullptr with the ctx error set
// when the build failed on the device (non-finite coordinates).
const IndexHeader* indexHeader(b200reg_ctx* ctx, const b200reg_index* cix) {
  b200reg_index* ix = const_cast<b200reg_index*>(cix);
  if (!ix->hdrFetched) {
    const cudaError_t e = cudaEventSynchronize(ix->hdrReady);
    if (e != cudaSuccess) { setErr(ctx, B200REG_E_CUDA, "index_build failed: %s", cudaGetErrorString(e)); return nullptr; }
    ix->hdrFetched = true;
  }
  if (ix->hHdr->status != 0) {
    setErr(ctx, B200REG_E_INVALID_PARAM, "index_build: non-finite coordinates in the reference cloud");
    return nullptr;
  }
  return ix->hHdr;
}

extern "C" {

int32_t b200reg_index_build(b200reg_ctx* ctx, const b200reg_cloud* ref, float cell, b200reg_index** out) {
  if (!ctx || !ref || !out) return setErr(ctx, B200REG_E_INVALID_PARAM, "index_build: bad arguments");
  *out = nullptr;
  if (ref->n <= 0) return setErr(ctx, B200REG_E_NO_POINTS, "index_build: reference cloud is empty");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  b200reg_index* ix = nullptr;
  TRY(buildIndex(ctx, ref, 0, cell, &ix));
  if (!indexHeader(ctx, ix)) {   // this entry point reports device-side failures at once
    indexRelease(ctx, ix);
    return B200REG_E_INVALID_PARAM;
  }
  *out = ix;
  return B200REG_OK;
}

void b200reg_index_free(b200reg_ctx* ctx, b200reg_index* index) {
  if (!index) return;
  if (ctx) cudaSetDevice(ctx->device);
  indexRelease(ctx, index);   // stream-ordered frees: kernels already enqueued on the ctx stream keep their memory
}

int64_t b200reg_index_size(const b200reg_index* index) { return index ? index->m : 0; }
float b200reg_index_cell(const b200reg_index* index) {
  if (!index) return 0.f;
  const IndexHeader* h = indexHeader(nullptr, index);
  return h ? h->g.cell : 0.f;
}

}  // extern "C"
