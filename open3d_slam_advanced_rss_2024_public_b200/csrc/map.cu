// csrc/map.cu — open3d_slam voxel-map crop / insert on the device, fp64 (paths relative to
// /root/reference/open3d_slam_rsl/open3d_slam/open3d_slam/):
//   CroppingVolume::isWithinVolume / crop        src/croppers.cpp:53-167
//   o3d_slam::transform                          src/helpers.cpp:283-318
//   Submap::insertScan                           src/Submap.cpp:39-96
//   voxelizeWithinCroppingVolume                 src/helpers.cpp:117-192, key = int(floor(p * (1/v))) (VoxelHashMap.hpp:43-51)
//   ScanToMapRegistration::cropSubmap            src/ScanToMapRegistration.cpp:90-96
//
// The map stays resident as plain fp64 SoA-of-3 arrays (a cloud of per-voxel means, exactly the reference's
// `mapCloud_`).  Insert = transform + append + [predicate -> compaction of pass-through points | 63-bit key radix sort of
// the points inside the volume -> one thread per voxel accumulates its points in ascending input order (the reference's
// accumulation order, so means are bit-exact)].  Output order: pass-through points in input order, then voxels in
// ascending (x,y,z) key order (the reference's is unordered_map order — implementation-defined).
// Compiled with --fmad=false (each fp64 op individually rounded, like the oracle).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cmath>

#include "common.cuh"

struct b200reg_map {
  double voxel = 0.0;
  int64_t n = 0, cap = 0;
  double* xyz = nullptr;   // 3 doubles per point
  double* nrm = nullptr;   // 3 doubles per point (allocated with xyz; valid iff hasNormals)
  double* xyz2 = nullptr;  // double buffer for the rebuilt map
  double* nrm2 = nullptr;
  bool hasNormals = false;
};

namespace {

typedef unsigned long long u64;

__device__ __forceinline__ bool withinVolume(const b200reg_crop_spec& s, double x, double y, double z) {
  const double dx = x - s.center[0], dy = y - s.center[1], dz = z - s.center[2];
  bool in;
  if (s.kind == B200REG_CROP_MAX_RADIUS) in = sqrt(dx * dx + dy * dy + dz * dz) <= s.rmax;
  else if (s.kind == B200REG_CROP_MIN_RADIUS) in = sqrt(dx * dx + dy * dy + dz * dz) >= s.rmin;
  else if (s.kind == B200REG_CROP_MINMAX_RADIUS) { const double d = sqrt(dx * dx + dy * dy + dz * dz); in = d <= s.rmax && d >= s.rmin; }
  else in = z >= s.zmin && z <= s.zmax && sqrt(dx * dx + dy * dy) <= s.rmax;   // cylinder: absolute z range (croppers.cpp:166)
  return s.invert ? !in : in;
}

// p' = (T [p;1]).head3 / w ; n' = (T [n;0]).head3   (helpers.cpp:296-309)
__global__ void transformF64Kernel(const double* __restrict__ xyz, const double* __restrict__ nrm, int64_t n, const double* __restrict__ Tm,
                                   double* __restrict__ oxyz, double* __restrict__ onrm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double T[16];
  for (int j = 0; j < 16; ++j) T[j] = Tm[j];
  const double x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
  double q[4];
  for (int r = 0; r < 4; ++r) q[r] = ((T[4 * r] * x + T[4 * r + 1] * y) + T[4 * r + 2] * z) + T[4 * r + 3] * 1.0;
  for (int r = 0; r < 3; ++r) oxyz[3 * i + r] = q[r] / q[3];
  if (nrm && onrm) {
    const double a = nrm[3 * i], b = nrm[3 * i + 1], c = nrm[3 * i + 2];
    for (int r = 0; r < 3; ++r) onrm[3 * i + r] = ((T[4 * r] * a + T[4 * r + 1] * b) + T[4 * r + 2] * c) + T[4 * r + 3] * 0.0;
  }
}

// inside flag (+ voxel key for inside points); outFlag = 1 - inside for the pass-through compaction
__global__ void classifyKernel(const double* __restrict__ xyz, int64_t n, b200reg_crop_spec spec, double inv, unsigned* __restrict__ inFlag,
                               unsigned* __restrict__ outFlag, u64* __restrict__ keys, int* __restrict__ err) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
  const bool in = withinVolume(spec, x, y, z);
  inFlag[i] = in ? 1u : 0u;
  if (outFlag) outFlag[i] = in ? 0u : 1u;
  if (keys) {
    u64 key = ~0ull;
    if (in) {
      const double fx = floor(x * inv), fy = floor(y * inv), fz = floor(z * inv);   // getVoxelIdx (VoxelHashMap.hpp:48-51)
      const double lim = 1048575.0;
      if (!(fabs(fx) <= lim) || !(fabs(fy) <= lim) || !(fabs(fz) <= lim)) { *err = 1; }
      else key = ((u64)((long long)fx + 1048576ll) << 42) | ((u64)((long long)fy + 1048576ll) << 21) | (u64)((long long)fz + 1048576ll);
    }
    keys[i] = key;
  }
}

// How the three voxel coordinates of a sort key are packed: x above y above z (ascending key = ascending (x, y, z), the emission
// order), each in the bits its range needs.  shy / shx = bits of z / of y + z; total = the sort's end bit.
struct KeyPack {
  long long k0[3];   // subtracted from the (offset) coordinates before packing
  int shx, shy, total;
};
inline int bitsFor(double span) {   // bits holding 0 .. span
  int b = 1;
  while (b < 21 && (double)(1ll << b) <= span) ++b;
  return b;
}

__global__ void compactPairsKernel(const unsigned* __restrict__ inFlag, const unsigned* __restrict__ inOff, const u64* __restrict__ keys,
                                   int64_t n, KeyPack kp, u64* __restrict__ ckeys, unsigned* __restrict__ cvals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !inFlag[i]) return;
  const unsigned o = inOff[i];
  u64 key = keys[i];
  if (kp.total < 63) {   // 3 x 21 bits -> the bits the cropping volume needs (same order: per-axis shifts keep (x, y, z) lexicographic)
    const long long kx = (long long)(key >> 42) - kp.k0[0], ky = (long long)((key >> 21) & 0x1fffffull) - kp.k0[1], kz = (long long)(key & 0x1fffffull) - kp.k0[2];
    key = ((u64)kx << kp.shx) | ((u64)ky << kp.shy) | (u64)kz;
  }
  ckeys[o] = key;
  cvals[o] = (unsigned)i;
}

// Key packing for the points inside a cropping volume: voxel coordinate c = floor(p / voxel) + 2^20 (classifyKernel) of a point inside a
// sphere / cylinder of radius rmax around `center` lies within floor((center -+ rmax) / voxel) + 2^20, one voxel of margin on each side.
// Unbounded volumes (inverted, min-radius) keep the 63-bit keys.
KeyPack keyPackFor(const b200reg_crop_spec& s, double inv) {
  KeyPack kp;
  kp.k0[0] = kp.k0[1] = kp.k0[2] = 0;
  kp.shx = 42; kp.shy = 21; kp.total = 63;
  if (s.invert || s.kind == B200REG_CROP_MIN_RADIUS || !(s.rmax >= 0.0) || !(s.rmax < 1e300)) return kp;
  double lo[3], hi[3];
  for (int d = 0; d < 3; ++d) { lo[d] = s.center[d] - s.rmax; hi[d] = s.center[d] + s.rmax; }
  if (s.kind == B200REG_CROP_CYLINDER) { lo[2] = s.zmin; hi[2] = s.zmax; }   // the cylinder's z range is absolute
  int bits[3];
  for (int d = 0; d < 3; ++d) {
    const double a = floor(lo[d] * inv) - 1.0, b = floor(hi[d] * inv) + 1.0;
    if (!(fabs(a) < 1048000.0) || !(fabs(b) < 1048000.0) || !(b >= a)) return kp;
    kp.k0[d] = (long long)a + 1048576ll;
    bits[d] = bitsFor(b - a);
  }
  kp.shy = bits[2];
  kp.shx = bits[2] + bits[1];
  kp.total = bits[0] + bits[1] + bits[2];
  return kp;
}

__global__ void headFlagKernel(const u64* __restrict__ skeys, int64_t m, unsigned* __restrict__ head) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  head[j] = (j == 0 || skeys[j] != skeys[j - 1]) ? 1u : 0u;
}

// AccumulatedPoint (helpers.cpp:30-73): one thread per voxel head walks the voxel's points in ascending input index
__global__ void voxelReduceKernel(const u64* __restrict__ skeys, const unsigned* __restrict__ svals, const unsigned* __restrict__ head,
                                  const unsigned* __restrict__ headOff, int64_t m, const double* __restrict__ xyz,
                                  const double* __restrict__ nrm, int64_t outBase, double* __restrict__ oxyz, double* __restrict__ onrm,
                                  int normalize) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m || !head[j]) return;
  const u64 key = skeys[j];
  double p[3] = {0, 0, 0}, a[3] = {0, 0, 0};
  int cnt = 0;
  for (int64_t t = j; t < m && skeys[t] == key; ++t) {
    const unsigned i = svals[t];
    p[0] += xyz[3 * (size_t)i]; p[1] += xyz[3 * (size_t)i + 1]; p[2] += xyz[3 * (size_t)i + 2];
    if (nrm) {
      const double nx = nrm[3 * (size_t)i], ny = nrm[3 * (size_t)i + 1], nz = nrm[3 * (size_t)i + 2];
      if (!isnan(nx) && !isnan(ny) && !isnan(nz)) { a[0] += nx; a[1] += ny; a[2] += nz; }   // helpers.cpp:35-37
    }
    ++cnt;
  }
  const int64_t o = outBase + headOff[j];
  const double c = (double)cnt;
  oxyz[3 * o] = p[0] / c; oxyz[3 * o + 1] = p[1] / c; oxyz[3 * o + 2] = p[2] / c;
  if (nrm && onrm) {
    const double m0 = a[0] / c, m1 = a[1] / c, m2 = a[2] / c;
    const double nn = normalize ? sqrt(m0 * m0 + m1 * m1 + m2 * m2) : 0.0;   // .normalized() (helpers.cpp:181); Open3D's VoxelDownSample keeps the plain mean
    onrm[3 * o] = (nn > 0.0) ? m0 / nn : m0;
    onrm[3 * o + 1] = (nn > 0.0) ? m1 / nn : m1;
    onrm[3 * o + 2] = (nn > 0.0) ? m2 / nn : m2;
  }
}

__global__ void scatterSelectedKernel(const unsigned* __restrict__ flag, const unsigned* __restrict__ off, int64_t n,
                                      const double* __restrict__ xyz, const double* __restrict__ nrm, double* __restrict__ oxyz,
                                      double* __restrict__ onrm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !flag[i]) return;
  const size_t o = off[i];
  for (int d = 0; d < 3; ++d) oxyz[3 * o + d] = xyz[3 * (size_t)i + d];
  if (nrm && onrm) for (int d = 0; d < 3; ++d) onrm[3 * o + d] = nrm[3 * (size_t)i + d];
}

// crop + open3dToPointmatcher (open3d_conversions.cpp:57-118): fp64 -> fp32 cloud, pad = 1
__global__ void scatterSelectedF32Kernel(const unsigned* __restrict__ flag, const unsigned* __restrict__ off, int64_t n,
                                         const double* __restrict__ xyz, const double* __restrict__ nrm, float4* __restrict__ oxyz,
                                         float4* __restrict__ onrm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !flag[i]) return;
  const size_t o = off[i];
  oxyz[o] = make_float4((float)xyz[3 * (size_t)i], (float)xyz[3 * (size_t)i + 1], (float)xyz[3 * (size_t)i + 2], 1.f);
  if (nrm && onrm) onrm[o] = make_float4((float)nrm[3 * (size_t)i], (float)nrm[3 * (size_t)i + 1], (float)nrm[3 * (size_t)i + 2], 0.f);
}

// per-block minimum (Open3D GetMinBound) and maximum of the coordinates: partial[6 b + 0..2] = min, [3..5] = max.  The maximum only
// sizes the sort: with the extent known the voxel keys pack into the bits they need and the radix sort runs 4-5 passes instead of 8.
__global__ void minBoundKernel(const double* __restrict__ xyz, int64_t n, double* __restrict__ partial) {
  double m[6] = {CUDART_INF, CUDART_INF, CUDART_INF, -CUDART_INF, -CUDART_INF, -CUDART_INF};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    for (int d = 0; d < 3; ++d) { const double v = xyz[3 * i + d]; m[d] = fmin(m[d], v); m[3 + d] = fmax(m[3 + d], v); }
  for (int o = 16; o > 0; o >>= 1)
    for (int d = 0; d < 3; ++d) {
      m[d] = fmin(m[d], __shfl_xor_sync(0xffffffffu, m[d], o));
      m[3 + d] = fmax(m[3 + d], __shfl_xor_sync(0xffffffffu, m[3 + d], o));
    }
  __shared__ double sm[8][6];
  if ((threadIdx.x & 31) == 0) for (int d = 0; d < 6; ++d) sm[threadIdx.x >> 5][d] = m[d];
  __syncthreads();
  if (threadIdx.x < 6) {
    double v = sm[0][threadIdx.x];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) v = (threadIdx.x < 3) ? fmin(v, sm[w][threadIdx.x]) : fmax(v, sm[w][threadIdx.x]);
    partial[6 * blockIdx.x + threadIdx.x] = v;
  }
}

// Open3D VoxelDownSample key: floor((p - voxel_min_bound) / voxel), voxel_min_bound = min_bound - voxel / 2
__global__ void o3dKeyKernel(const double* __restrict__ xyz, int64_t n, double mx, double my, double mz, double voxel, int shx, int shy,
                             u64* __restrict__ keys, unsigned* __restrict__ vals, int* __restrict__ err) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double fx = floor((xyz[3 * i] - mx) / voxel), fy = floor((xyz[3 * i + 1] - my) / voxel), fz = floor((xyz[3 * i + 2] - mz) / voxel);
  const double lim = 2097151.0;   // 21 bits per axis, keys are non-negative by construction
  u64 key = ~0ull >> 1;
  if (!(fx >= 0.0 && fx <= lim) || !(fy >= 0.0 && fy <= lim) || !(fz >= 0.0 && fz <= lim)) *err = 1;
  else key = ((u64)fx << shx) | ((u64)fy << shy) | (u64)fz;   // (a coordinate beyond its field could only come from a max bound that is not one: impossible)
  keys[i] = key;
  vals[i] = (unsigned)i;
}

int scanFlags(b200reg_ctx* ctx, const unsigned* flags, unsigned* offs, int64_t n) {
  size_t tmp = 0;
  CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp, flags, offs, (int)n, ctx->stream));
  TRY(ensureBuf(ctx, ctx->cubtmp, tmp));
  tmp = ctx->cubtmp.cap;
  CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(ctx->cubtmp.p, tmp, flags, offs, (int)n, ctx->stream));
  ctx->launches += 2;
  return B200REG_OK;
}

// total of an exclusive scan: offs[n-1] + flags[n-1] (synchronises)
int scanTotal(b200reg_ctx* ctx, const unsigned* flags, const unsigned* offs, int64_t n, int64_t* total) {
  unsigned* h = (unsigned*)ctx->pinned;
  CUDA_TRY(ctx, cudaMemcpyAsync(h, offs + (n - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(h + 1, flags + (n - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  *total = (int64_t)h[0] + (int64_t)h[1];
  return B200REG_OK;
}

int mapReserve(b200reg_ctx* ctx, b200reg_map* map, int64_t want) {
  if (want <= map->cap) return B200REG_OK;
  const int64_t cap = want + want / 2 + 1024;
  double* nx = nullptr; double* nn = nullptr; double* nx2 = nullptr; double* nn2 = nullptr;
  cudaError_t e = devAlloc(ctx, &nx, sizeof(double) * 3 * (size_t)cap);
  if (e == cudaSuccess) e = devAlloc(ctx, &nn, sizeof(double) * 3 * (size_t)cap);
  if (e == cudaSuccess) e = devAlloc(ctx, &nx2, sizeof(double) * 3 * (size_t)cap);
  if (e == cudaSuccess) e = devAlloc(ctx, &nn2, sizeof(double) * 3 * (size_t)cap);
  if (e != cudaSuccess) {
    devFree(ctx, nx); devFree(ctx, nn); devFree(ctx, nx2); devFree(ctx, nn2);
    return setErr(ctx, B200REG_E_OOM, "map: cannot reserve %lld points: %s", (long long)cap, cudaGetErrorString(e));
  }
  if (map->n > 0) {
    CUDA_TRY(ctx, cudaMemcpyAsync(nx, map->xyz, sizeof(double) * 3 * (size_t)map->n, cudaMemcpyDeviceToDevice, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(nn, map->nrm, sizeof(double) * 3 * (size_t)map->n, cudaMemcpyDeviceToDevice, ctx->stream));
  }
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  devFree(ctx, map->xyz); devFree(ctx, map->nrm); devFree(ctx, map->xyz2); devFree(ctx, map->nrm2);
  map->xyz = nx; map->nrm = nn; map->xyz2 = nx2; map->nrm2 = nn2;
  map->cap = cap;
  return B200REG_OK;
}

// voxelizeWithinCroppingVolume over `total` points of (xyz,nrm) -> (oxyz,onrm); returns the new size
int revoxelize(b200reg_ctx* ctx, double voxel, const b200reg_crop_spec& spec, const double* xyz, const double* nrm, int64_t total,
               double* oxyz, double* onrm, int64_t* newSize) {
  const size_t u4 = sizeof(unsigned) * (size_t)total, k8 = sizeof(u64) * (size_t)total;
  TRY(ensureBuf(ctx, ctx->flags, 6 * u4 + 64));
  unsigned* inFlag = (unsigned*)ctx->flags.p;
  unsigned* outFlag = inFlag + total;
  unsigned* inOff = outFlag + total;
  unsigned* outOff = inOff + total;
  unsigned* head = outOff + total;
  unsigned* headOff = head + total;
  TRY(ensureBuf(ctx, ctx->keys0, k8));
  TRY(ensureBuf(ctx, ctx->keys1, k8));
  TRY(ensureBuf(ctx, ctx->vals0, u4));
  TRY(ensureBuf(ctx, ctx->vals1, u4));
  TRY(ensureBuf(ctx, ctx->misc, k8));
  u64* keysAll = (u64*)ctx->misc.p;
  int* dErr = (int*)((char*)ctx->dstate + 32768 + 64);
  CUDA_TRY(ctx, cudaMemsetAsync(dErr, 0, sizeof(int), ctx->stream));
  classifyKernel<<<divUp(total, 256), 256, 0, ctx->stream>>>(xyz, total, spec, 1.0 / voxel, inFlag, outFlag, keysAll, dErr);
  LAUNCH_CHECK(ctx);
  TRY(scanFlags(ctx, inFlag, inOff, total));
  TRY(scanFlags(ctx, outFlag, outOff, total));
  int64_t nIn = 0, nOut = 0;
  TRY(scanTotal(ctx, inFlag, inOff, total, &nIn));
  nOut = total - nIn;
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pinned, dErr, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (*(const int*)ctx->pinned) return setErr(ctx, B200REG_E_INVALID_PARAM, "map: voxel index outside +-2^20 (voxel size too small for the map extent)");
  // pass-through points first, in input order (helpers.cpp:162-174)
  scatterSelectedKernel<<<divUp(total, 256), 256, 0, ctx->stream>>>(outFlag, outOff, total, xyz, nrm, oxyz, onrm);
  LAUNCH_CHECK(ctx);
  int64_t nVox = 0;
  if (nIn > 0) {
    const KeyPack kp = keyPackFor(spec, 1.0 / voxel);
    compactPairsKernel<<<divUp(total, 256), 256, 0, ctx->stream>>>(inFlag, inOff, keysAll, total, kp, (u64*)ctx->keys0.p, (unsigned*)ctx->vals0.p);
    LAUNCH_CHECK(ctx);
    cub::DoubleBuffer<u64> dk((u64*)ctx->keys0.p, (u64*)ctx->keys1.p);
    cub::DoubleBuffer<unsigned> dv((unsigned*)ctx->vals0.p, (unsigned*)ctx->vals1.p);
    size_t tmp = 0;
    CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp, dk, dv, (int)nIn, 0, kp.total, ctx->stream));
    TRY(ensureBuf(ctx, ctx->cubtmp, tmp));
    tmp = ctx->cubtmp.cap;
    CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(ctx->cubtmp.p, tmp, dk, dv, (int)nIn, 0, kp.total, ctx->stream));
    ctx->launches += (uint64_t)((kp.total + 7) / 8 + 1);
    headFlagKernel<<<divUp(nIn, 256), 256, 0, ctx->stream>>>(dk.Current(), nIn, head);
    LAUNCH_CHECK(ctx);
    TRY(scanFlags(ctx, head, headOff, nIn));
    voxelReduceKernel<<<divUp(nIn, 256), 256, 0, ctx->stream>>>(dk.Current(), dv.Current(), head, headOff, nIn, xyz, nrm, nOut, oxyz, onrm, 1);
    LAUNCH_CHECK(ctx);
    TRY(scanTotal(ctx, head, headOff, nIn, &nVox));
  }
  *newSize = nOut + nVox;
  return B200REG_OK;
}

int cloud64Reserve(b200reg_ctx* ctx, b200reg_cloud64* c, int64_t want) {
  if (want <= c->cap) return B200REG_OK;
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  devFree(ctx, c->xyz); devFree(ctx, c->nrm);
  c->xyz = c->nrm = nullptr;
  c->cap = 0;
  const int64_t cap = want + want / 4 + 1024;
  cudaError_t e = devAlloc(ctx, &c->xyz, sizeof(double) * 3 * (size_t)cap);
  if (e == cudaSuccess) e = devAlloc(ctx, &c->nrm, sizeof(double) * 3 * (size_t)cap);
  if (e != cudaSuccess) { devFree(ctx, c->xyz); c->xyz = nullptr; return setErr(ctx, B200REG_E_OOM, "cloud64: cannot reserve %lld points: %s", (long long)cap, cudaGetErrorString(e)); }
  c->cap = cap;
  return B200REG_OK;
}

// *out == nullptr: a new cloud; otherwise its storage is reused (grown when too small)
int cloud64Out(b200reg_ctx* ctx, b200reg_cloud64** out, int64_t want) {
  if (!*out) *out = new b200reg_cloud64();
  (*out)->n = 0;
  return cloud64Reserve(ctx, *out, want);
}

// CroppingVolume::crop (croppers.cpp:76-106) on device arrays: order-preserving compaction; synchronises for the count
int cropDevice(b200reg_ctx* ctx, const b200reg_crop_spec& spec, const double* dIn, const double* dInN, int64_t n, double* dOut, double* dOutN,
               int64_t* m) {
  TRY(ensureBuf(ctx, ctx->flags, 2 * sizeof(unsigned) * (size_t)n + 64));
  unsigned* flag = (unsigned*)ctx->flags.p;
  unsigned* off = flag + n;
  classifyKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(dIn, n, spec, 0.0, flag, nullptr, nullptr, nullptr);
  LAUNCH_CHECK(ctx);
  TRY(scanFlags(ctx, flag, off, n));
  scatterSelectedKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(flag, off, n, dIn, dInN, dOut, dOutN);
  LAUNCH_CHECK(ctx);
  return scanTotal(ctx, flag, off, n, m);
}

// Open3D v0.15.1 PointCloud::VoxelDownSample on device arrays (see b200reg_voxel_down_sample_f64); synchronises twice
int voxelDownDevice(b200reg_ctx* ctx, double voxel, const double* dIn, const double* dInN, int64_t n, double* dOut, double* dOutN, int64_t* nVox) {
  const size_t u4 = sizeof(unsigned) * (size_t)n, k8 = sizeof(u64) * (size_t)n;
  TRY(ensureBuf(ctx, ctx->flags, 2 * u4 + 64));
  unsigned* head = (unsigned*)ctx->flags.p;
  unsigned* headOff = head + n;
  TRY(ensureBuf(ctx, ctx->keys0, k8));
  TRY(ensureBuf(ctx, ctx->keys1, k8));
  TRY(ensureBuf(ctx, ctx->vals0, u4));
  TRY(ensureBuf(ctx, ctx->vals1, u4));
  const int mb = (int)fmin((double)(2 * ctx->sm_count), (double)divUp(n, 256));
  TRY(ensureBuf(ctx, ctx->partials, sizeof(double) * 6 * (size_t)mb));
  minBoundKernel<<<mb, 256, 0, ctx->stream>>>(dIn, n, (double*)ctx->partials.p);
  LAUNCH_CHECK(ctx);
  TRY(ensurePinned(ctx, sizeof(double) * 6 * (size_t)mb + 64));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pinned, ctx->partials.p, sizeof(double) * 6 * (size_t)mb, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  double mn[3] = {INFINITY, INFINITY, INFINITY}, mxb[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int b = 0; b < mb; ++b)
    for (int d = 0; d < 3; ++d) { mn[d] = fmin(mn[d], ((const double*)ctx->pinned)[6 * b + d]); mxb[d] = fmax(mxb[d], ((const double*)ctx->pinned)[6 * b + 3 + d]); }
  for (int d = 0; d < 3; ++d) mn[d] = mn[d] - voxel * 0.5;
  // bits per axis from the extent (one voxel of margin; a non-finite extent keeps the full 21): the sort runs over the bits in use
  int bits[3];
  for (int d = 0; d < 3; ++d) {
    const double span = floor((mxb[d] - mn[d]) / voxel) + 1.0;
    bits[d] = (span >= 0.0 && span < 2097151.0) ? bitsFor(span) : 21;
  }
  const int shy = bits[2], shx = bits[2] + bits[1], endBit = bits[0] + bits[1] + bits[2];
  int* dErr = (int*)((char*)ctx->dstate + 32768 + 64);
  CUDA_TRY(ctx, cudaMemsetAsync(dErr, 0, sizeof(int), ctx->stream));
  o3dKeyKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(dIn, n, mn[0], mn[1], mn[2], voxel, shx, shy, (u64*)ctx->keys0.p, (unsigned*)ctx->vals0.p, dErr);
  LAUNCH_CHECK(ctx);
  cub::DoubleBuffer<u64> dk((u64*)ctx->keys0.p, (u64*)ctx->keys1.p);
  cub::DoubleBuffer<unsigned> dv((unsigned*)ctx->vals0.p, (unsigned*)ctx->vals1.p);
  size_t tmp = 0;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp, dk, dv, (int)n, 0, endBit, ctx->stream));
  TRY(ensureBuf(ctx, ctx->cubtmp, tmp));
  tmp = ctx->cubtmp.cap;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(ctx->cubtmp.p, tmp, dk, dv, (int)n, 0, endBit, ctx->stream));
  ctx->launches += (uint64_t)((endBit + 7) / 8 + 1);
  headFlagKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(dk.Current(), n, head);
  LAUNCH_CHECK(ctx);
  TRY(scanFlags(ctx, head, headOff, n));
  voxelReduceKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(dk.Current(), dv.Current(), head, headOff, n, dIn, dInN, 0, dOut, dOutN, 0);
  LAUNCH_CHECK(ctx);
  CUDA_TRY(ctx, cudaMemcpyAsync((int*)ctx->pinned + 8, dErr, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  TRY(scanTotal(ctx, head, headOff, n, nVox));
  if (((const int*)ctx->pinned)[8]) return setErr(ctx, B200REG_E_INVALID_PARAM, "voxel_down_sample: voxel size too small for the cloud extent (index beyond 2^21)");
  return B200REG_OK;
}

// Submap::insertScan on a scan already in device memory (dIn / dInN may live in ctx->misc: it is reused only after the sync)
int mapInsertDevice(b200reg_ctx* ctx, b200reg_map* map, const double* dIn, const double* dInN, int64_t n, const double T_map_scan[16],
                    const b200reg_crop_spec& volume) {
  const int64_t total = map->n + n;
  if (total >= (int64_t)0x7fffffff) return setErr(ctx, B200REG_E_INVALID_PARAM, "map_insert: map too large");
  TRY(mapReserve(ctx, map, total));
  // `mapCloud_ += transformedCloud` (Submap.cpp:84): Open3D keeps normals only if both sides have them
  const bool normals = (map->n == 0 || map->hasNormals) && dInN != nullptr;
  double* dT = (double*)((char*)ctx->dstate + 40960);
  CUDA_TRY(ctx, cudaMemcpyAsync(dT, T_map_scan, sizeof(double) * 16, cudaMemcpyHostToDevice, ctx->stream));
  transformF64Kernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(dIn, normals ? dInN : nullptr, n, dT, map->xyz + 3 * (size_t)map->n,
                                                             normals ? map->nrm + 3 * (size_t)map->n : nullptr);
  LAUNCH_CHECK(ctx);
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));   // T_map_scan (host) and the caller's staging may be reused
  map->hasNormals = normals;
  if (!(map->voxel > 0.0)) {   // voxel_size <= 0: plain append (helpers.cpp:121-124)
    map->n = total;
    return B200REG_OK;
  }
  int64_t newSize = 0;
  TRY(revoxelize(ctx, map->voxel, volume, map->xyz, normals ? map->nrm : nullptr, total, map->xyz2, normals ? map->nrm2 : nullptr, &newSize));
  double* t = map->xyz; map->xyz = map->xyz2; map->xyz2 = t;
  t = map->nrm; map->nrm = map->nrm2; map->nrm2 = t;
  map->n = newSize;
  return B200REG_OK;
}

__global__ void fillKernel(unsigned* __restrict__ a, int64_t n, unsigned v) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = v;
}

// getIdxsOfCarvedPoints (helpers.cpp:245-281), one thread per scan ray.  The VoxelMap over the map points is the sorted
// (key, index) table; a sampled position looks its voxel up by binary search.  Consecutive samples that fall into the same
// voxel are looked up once (the reference inserts the same ids into a set again).  Removal = keep[id] <- 0 (idempotent).
__global__ void carveRayKernel(const double* __restrict__ scan, int64_t n, double sx, double sy, double sz, const u64* __restrict__ skeys,
                               const unsigned* __restrict__ svals, int64_t m, const double* __restrict__ nrm, double inv, double step,
                               double maxRay, double trunc, double minDot, unsigned* __restrict__ keep) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double d0 = scan[3 * i] - sx, d1 = scan[3 * i + 1] - sy, d2 = scan[3 * i + 2] - sz;
  const double length = sqrt((d0 * d0 + d1 * d1) + d2 * d2);
  const double dx = d0 / length, dy = d1 / length, dz = d2 / length;
  const double lim = fmin(length - trunc, maxRay);          // std::min(length - truncation, maxRaytracingLength)
  const double maxPath = (step < lim) ? lim : step;         // std::max(voxelSize, .)
  double distance = 0.0;
  u64 prev = ~0ull;
  while (distance < maxPath) {
    const double fx = floor((distance * dx + sx) * inv), fy = floor((distance * dy + sy) * inv), fz = floor((distance * dz + sz) * inv);
    const double kl = 1048575.0;
    if (fabs(fx) <= kl && fabs(fy) <= kl && fabs(fz) <= kl) {
      const u64 key = ((u64)((long long)fx + 1048576ll) << 42) | ((u64)((long long)fy + 1048576ll) << 21) | (u64)((long long)fz + 1048576ll);
      if (key != prev) {
        prev = key;
        int64_t lo = 0, hi = m;
        while (lo < hi) {
          const int64_t mid = (lo + hi) >> 1;
          if (skeys[mid] < key) lo = mid + 1; else hi = mid;
        }
        for (int64_t t = lo; t < m && skeys[t] == key; ++t) {
          const unsigned id = svals[t];
          bool rm = true;
          if (nrm) {
            double nx = nrm[3 * (size_t)id], ny = nrm[3 * (size_t)id + 1], nz = nrm[3 * (size_t)id + 2];
            const double n2 = (nx * nx + ny * ny) + nz * nz;
            if (n2 > 0.0) { const double nn = sqrt(n2); nx /= nn; ny /= nn; nz /= nn; }   // Eigen normalized()
            rm = fabs((dx * nx + dy * ny) + dz * nz) > minDot;
          }
          if (rm) keep[id] = 0u;
        }
      }
    }
    distance += step;
  }
}

// Submap::carve (Submap.cpp:130-144) on the resident map; dScan: the raw scan in the sensor frame, device memory
int carveDevice(b200reg_ctx* ctx, b200reg_map* map, const double* dScan, int64_t n, const double T_map_scan[16], const b200reg_crop_spec& volume,
                const b200reg_carve_params& cp, int64_t* removed) {
  *removed = 0;
  const int64_t N = map->n;
  if (N == 0 || n == 0) return B200REG_OK;
  if (!(cp.voxel_size > 0.0)) return setErr(ctx, B200REG_E_INVALID_PARAM, "carve: voxel_size must be > 0");
  const size_t u4 = sizeof(unsigned) * (size_t)N, k8 = sizeof(u64) * (size_t)N;
  TRY(ensureBuf(ctx, ctx->flags, 4 * u4 + 64));
  unsigned* inFlag = (unsigned*)ctx->flags.p;
  unsigned* inOff = inFlag + N;
  unsigned* keep = inOff + N;
  unsigned* keepOff = keep + N;
  TRY(ensureBuf(ctx, ctx->keys0, k8));
  TRY(ensureBuf(ctx, ctx->keys1, k8));
  TRY(ensureBuf(ctx, ctx->vals0, u4));
  TRY(ensureBuf(ctx, ctx->vals1, u4));
  TRY(ensureBuf(ctx, ctx->misc, k8 + sizeof(double) * 3 * (size_t)n + 64));
  u64* keysAll = (u64*)ctx->misc.p;
  double* dScanT = (double*)((char*)ctx->misc.p + ((k8 + 63) & ~(size_t)63));
  double* dT = (double*)((char*)ctx->dstate + 40960);
  int* dErr = (int*)((char*)ctx->dstate + 32768 + 64);
  CUDA_TRY(ctx, cudaMemcpyAsync(dT, T_map_scan, sizeof(double) * 16, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(dErr, 0, sizeof(int), ctx->stream));
  transformF64Kernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(dScan, nullptr, n, dT, dScanT, nullptr);   // o3d_slam::transform(mapToRangeSensor, rawScan)
  LAUNCH_CHECK(ctx);
  classifyKernel<<<divUp(N, 256), 256, 0, ctx->stream>>>(map->xyz, N, volume, 1.0 / cp.voxel_size, inFlag, nullptr, keysAll, dErr);   // getIndicesWithinVolume + getKey
  LAUNCH_CHECK(ctx);
  TRY(scanFlags(ctx, inFlag, inOff, N));
  int64_t nIn = 0;
  CUDA_TRY(ctx, cudaMemcpyAsync((int*)ctx->pinned + 8, dErr, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  TRY(scanTotal(ctx, inFlag, inOff, N, &nIn));   // synchronises (T_map_scan may be reused by the caller)
  if (((const int*)ctx->pinned)[8]) return setErr(ctx, B200REG_E_INVALID_PARAM, "carve: voxel index outside +-2^20 (voxel size too small for the map extent)");
  if (nIn == 0) return B200REG_OK;
  KeyPack kp63;   // the ray march looks voxels up by their 3 x 21-bit key: no re-packing here
  kp63.k0[0] = kp63.k0[1] = kp63.k0[2] = 0; kp63.shx = 42; kp63.shy = 21; kp63.total = 63;
  compactPairsKernel<<<divUp(N, 256), 256, 0, ctx->stream>>>(inFlag, inOff, keysAll, N, kp63, (u64*)ctx->keys0.p, (unsigned*)ctx->vals0.p);
  LAUNCH_CHECK(ctx);
  cub::DoubleBuffer<u64> dk((u64*)ctx->keys0.p, (u64*)ctx->keys1.p);
  cub::DoubleBuffer<unsigned> dv((unsigned*)ctx->vals0.p, (unsigned*)ctx->vals1.p);
  size_t tmp = 0;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp, dk, dv, (int)nIn, 0, 63, ctx->stream));
  TRY(ensureBuf(ctx, ctx->cubtmp, tmp));
  tmp = ctx->cubtmp.cap;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(ctx->cubtmp.p, tmp, dk, dv, (int)nIn, 0, 63, ctx->stream));
  ctx->launches += 9;
  fillKernel<<<divUp(N, 256), 256, 0, ctx->stream>>>(keep, N, 1u);
  LAUNCH_CHECK(ctx);
  carveRayKernel<<<divUp(n, 128), 128, 0, ctx->stream>>>(dScanT, n, T_map_scan[3], T_map_scan[7], T_map_scan[11], dk.Current(), dv.Current(), nIn,
                                                         map->hasNormals ? map->nrm : nullptr, 1.0 / cp.voxel_size, cp.voxel_size,
                                                         cp.max_raytracing_length, cp.truncation_distance, cp.min_dot_product_with_normal, keep);
  LAUNCH_CHECK(ctx);
  TRY(scanFlags(ctx, keep, keepOff, N));
  int64_t kept = 0;
  TRY(scanTotal(ctx, keep, keepOff, N, &kept));
  if (kept == N) return B200REG_OK;   // removeByIds: nothing to remove (helpers.cpp:230-233)
  scatterSelectedKernel<<<divUp(N, 256), 256, 0, ctx->stream>>>(keep, keepOff, N, map->xyz, map->hasNormals ? map->nrm : nullptr, map->xyz2,
                                                                 map->hasNormals ? map->nrm2 : nullptr);
  LAUNCH_CHECK(ctx);
  double* t = map->xyz; map->xyz = map->xyz2; map->xyz2 = t;
  t = map->nrm; map->nrm = map->nrm2; map->nrm2 = t;
  map->n = kept;
  *removed = N - kept;
  return B200REG_OK;
}

// open3dToPointmatcher (open3d_conversions.cpp:57-118) and back: fp64 SoA-of-3 <-> float4 clouds
__global__ void toF32Kernel(const double* __restrict__ xyz, const double* __restrict__ nrm, int64_t n, float4* __restrict__ oxyz, float4* __restrict__ onrm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  oxyz[i] = make_float4((float)xyz[3 * i], (float)xyz[3 * i + 1], (float)xyz[3 * i + 2], 1.f);
  if (nrm && onrm) onrm[i] = make_float4((float)nrm[3 * i], (float)nrm[3 * i + 1], (float)nrm[3 * i + 2], 0.f);
}
__global__ void normalsToF64Kernel(const float4* __restrict__ nrm, int64_t n, double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 v = nrm[i];
  out[3 * i] = (double)v.x; out[3 * i + 1] = (double)v.y; out[3 * i + 2] = (double)v.z;
}


// sensor_msgs/PointCloud2 bytes -> device clouds.  Both reference deserialisers read the fields through
// sensor_msgs::PointCloud2ConstIterator<float>: point i's value is the 4 bytes at data + i * point_step + field.offset taken as a
// host-order float, whatever datatype the field declares; row_step and is_bigendian are not consulted.
__device__ __forceinline__ float pc2Float(const unsigned char* __restrict__ p, int aligned) {
  if (aligned) return *reinterpret_cast<const float*>(p);
  const unsigned v = (unsigned)p[0] | ((unsigned)p[1] << 8) | ((unsigned)p[2] << 16) | ((unsigned)p[3] << 24);
  return __uint_as_float(v);
}
struct Pc2Dev {
  unsigned step;
  int ox, oy, oz, onx, ony, onz, aligned;
};
// RosPointCloud2Deserializer<float>::deserialize (RosPointCloud2Deserializer.cpp:9-38,213-232): features x y z pad = 1, "normals"
__global__ void pc2ToF32Kernel(const unsigned char* __restrict__ data, int64_t n, Pc2Dev L, float4* __restrict__ xyz, float4* __restrict__ nrm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned char* p = data + (size_t)i * L.step;
  xyz[i] = make_float4(pc2Float(p + L.ox, L.aligned), pc2Float(p + L.oy, L.aligned), pc2Float(p + L.oz, L.aligned), 1.f);
  if (nrm) nrm[i] = make_float4(pc2Float(p + L.onx, L.aligned), pc2Float(p + L.ony, L.aligned), pc2Float(p + L.onz, L.aligned), 0.f);
}
// open3d_conversions::rosToOpen3d (open3d_conversions.cpp:203-244): Vector3d(float x, float y, float z), normals likewise
__global__ void pc2ToF64Kernel(const unsigned char* __restrict__ data, int64_t n, Pc2Dev L, double* __restrict__ xyz, double* __restrict__ nrm) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned char* p = data + (size_t)i * L.step;
  xyz[3 * i] = (double)pc2Float(p + L.ox, L.aligned);
  xyz[3 * i + 1] = (double)pc2Float(p + L.oy, L.aligned);
  xyz[3 * i + 2] = (double)pc2Float(p + L.oz, L.aligned);
  if (nrm) {
    nrm[3 * i] = (double)pc2Float(p + L.onx, L.aligned);
    nrm[3 * i + 1] = (double)pc2Float(p + L.ony, L.aligned);
    nrm[3 * i + 2] = (double)pc2Float(p + L.onz, L.aligned);
  }
}

// validates the layout, uploads the message bytes into ctx->misc; *n = width * height, *hasN = all three normal fields present
int pc2Stage(b200reg_ctx* ctx, const uint8_t* data, size_t bytes, const b200reg_pc2_layout* L, Pc2Dev* dev, int64_t* n, bool* hasN) {
  *n = (int64_t)L->width * (int64_t)L->height;
  if (*n >= (int64_t)0x7fffffff) return setErr(ctx, B200REG_E_INVALID_PARAM, "pointcloud2: too many points");
  *hasN = L->off_normal_x >= 0 && L->off_normal_y >= 0 && L->off_normal_z >= 0;
  const int offs[6] = {L->off_x, L->off_y, L->off_z, L->off_normal_x, L->off_normal_y, L->off_normal_z};
  if (L->off_x < 0 || L->off_y < 0 || L->off_z < 0)   // the iterator throws "Field x does not exist" (point_cloud2_iterator.h)
    return setErr(ctx, B200REG_E_INVALID_FIELD, "pointcloud2: fields x, y, z are required (3-D clouds)");
  int aligned = (L->point_step % 4 == 0) ? 1 : 0;
  for (int k = 0; k < (*hasN ? 6 : 3); ++k) {
    if ((uint32_t)offs[k] + 4u > L->point_step) return setErr(ctx, B200REG_E_INVALID_PARAM, "pointcloud2: field offset %d beyond point_step %u", offs[k], L->point_step);
    if (offs[k] % 4 != 0) aligned = 0;
  }
  if (*n > 0 && (!data || bytes < (size_t)*n * L->point_step)) return setErr(ctx, B200REG_E_INVALID_PARAM, "pointcloud2: data holds %zu bytes, width * height * point_step = %zu", bytes, (size_t)*n * L->point_step);
  *dev = Pc2Dev{L->point_step, offs[0], offs[1], offs[2], offs[3], offs[4], offs[5], aligned};
  if (*n == 0) return B200REG_OK;
  const size_t need = (size_t)*n * L->point_step;
  TRY(ensureBuf(ctx, ctx->misc, need + 64));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->misc.p, data, need, cudaMemcpyHostToDevice, ctx->stream));
  return B200REG_OK;
}

}  // namespace

extern "C" {

int32_t b200reg_map_create(b200reg_ctx* ctx, double voxel_size, b200reg_map** out) {
  if (!ctx || !out) return setErr(ctx, B200REG_E_INVALID_PARAM, "map_create: bad arguments");
  b200reg_map* m = new b200reg_map();
  m->voxel = voxel_size;
  *out = m;
  return B200REG_OK;
}

void b200reg_map_destroy(b200reg_ctx* ctx, b200reg_map* map) {
  if (!map) return;
  if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
  devFree(ctx, map->xyz); devFree(ctx, map->nrm); devFree(ctx, map->xyz2); devFree(ctx, map->nrm2);
  delete map;
}

int64_t b200reg_map_size(const b200reg_map* map) { return map ? map->n : 0; }
int32_t b200reg_map_has_normals(const b200reg_map* map) { return (map && map->hasNormals) ? 1 : 0; }

int32_t b200reg_map_insert(b200reg_ctx* ctx, b200reg_map* map, const double* scan_xyz, const double* scan_nrm, int64_t n,
                           const double T_map_scan[16], const b200reg_crop_spec* volume) {
  if (!ctx || !map || n < 0 || (n > 0 && !scan_xyz) || !T_map_scan || !volume) return setErr(ctx, B200REG_E_INVALID_PARAM, "map_insert: bad arguments");
  if (n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const size_t fb = sizeof(double) * 3 * (size_t)n;
  TRY(ensureBuf(ctx, ctx->misc, 2 * fb + 256));
  double* dIn = (double*)ctx->misc.p;
  double* dInN = scan_nrm ? dIn + 3 * (size_t)n : nullptr;
  CUDA_TRY(ctx, cudaMemcpyAsync(dIn, scan_xyz, fb, cudaMemcpyHostToDevice, ctx->stream));
  if (scan_nrm) CUDA_TRY(ctx, cudaMemcpyAsync(dInN, scan_nrm, fb, cudaMemcpyHostToDevice, ctx->stream));
  // (the core synchronises after the transform, before misc is reused by the re-voxelisation)
  return mapInsertDevice(ctx, map, dIn, dInN, n, T_map_scan, *volume);
}

int32_t b200reg_map_insert_cloud64(b200reg_ctx* ctx, b200reg_map* map, const b200reg_cloud64* scan, const double T_map_scan[16],
                                   const b200reg_crop_spec* volume) {
  if (!ctx || !map || !scan || !T_map_scan || !volume) return setErr(ctx, B200REG_E_INVALID_PARAM, "map_insert_cloud64: bad arguments");
  if (scan->n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  return mapInsertDevice(ctx, map, scan->xyz, scan->hasNormals ? scan->nrm : nullptr, scan->n, T_map_scan, *volume);
}

int32_t b200reg_map_carve(b200reg_ctx* ctx, b200reg_map* map, const double* scan_xyz, int64_t n, const double T_map_scan[16],
                          const b200reg_crop_spec* volume, const b200reg_carve_params* params, int64_t* removed) {
  if (!ctx || !map || n < 0 || (n > 0 && !scan_xyz) || !T_map_scan || !volume || !params) return setErr(ctx, B200REG_E_INVALID_PARAM, "map_carve: bad arguments");
  int64_t r = 0;
  if (removed) *removed = 0;
  if (n == 0 || map->n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  double* dIn = nullptr;
  CUDA_TRY(ctx, devAlloc(ctx, &dIn, sizeof(double) * 3 * (size_t)n));
  cudaError_t e = cudaMemcpyAsync(dIn, scan_xyz, sizeof(double) * 3 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream);
  int code = (e == cudaSuccess) ? carveDevice(ctx, map, dIn, n, T_map_scan, *volume, *params, &r) : setErr(ctx, B200REG_E_CUDA, "map_carve: %s", cudaGetErrorString(e));
  cudaStreamSynchronize(ctx->stream);   // the host scan may be reused
  devFree(ctx, dIn);
  if (removed) *removed = r;
  return code;
}

int32_t b200reg_map_carve_cloud64(b200reg_ctx* ctx, b200reg_map* map, const b200reg_cloud64* scan, const double T_map_scan[16],
                                  const b200reg_crop_spec* volume, const b200reg_carve_params* params, int64_t* removed) {
  if (!ctx || !map || !scan || !T_map_scan || !volume || !params) return setErr(ctx, B200REG_E_INVALID_PARAM, "map_carve_cloud64: bad arguments");
  int64_t r = 0;
  if (removed) *removed = 0;
  if (scan->n == 0 || map->n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const int code = carveDevice(ctx, map, scan->xyz, scan->n, T_map_scan, *volume, *params, &r);
  if (removed) *removed = r;
  return code;
}

int32_t b200reg_map_crop(b200reg_ctx* ctx, const b200reg_map* map, const b200reg_crop_spec* volume, b200reg_cloud** out) {
  if (!ctx || !map || !volume || !out) return setErr(ctx, B200REG_E_INVALID_PARAM, "map_crop: bad arguments");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const int64_t n = map->n;
  int64_t m = 0;
  unsigned* flag = nullptr; unsigned* off = nullptr;
  if (n > 0) {
    TRY(ensureBuf(ctx, ctx->flags, 2 * sizeof(unsigned) * (size_t)n + 64));
    flag = (unsigned*)ctx->flags.p;
    off = flag + n;
    classifyKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(map->xyz, n, *volume, 0.0, flag, nullptr, nullptr, nullptr);
    LAUNCH_CHECK(ctx);
    TRY(scanFlags(ctx, flag, off, n));
    TRY(scanTotal(ctx, flag, off, n, &m));
  }
  b200reg_cloud* c = new b200reg_cloud();
  c->n = m;
  c->cap = m > 0 ? m : 1;
  cudaError_t e = devAlloc(ctx, &c->xyz, sizeof(float4) * (size_t)c->cap);
  if (e == cudaSuccess && map->hasNormals) e = devAlloc(ctx, &c->nrm, sizeof(float4) * (size_t)c->cap);
  if (e != cudaSuccess) { devFree(ctx, c->xyz); delete c; return setErr(ctx, B200REG_E_OOM, "map_crop: %s", cudaGetErrorString(e)); }
  if (m > 0) {
    scatterSelectedF32Kernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(flag, off, n, map->xyz, map->hasNormals ? map->nrm : nullptr, c->xyz, c->nrm);
    LAUNCH_CHECK(ctx);
  }
  *out = c;
  return B200REG_OK;
}

int32_t b200reg_map_download(b200reg_ctx* ctx, const b200reg_map* map, double* xyz, double* nrm) {
  if (!ctx || !map) return setErr(ctx, B200REG_E_INVALID_PARAM, "map_download: bad arguments");
  if (map->n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (xyz) CUDA_TRY(ctx, cudaMemcpyAsync(xyz, map->xyz, sizeof(double) * 3 * (size_t)map->n, cudaMemcpyDeviceToHost, ctx->stream));
  if (nrm) {
    if (!map->hasNormals) return setErr(ctx, B200REG_E_INVALID_FIELD, "map has no normals");
    CUDA_TRY(ctx, cudaMemcpyAsync(nrm, map->nrm, sizeof(double) * 3 * (size_t)map->n, cudaMemcpyDeviceToHost, ctx->stream));
  }
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return B200REG_OK;
}

int32_t b200reg_crop_f64(b200reg_ctx* ctx, const b200reg_crop_spec* volume, const double* xyz, const double* nrm, int64_t n,
                         double* xyz_out, double* nrm_out, int64_t* n_out) {
  if (!ctx || !volume || n < 0 || (n > 0 && !xyz) || !n_out) return setErr(ctx, B200REG_E_INVALID_PARAM, "crop_f64: bad arguments");
  *n_out = 0;
  if (n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const size_t fb = sizeof(double) * 3 * (size_t)n;
  TRY(ensureBuf(ctx, ctx->misc, 4 * fb + 64));
  double* dIn = (double*)ctx->misc.p;
  double* dInN = dIn + 3 * (size_t)n;
  double* dOut = dInN + 3 * (size_t)n;
  double* dOutN = dOut + 3 * (size_t)n;
  CUDA_TRY(ctx, cudaMemcpyAsync(dIn, xyz, fb, cudaMemcpyHostToDevice, ctx->stream));
  if (nrm) CUDA_TRY(ctx, cudaMemcpyAsync(dInN, nrm, fb, cudaMemcpyHostToDevice, ctx->stream));
  int64_t m = 0;
  TRY(cropDevice(ctx, *volume, dIn, nrm ? dInN : nullptr, n, dOut, dOutN, &m));
  if (m > 0 && xyz_out) CUDA_TRY(ctx, cudaMemcpyAsync(xyz_out, dOut, sizeof(double) * 3 * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
  if (m > 0 && nrm && nrm_out) CUDA_TRY(ctx, cudaMemcpyAsync(nrm_out, dOutN, sizeof(double) * 3 * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  *n_out = m;
  return B200REG_OK;
}

// o3d_slam::voxelize (helpers.cpp:108-115) -> Open3D v0.15.1 PointCloud::VoxelDownSample on a HOST fp64 cloud: mean position
// and mean normal (not re-normalised) per voxel of the grid anchored at min_bound - voxel/2; output in ascending (x,y,z) key
// order (Open3D's is hash order — unpinned).  xyz_out / nrm_out hold n points; *n_out = voxels.
int32_t b200reg_voxel_down_sample_f64(b200reg_ctx* ctx, double voxel, const double* xyz, const double* nrm, int64_t n, double* xyz_out,
                                      double* nrm_out, int64_t* n_out) {
  if (!ctx || n < 0 || (n > 0 && (!xyz || !xyz_out)) || !n_out) return setErr(ctx, B200REG_E_INVALID_PARAM, "voxel_down_sample: bad arguments");
  if (!(voxel > 0.0)) return setErr(ctx, B200REG_E_INVALID_PARAM, "voxel_down_sample: voxel_size <= 0");   // Open3D throws too
  *n_out = 0;
  if (n == 0) return B200REG_OK;
  if (n >= (int64_t)0x7fffffff) return setErr(ctx, B200REG_E_INVALID_PARAM, "voxel_down_sample: too many points");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const size_t fb = sizeof(double) * 3 * (size_t)n;
  TRY(ensureBuf(ctx, ctx->misc, 4 * fb + 64));
  double* dIn = (double*)ctx->misc.p;
  double* dInN = dIn + 3 * (size_t)n;
  double* dOut = dInN + 3 * (size_t)n;
  double* dOutN = dOut + 3 * (size_t)n;
  CUDA_TRY(ctx, cudaMemcpyAsync(dIn, xyz, fb, cudaMemcpyHostToDevice, ctx->stream));
  if (nrm) CUDA_TRY(ctx, cudaMemcpyAsync(dInN, nrm, fb, cudaMemcpyHostToDevice, ctx->stream));
  int64_t nVox = 0;
  TRY(voxelDownDevice(ctx, voxel, dIn, nrm ? dInN : nullptr, n, dOut, nrm ? dOutN : nullptr, &nVox));
  CUDA_TRY(ctx, cudaMemcpyAsync(xyz_out, dOut, sizeof(double) * 3 * (size_t)nVox, cudaMemcpyDeviceToHost, ctx->stream));
  if (nrm && nrm_out) CUDA_TRY(ctx, cudaMemcpyAsync(nrm_out, dOutN, sizeof(double) * 3 * (size_t)nVox, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  *n_out = nVox;
  return B200REG_OK;
}

// ---- the same steps on device-resident fp64 clouds (no host round trip between them) ------------------------------------
int32_t b200reg_cloud64_upload(b200reg_ctx* ctx, const double* xyz, const double* nrm, int64_t n, b200reg_cloud64** out) {
  if (!ctx || !out || n < 0 || (n > 0 && !xyz)) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud64_upload: bad arguments");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  TRY(cloud64Out(ctx, out, n));
  b200reg_cloud64* c = *out;
  if (n > 0) {
    CUDA_TRY(ctx, cudaMemcpyAsync(c->xyz, xyz, sizeof(double) * 3 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    if (nrm) CUDA_TRY(ctx, cudaMemcpyAsync(c->nrm, nrm, sizeof(double) * 3 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));   // pageable host memory may be reused by the caller
  }
  c->n = n;
  c->hasNormals = nrm != nullptr;
  return B200REG_OK;
}

int32_t b200reg_cloud64_download(b200reg_ctx* ctx, const b200reg_cloud64* c, double* xyz, double* nrm) {
  if (!ctx || !c) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud64_download: bad arguments");
  if (nrm && !c->hasNormals) return setErr(ctx, B200REG_E_INVALID_FIELD, "cloud64 has no normals");
  if (c->n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (xyz) CUDA_TRY(ctx, cudaMemcpyAsync(xyz, c->xyz, sizeof(double) * 3 * (size_t)c->n, cudaMemcpyDeviceToHost, ctx->stream));
  if (nrm) CUDA_TRY(ctx, cudaMemcpyAsync(nrm, c->nrm, sizeof(double) * 3 * (size_t)c->n, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return B200REG_OK;
}

// Next-scan prefetch: the same upload on the ctx's COPY stream, so that it overlaps the kernels of the scan being processed
// (a pinned host buffer makes the copy truly asynchronous).  The copy is ordered behind everything already enqueued on the
// compute stream (which may still read *out's storage); b200reg_cloud64_acquire waits for it before the cloud is used.
int32_t b200reg_cloud64_prefetch(b200reg_ctx* ctx, const double* xyz, const double* nrm, int64_t n, b200reg_cloud64** out) {
  if (!ctx || !out || n < 0 || (n > 0 && !xyz)) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud64_prefetch: bad arguments");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (!ctx->copyStream) {
    CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->copyStream, cudaStreamNonBlocking));
    CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->copyFence, cudaEventDisableTiming));
  }
  if (*out && (*out)->prefetched) TRY(b200reg_cloud64_acquire(ctx, *out));
  TRY(cloud64Out(ctx, out, n));
  b200reg_cloud64* c = *out;
  if (!c->ready) CUDA_TRY(ctx, cudaEventCreateWithFlags(&c->ready, cudaEventDisableTiming));
  CUDA_TRY(ctx, cudaEventRecord(ctx->copyFence, ctx->stream));
  CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->copyStream, ctx->copyFence, 0));
  if (n > 0) {
    CUDA_TRY(ctx, cudaMemcpyAsync(c->xyz, xyz, sizeof(double) * 3 * (size_t)n, cudaMemcpyHostToDevice, ctx->copyStream));
    if (nrm) CUDA_TRY(ctx, cudaMemcpyAsync(c->nrm, nrm, sizeof(double) * 3 * (size_t)n, cudaMemcpyHostToDevice, ctx->copyStream));
  }
  CUDA_TRY(ctx, cudaEventRecord(c->ready, ctx->copyStream));
  c->n = n;
  c->hasNormals = nrm != nullptr;
  c->prefetched = true;
  return B200REG_OK;
}

// The host waits for the prefetch's copy (normally long finished) and the compute stream is ordered behind it: afterwards the
// cloud behaves like an uploaded one and the host buffer may be reused.
int32_t b200reg_cloud64_acquire(b200reg_ctx* ctx, b200reg_cloud64* c) {
  if (!ctx || !c) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud64_acquire: bad arguments");
  if (!c->prefetched) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaEventSynchronize(c->ready));
  CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, c->ready, 0));
  c->prefetched = false;
  return B200REG_OK;
}

int64_t b200reg_cloud64_size(const b200reg_cloud64* c) { return c ? c->n : 0; }
int32_t b200reg_cloud64_has_normals(const b200reg_cloud64* c) { return (c && c->hasNormals) ? 1 : 0; }

void b200reg_cloud64_free(b200reg_ctx* ctx, b200reg_cloud64* c) {
  if (!c) return;
  if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
  if (c->ready) { cudaEventSynchronize(c->ready); cudaEventDestroy(c->ready); }
  devFree(ctx, c->xyz); devFree(ctx, c->nrm);
  delete c;
}

int32_t b200reg_cloud64_crop(b200reg_ctx* ctx, const b200reg_crop_spec* volume, const b200reg_cloud64* in, b200reg_cloud64** out) {
  if (!ctx || !volume || !in || !out || *out == in) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud64_crop: bad arguments");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  TRY(cloud64Out(ctx, out, in->n));
  b200reg_cloud64* c = *out;
  c->hasNormals = in->hasNormals;
  if (in->n == 0) return B200REG_OK;
  int64_t m = 0;
  TRY(cropDevice(ctx, *volume, in->xyz, in->hasNormals ? in->nrm : nullptr, in->n, c->xyz, c->nrm, &m));
  c->n = m;
  return B200REG_OK;
}

int32_t b200reg_cloud64_voxel_down_sample(b200reg_ctx* ctx, double voxel, const b200reg_cloud64* in, b200reg_cloud64** out) {
  if (!ctx || !in || !out || *out == in) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud64_voxel_down_sample: bad arguments");
  if (!(voxel > 0.0)) return setErr(ctx, B200REG_E_INVALID_PARAM, "voxel_down_sample: voxel_size <= 0");
  if (in->n >= (int64_t)0x7fffffff) return setErr(ctx, B200REG_E_INVALID_PARAM, "voxel_down_sample: too many points");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  TRY(cloud64Out(ctx, out, in->n));
  b200reg_cloud64* c = *out;
  c->hasNormals = in->hasNormals;
  if (in->n == 0) return B200REG_OK;
  int64_t nVox = 0;
  TRY(voxelDownDevice(ctx, voxel, in->xyz, in->hasNormals ? in->nrm : nullptr, in->n, c->xyz, in->hasNormals ? c->nrm : nullptr, &nVox));
  c->n = nVox;
  return B200REG_OK;
}

int32_t b200reg_cloud64_to_f32(b200reg_ctx* ctx, const b200reg_cloud64* in, b200reg_cloud** out) {
  if (!ctx || !in || !out) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud64_to_f32: bad arguments");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const bool fresh = (*out == nullptr);
  b200reg_cloud* c = fresh ? new b200reg_cloud() : *out;
  if (c->cap < in->n || c->cap == 0 || (in->hasNormals && !c->nrm)) {
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    devFree(ctx, c->xyz); devFree(ctx, c->nrm);
    c->xyz = c->nrm = nullptr;
    c->n = 0;
    c->cap = in->n + in->n / 4 + 1024;
    cudaError_t e = devAlloc(ctx, &c->xyz, sizeof(float4) * (size_t)c->cap);
    if (e == cudaSuccess) e = devAlloc(ctx, &c->nrm, sizeof(float4) * (size_t)c->cap);
    if (e != cudaSuccess) {
      devFree(ctx, c->xyz);
      c->xyz = nullptr;
      c->cap = 0;
      if (fresh) delete c;
      return setErr(ctx, B200REG_E_OOM, "cloud64_to_f32: %s", cudaGetErrorString(e));
    }
  }
  if (!in->hasNormals && c->nrm) {   // DataPoints without a normals descriptor
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    devFree(ctx, c->nrm);
    c->nrm = nullptr;
  }
  c->n = in->n;
  *out = c;
  if (in->n > 0) {
    toF32Kernel<<<divUp(in->n, 256), 256, 0, ctx->stream>>>(in->xyz, in->hasNormals ? in->nrm : nullptr, in->n, c->xyz, c->nrm);
    LAUNCH_CHECK(ctx);
  }
  return B200REG_OK;
}

// estimateNormalsOrCovariancesIfNeeded (CloudRegistration.cpp:62-75): Open3D's EstimateNormals + orientation toward `origin`, in fp64.
// The fp32 copy of the points only feeds the index that proposes neighbour candidates; distances, the neighbour selection, the
// covariance and the eigenvector are fp64 on the cloud's own coordinates (normals.cu estimateNormals64).
int32_t b200reg_cloud64_estimate_normals(b200reg_ctx* ctx, b200reg_cloud64* inout, int32_t knn, double radius, const double origin[3],
                                         int32_t* degenerate_count) {
  if (!ctx || !inout) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud64_estimate_normals: bad arguments");
  if (degenerate_count) *degenerate_count = 0;
  if (inout->n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  b200reg_cloud* tmp = nullptr;
  inout->hasNormals = false;
  TRY(b200reg_cloud64_to_f32(ctx, inout, &tmp));
  const int code = estimateNormals64(ctx, inout, tmp, knn, radius, origin, degenerate_count);
  if (code == B200REG_OK) inout->hasNormals = true;
  b200reg_cloud_free(ctx, tmp);
  return code;
}
int32_t b200reg_cloud_from_pointcloud2(b200reg_ctx* ctx, const uint8_t* data, size_t data_bytes, const b200reg_pc2_layout* layout,
                                       b200reg_cloud** out) {
  if (!ctx || !layout || !out) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud_from_pointcloud2: bad arguments");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Pc2Dev dev;
  int64_t n = 0;
  bool hasN = false;
  TRY(pc2Stage(ctx, data, data_bytes, layout, &dev, &n, &hasN));
  const bool fresh = (*out == nullptr);
  b200reg_cloud* c = fresh ? new b200reg_cloud() : *out;
  if (c->cap < n || c->cap == 0 || (hasN && !c->nrm)) {
    devFree(ctx, c->xyz); devFree(ctx, c->nrm);
    c->xyz = c->nrm = nullptr;
    c->n = 0;
    c->cap = n + n / 4 + 1024;
    cudaError_t e = devAlloc(ctx, &c->xyz, sizeof(float4) * (size_t)c->cap);
    if (e == cudaSuccess && hasN) e = devAlloc(ctx, &c->nrm, sizeof(float4) * (size_t)c->cap);
    if (e != cudaSuccess) {
      devFree(ctx, c->xyz);
      c->xyz = nullptr;
      c->cap = 0;
      if (fresh) delete c;
      return setErr(ctx, B200REG_E_OOM, "cloud_from_pointcloud2: %s", cudaGetErrorString(e));
    }
  }
  if (!hasN && c->nrm) { devFree(ctx, c->nrm); c->nrm = nullptr; }
  c->n = n;
  *out = c;
  if (n > 0) {
    pc2ToF32Kernel<<<divUp(n, 256), 256, 0, ctx->stream>>>((const unsigned char*)ctx->misc.p, n, dev, c->xyz, hasN ? c->nrm : nullptr);
    LAUNCH_CHECK(ctx);
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));   // the message buffer may be reused by the caller
  }
  return B200REG_OK;
}

int32_t b200reg_cloud64_from_pointcloud2(b200reg_ctx* ctx, const uint8_t* data, size_t data_bytes, const b200reg_pc2_layout* layout,
                                         b200reg_cloud64** out) {
  if (!ctx || !layout || !out) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud64_from_pointcloud2: bad arguments");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Pc2Dev dev;
  int64_t n = 0;
  bool hasN = false;
  TRY(pc2Stage(ctx, data, data_bytes, layout, &dev, &n, &hasN));
  TRY(cloud64Out(ctx, out, n));
  b200reg_cloud64* c = *out;
  c->hasNormals = hasN;
  c->n = n;
  if (n > 0) {
    pc2ToF64Kernel<<<divUp(n, 256), 256, 0, ctx->stream>>>((const unsigned char*)ctx->misc.p, n, dev, c->xyz, hasN ? c->nrm : nullptr);
    LAUNCH_CHECK(ctx);
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return B200REG_OK;
}

}  // extern "C"
