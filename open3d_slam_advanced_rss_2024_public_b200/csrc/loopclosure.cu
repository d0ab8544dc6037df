// csrc/loopclosure.cu — support operations of the loop-closure refinement around the batched registration,
// paths relative to /root/reference/open3d_slam_rsl/open3d_slam/open3d_slam/:
//   computeIndicesOfOverlappingPoints            src/helpers.cpp:320-345   (call sites PlaceRecognition.cpp:100-106)
//   GetInformationMatrixFromPointClouds          Open3D v0.15.1 pipelines/registration/Registration.cpp (external, pinned by
//                                                open3d_catkin/CMakeLists.txt:106-107); call sites PlaceRecognition.cpp:148-149,
//                                                constraint_builders.cpp:71-72
// Both are "voxel join / nearest neighbour + small reduction": the kernels of the scan-to-map path with different outputs.
// Compiled with --fmad=false (every fp64 operation individually rounded, like the oracle).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cmath>

#include "common.cuh"

int runKnn(b200reg_ctx* ctx, const b200reg_index* ix, const float4* queries, int64_t n, const float* T, int k, float maxDist, int allowSelf,
           int seeded);

namespace {

typedef unsigned long long u64;

// key = int(floor(p * (1/voxel))) per axis (VoxelHashMap.hpp:48-51,127), 21 bits per axis; the source cloud is moved by
// PointCloud::Transform first: (T [p;1]).head3 / w in column order ((T0 x + T1 y) + T2 z) + T3
__global__ void overlapKeyKernel(const double* __restrict__ xyz, int64_t n, const double* __restrict__ Tm, double inv, u64* __restrict__ keys,
                                 int* __restrict__ err) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
  if (Tm) {
    double q[4];
    for (int r = 0; r < 4; ++r) q[r] = ((Tm[4 * r] * x + Tm[4 * r + 1] * y) + Tm[4 * r + 2] * z) + Tm[4 * r + 3] * 1.0;
    x = q[0] / q[3]; y = q[1] / q[3]; z = q[2] / q[3];
  }
  const double fx = floor(x * inv), fy = floor(y * inv), fz = floor(z * inv);
  const double lim = 1048575.0;
  u64 key = ~0ull;
  if (!(fabs(fx) <= lim) || !(fabs(fy) <= lim) || !(fabs(fz) <= lim)) *err = 1;
  else key = ((u64)((long long)fx + 1048576ll) << 42) | ((u64)((long long)fy + 1048576ll) << 21) | (u64)((long long)fz + 1048576ll);
  keys[i] = key;
}

__device__ __forceinline__ int64_t lowerBound(const u64* __restrict__ a, int64_t m, u64 key) {
  int64_t lo = 0, hi = m;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (a[mid] < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ int64_t countOf(const u64* __restrict__ a, int64_t m, u64 key) {
  const int64_t lo = lowerBound(a, m, key);
  if (lo >= m || a[lo] != key) return 0;
  return lowerBound(a, m, key + 1ull) - lo;
}

// a point belongs to the overlap iff its voxel holds >= minPts source points AND >= minPts target points (helpers.cpp:339-343)
__global__ void overlapFlagKernel(const u64* __restrict__ keys, int64_t n, const u64* __restrict__ sortedOwn, int64_t nOwn,
                                  const u64* __restrict__ sortedOther, int64_t nOther, int64_t minPts, unsigned* __restrict__ flag) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const u64 key = keys[i];
  flag[i] = (countOf(sortedOther, nOther, key) >= minPts && countOf(sortedOwn, nOwn, key) >= minPts) ? 1u : 0u;
}

__global__ void flaggedIndicesKernel(const unsigned* __restrict__ flag, const unsigned* __restrict__ off, int64_t n, long long* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && flag[i]) out[off[i]] = (long long)i;
}

int sortKeys(b200reg_ctx* ctx, const u64* in, u64* out, int64_t n) {
  size_t tmp = 0;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortKeys(nullptr, tmp, in, out, (int)n, 0, 64, ctx->stream));
  TRY(ensureBuf(ctx, ctx->cubtmp, tmp));
  tmp = ctx->cubtmp.cap;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortKeys(ctx->cubtmp.p, tmp, in, out, (int)n, 0, 64, ctx->stream));
  ctx->launches += 9;
  return B200REG_OK;
}

int exclusiveScan(b200reg_ctx* ctx, const unsigned* flags, unsigned* offs, int64_t n) {
  size_t tmp = 0;
  CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp, flags, offs, (int)n, ctx->stream));
  TRY(ensureBuf(ctx, ctx->cubtmp, tmp));
  tmp = ctx->cubtmp.cap;
  CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(ctx->cubtmp.p, tmp, flags, offs, (int)n, ctx->stream));
  ctx->launches += 2;
  return B200REG_OK;
}

}  // namespace

extern "C" int32_t b200reg_overlap_indices(b200reg_ctx* ctx, const b200reg_cloud64* source, const b200reg_cloud64* target,
                                           const double T_source_to_target[16], double voxel_size, int64_t min_points_per_voxel,
                                           int64_t* idx_source_out, int64_t* n_source_out, int64_t* idx_target_out, int64_t* n_target_out) {
  if (!ctx || !source || !target || !T_source_to_target || !n_source_out || !n_target_out)
    return setErr(ctx, B200REG_E_INVALID_PARAM, "overlap_indices: bad arguments");
  *n_source_out = *n_target_out = 0;
  if (!(voxel_size > 0.0)) return setErr(ctx, B200REG_E_INVALID_PARAM, "overlap_indices: voxel_size must be > 0");
  if (min_points_per_voxel < 1) return setErr(ctx, B200REG_E_INVALID_PARAM, "overlap_indices: minNumPointsPerVoxel must be >= 1");   // assert_ge (helpers.cpp:323)
  const int64_t nS = source->n, nT = target->n;
  if (nS == 0 || nT == 0) return B200REG_OK;
  if (nS >= (int64_t)0x7fffffff || nT >= (int64_t)0x7fffffff) return setErr(ctx, B200REG_E_INVALID_PARAM, "overlap_indices: too many points");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const int64_t nMax = nS > nT ? nS : nT;
  u64 *kS = nullptr, *kT = nullptr, *sS = nullptr, *sT = nullptr;
  unsigned *flag = nullptr, *off = nullptr;
  long long* dIdx = nullptr;
  cudaError_t e = devAlloc(ctx, &kS, sizeof(u64) * (size_t)nS);
  if (e == cudaSuccess) e = devAlloc(ctx, &sS, sizeof(u64) * (size_t)nS);
  if (e == cudaSuccess) e = devAlloc(ctx, &kT, sizeof(u64) * (size_t)nT);
  if (e == cudaSuccess) e = devAlloc(ctx, &sT, sizeof(u64) * (size_t)nT);
  if (e == cudaSuccess) e = devAlloc(ctx, &flag, sizeof(unsigned) * (size_t)nMax);
  if (e == cudaSuccess) e = devAlloc(ctx, &off, sizeof(unsigned) * (size_t)nMax);
  if (e == cudaSuccess) e = devAlloc(ctx, &dIdx, sizeof(long long) * (size_t)nMax);
  int code = B200REG_OK;
  auto cleanup = [&]() {
    devFree(ctx, kS); devFree(ctx, sS); devFree(ctx, kT); devFree(ctx, sT); devFree(ctx, flag); devFree(ctx, off); devFree(ctx, dIdx);
  };
  if (e != cudaSuccess) { cleanup(); return setErr(ctx, B200REG_E_OOM, "overlap_indices: %s", cudaGetErrorString(e)); }
  double* dT = (double*)((char*)ctx->dstate + 40960);
  int* dErr = (int*)((char*)ctx->dstate + 32768 + 64);
  cudaMemcpyAsync(dT, T_source_to_target, sizeof(double) * 16, cudaMemcpyHostToDevice, ctx->stream);
  cudaMemsetAsync(dErr, 0, sizeof(int), ctx->stream);
  const double inv = 1.0 / voxel_size;
  overlapKeyKernel<<<divUp(nS, 256), 256, 0, ctx->stream>>>(source->xyz, nS, dT, inv, kS, dErr);
  overlapKeyKernel<<<divUp(nT, 256), 256, 0, ctx->stream>>>(target->xyz, nT, nullptr, inv, kT, dErr);
  ctx->launches += 2;
  code = sortKeys(ctx, kS, sS, nS);
  if (code == B200REG_OK) code = sortKeys(ctx, kT, sT, nT);
  unsigned* h = (unsigned*)ctx->pinned;
  for (int side = 0; side < 2 && code == B200REG_OK; ++side) {
    const int64_t n = side == 0 ? nS : nT;
    overlapFlagKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(side == 0 ? kS : kT, n, side == 0 ? sS : sT, side == 0 ? nS : nT,
                                                              side == 0 ? sT : sS, side == 0 ? nT : nS, min_points_per_voxel, flag);
    ctx->launches++;
    code = exclusiveScan(ctx, flag, off, n);
    if (code != B200REG_OK) break;
    flaggedIndicesKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(flag, off, n, dIdx);
    ctx->launches++;
    cudaMemcpyAsync(h, off + (n - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream);
    cudaMemcpyAsync(h + 1, flag + (n - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream);
    cudaMemcpyAsync(h + 2, dErr, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) { code = setErr(ctx, B200REG_E_CUDA, "overlap_indices: %s", cudaGetErrorString(cudaGetLastError())); break; }
    if (h[2]) { code = setErr(ctx, B200REG_E_INVALID_PARAM, "overlap_indices: voxel index outside +-2^20 (voxel size too small for the cloud extent)"); break; }
    const int64_t m = (int64_t)h[0] + (int64_t)h[1];
    int64_t* dst = side == 0 ? idx_source_out : idx_target_out;
    if (dst && m > 0) {
      cudaMemcpyAsync(dst, dIdx, sizeof(long long) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream);
      cudaStreamSynchronize(ctx->stream);
    }
    *(side == 0 ? n_source_out : n_target_out) = m;
  }
  cleanup();
  return code;
}

namespace {

__global__ void toF32MovedKernel(const double* __restrict__ xyz, int64_t n, const double* __restrict__ Tm, float4* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
  if (Tm) {   // PointCloud::Transform
    double q[4];
    for (int r = 0; r < 4; ++r) q[r] = ((Tm[4 * r] * x + Tm[4 * r + 1] * y) + Tm[4 * r + 2] * z) + Tm[4 * r + 3] * 1.0;
    x = q[0] / q[3]; y = q[1] / q[3]; z = q[2] / q[3];
  }
  out[i] = make_float4((float)x, (float)y, (float)z, 1.f);
}

// G^T G of GetInformationMatrixFromPointClouds depends on the matched target points only through 10 moments:
// count, sum x, y, z, xx, yy, zz, xy, xz, yz (fp64).  Per-block partials; the host adds them in block order (deterministic).
constexpr int INFO_MOM = 10;
__global__ void __launch_bounds__(256) infoReduceKernel(const int* __restrict__ ids, int64_t n, const double* __restrict__ txyz, double* __restrict__ partials) {
  double m[INFO_MOM];
#pragma unroll
  for (int j = 0; j < INFO_MOM; ++j) m[j] = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int id = ids[i];
    if (id < 0) continue;
    const double x = txyz[3 * (size_t)id], y = txyz[3 * (size_t)id + 1], z = txyz[3 * (size_t)id + 2];
    m[0] += 1.0; m[1] += x; m[2] += y; m[3] += z;
    m[4] += x * x; m[5] += y * y; m[6] += z * z;
    m[7] += x * y; m[8] += x * z; m[9] += y * z;
  }
  __shared__ double sm[8][INFO_MOM];
#pragma unroll
  for (int j = 0; j < INFO_MOM; ++j)
    for (int o = 16; o > 0; o >>= 1) m[j] += __shfl_xor_sync(0xffffffffu, m[j], o);
  if ((threadIdx.x & 31) == 0)
    for (int j = 0; j < INFO_MOM; ++j) sm[threadIdx.x >> 5][j] = m[j];
  __syncthreads();
  if (threadIdx.x < INFO_MOM) {
    double v = 0.0;
    for (int w = 0; w < 8; ++w) v += sm[w][threadIdx.x];
    partials[(size_t)blockIdx.x * INFO_MOM + threadIdx.x] = v;
  }
}

}  // namespace

extern "C" int32_t b200reg_information_matrix(b200reg_ctx* ctx, const b200reg_cloud64* source, const b200reg_cloud64* target,
                                              double max_correspondence_distance, const double T_source_to_target[16], double info_out[36],
                                              int64_t* n_correspondences) {
  if (!ctx || !source || !target || !T_source_to_target || !info_out) return setErr(ctx, B200REG_E_INVALID_PARAM, "information_matrix: bad arguments");
  if (!(max_correspondence_distance > 0.0)) return setErr(ctx, B200REG_E_INVALID_PARAM, "information_matrix: max_correspondence_distance must be > 0");
  for (int i = 0; i < 36; ++i) info_out[i] = 0.0;
  if (n_correspondences) *n_correspondences = 0;
  const int64_t nS = source->n, nT = target->n;
  if (nS == 0 || nT == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  b200reg_cloud src, tgt;
  src.n = src.cap = nS;
  tgt.n = tgt.cap = nT;
  cudaError_t e = devAlloc(ctx, &src.xyz, sizeof(float4) * (size_t)nS);
  if (e == cudaSuccess) e = devAlloc(ctx, &tgt.xyz, sizeof(float4) * (size_t)nT);
  b200reg_index* ix = nullptr;
  auto cleanup = [&]() {
    if (ix) b200reg_index_free(ctx, ix);
    devFree(ctx, src.xyz); devFree(ctx, tgt.xyz);
    src.xyz = tgt.xyz = nullptr;
  };
  if (e != cudaSuccess) { cleanup(); return setErr(ctx, B200REG_E_OOM, "information_matrix: %s", cudaGetErrorString(e)); }
  double* dT = (double*)((char*)ctx->dstate + 40960);
  cudaMemcpyAsync(dT, T_source_to_target, sizeof(double) * 16, cudaMemcpyHostToDevice, ctx->stream);
  toF32MovedKernel<<<divUp(nS, 256), 256, 0, ctx->stream>>>(source->xyz, nS, dT, src.xyz);
  toF32MovedKernel<<<divUp(nT, 256), 256, 0, ctx->stream>>>(target->xyz, nT, nullptr, tgt.xyz);
  ctx->launches += 2;
  int code = buildIndex(ctx, &tgt, 0, 0.f, &ix);
  // KDTreeFlann::SearchHybrid(point, max_correspondence_distance, 1): the nearest target point with d2 < r^2
  if (code == B200REG_OK) code = runKnn(ctx, ix, src.xyz, nS, nullptr, 1, (float)max_correspondence_distance, 1, 0);
  const int blocks = (int)fmin((double)(2 * ctx->sm_count), (double)divUp(nS, 256));
  if (code == B200REG_OK) code = ensureBuf(ctx, ctx->partials, sizeof(double) * INFO_MOM * (size_t)blocks);
  if (code == B200REG_OK) code = ensurePinned(ctx, sizeof(double) * INFO_MOM * (size_t)blocks + 64);
  if (code != B200REG_OK) { cleanup(); return code; }
  infoReduceKernel<<<blocks, 256, 0, ctx->stream>>>((const int*)ctx->matchIds.p, nS, target->xyz, (double*)ctx->partials.p);
  ctx->launches++;
  cudaMemcpyAsync(ctx->pinned, ctx->partials.p, sizeof(double) * INFO_MOM * (size_t)blocks, cudaMemcpyDeviceToHost, ctx->stream);
  e = cudaStreamSynchronize(ctx->stream);
  const bool indexOk = (e != cudaSuccess) || indexHeader(ctx, ix) != nullptr;
  cleanup();
  if (e != cudaSuccess) return setErr(ctx, B200REG_E_CUDA, "information_matrix: %s", cudaGetErrorString(e));
  if (!indexOk) return setErr(ctx, B200REG_E_INVALID_PARAM, "information_matrix: non-finite coordinates in the target cloud");
  double m[INFO_MOM] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  const double* h = (const double*)ctx->pinned;
  for (int b = 0; b < blocks; ++b)
    for (int j = 0; j < INFO_MOM; ++j) m[j] += h[(size_t)b * INFO_MOM + j];
  // rows of G per correspondence: [0 z -y 1 0 0], [-z 0 x 0 1 0], [y -x 0 0 0 1]; GTG = sum of their outer products
  const double c = m[0], sx = m[1], sy = m[2], sz = m[3], xx = m[4], yy = m[5], zz = m[6], xy = m[7], xz = m[8], yz = m[9];
  double* G = info_out;
  G[0] = zz + yy; G[1] = -xy;     G[2] = -xz;     G[3] = 0.0; G[4] = -sz; G[5] = sy;
  G[6] = -xy;     G[7] = zz + xx; G[8] = -yz;     G[9] = sz;  G[10] = 0.0; G[11] = -sx;
  G[12] = -xz;    G[13] = -yz;    G[14] = yy + xx; G[15] = -sy; G[16] = sx; G[17] = 0.0;
  G[18] = 0.0;    G[19] = sz;     G[20] = -sy;    G[21] = c;  G[22] = 0.0; G[23] = 0.0;
  G[24] = -sz;    G[25] = 0.0;    G[26] = sx;     G[27] = 0.0; G[28] = c;  G[29] = 0.0;
  G[30] = sy;     G[31] = -sx;    G[32] = 0.0;    G[33] = 0.0; G[34] = 0.0; G[35] = c;
  if (n_correspondences) *n_correspondences = (int64_t)c;
  return B200REG_OK;
}
