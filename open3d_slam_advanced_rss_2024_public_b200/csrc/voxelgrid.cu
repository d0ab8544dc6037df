// csrc/voxelgrid.cu — VoxelGridDataPointsFilter::inPlaceFilter (DataPointsFilters/VoxelGrid.cpp:71-344) on the device.
//
// Membership is bit-exact: the cell of a point is computed with the reference's fp32 operations in the reference's
// order ( floor(x / vSize - minBound), VoxelGrid.cpp:157-159; numDiv = unsigned(1 + maxBound - minBound), :115-121 ).
// Points are sorted by the 64-bit key (cell << 32 | pointIndex); one thread per voxel then walks its points in
// ascending input order and accumulates in fp32 exactly like the reference accumulates into the first point's slot
// (:192-222), so centroids are bit-exact as well.  Output order = ascending first-point index (:322-334).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cmath>

#include "common.cuh"


namespace {

struct VgParams {
  float vx, vy, vz;
  float minBx, minBy, minBz;
  unsigned nx, ny, nz;
};

__global__ void vgKeyKernel(const float4* __restrict__ xyz, int64_t n, VgParams P, unsigned long long* __restrict__ keys) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const float4 v = xyz[p];
  const unsigned i = (unsigned)floorf(__fsub_rn(__fdiv_rn(v.x, P.vx), P.minBx));
  const unsigned j = (unsigned)floorf(__fsub_rn(__fdiv_rn(v.y, P.vy), P.minBy));
  const unsigned k = (unsigned)floorf(__fsub_rn(__fdiv_rn(v.z, P.vz), P.minBz));
  const unsigned idx = i + j * P.nx + k * P.nx * P.ny;   // unsigned wrap-around like the reference
  keys[p] = ((unsigned long long)idx << 32) | (unsigned long long)(unsigned)p;
}

// one thread per sorted slot; segment heads reduce their voxel sequentially (ascending input index)
__global__ void vgReduceKernel(const unsigned long long* __restrict__ keys, int64_t n, const float4* __restrict__ xyz,
                               const float4* __restrict__ nrm, VgParams P, int useCentroid, int avgDesc, float4* __restrict__ tmpXyz,
                               float4* __restrict__ tmpNrm, unsigned* __restrict__ flags) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const unsigned long long key = keys[j];
  const unsigned idx = (unsigned)(key >> 32);
  if (j > 0 && (unsigned)(keys[j - 1] >> 32) == idx) return;   // not a head
  const unsigned first = (unsigned)key;
  float4 a = xyz[first];
  float4 an = nrm ? nrm[first] : make_float4(0, 0, 0, 0);
  unsigned count = 1;
  for (int64_t t = j + 1; t < n; ++t) {
    const unsigned long long kk = keys[t];
    if ((unsigned)(kk >> 32) != idx) break;
    const unsigned p = (unsigned)kk;
    ++count;
    if (useCentroid) {
      const float4 v = xyz[p];
      a.x = __fadd_rn(a.x, v.x); a.y = __fadd_rn(a.y, v.y); a.z = __fadd_rn(a.z, v.z);
    }
    if (avgDesc && nrm) {
      const float4 v = nrm[p];
      an.x = __fadd_rn(an.x, v.x); an.y = __fadd_rn(an.y, v.y); an.z = __fadd_rn(an.z, v.z);
    }
  }
  const float c = (float)count;
  if (useCentroid) {
    a.x = __fdiv_rn(a.x, c); a.y = __fdiv_rn(a.y, c); a.z = __fdiv_rn(a.z, c);
  } else {   // cell centre, corrected form of VoxelGrid.cpp:285-304: (minBound + index + 0.5) * vSize
    const unsigned kz = idx / (P.nx * P.ny);
    const unsigned jy = (idx - kz * P.nx * P.ny) / P.nx;
    const unsigned ix = idx - kz * P.nx * P.ny - jy * P.nx;
    a.x = __fmul_rn(__fadd_rn(__fadd_rn(P.minBx, (float)ix), 0.5f), P.vx);
    a.y = __fmul_rn(__fadd_rn(__fadd_rn(P.minBy, (float)jy), 0.5f), P.vy);
    a.z = __fmul_rn(__fadd_rn(__fadd_rn(P.minBz, (float)kz), 0.5f), P.vz);
  }
  if (avgDesc && nrm) { an.x = __fdiv_rn(an.x, c); an.y = __fdiv_rn(an.y, c); an.z = __fdiv_rn(an.z, c); }
  tmpXyz[first] = a;
  if (nrm) tmpNrm[first] = an;
  flags[first] = 1u;
}

__global__ void vgCompactKernel(const unsigned* __restrict__ flags, const unsigned* __restrict__ offs, int64_t n,
                                const float4* __restrict__ tmpXyz, const float4* __restrict__ tmpNrm, float4* __restrict__ xyz,
                                float4* __restrict__ nrm) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n || !flags[p]) return;
  const unsigned o = offs[p];
  xyz[o] = tmpXyz[p];
  if (nrm) nrm[o] = tmpNrm[p];
}

}  // namespace

extern "C" int32_t b200reg_voxelgrid(b200reg_ctx* ctx, b200reg_cloud* inout, float vx, float vy, float vz, int32_t use_centroid,
                                     int32_t average_descriptors) {
  if (!ctx || !inout) return setErr(ctx, B200REG_E_INVALID_PARAM, "voxelgrid: bad arguments");
  if (!(vx >= 0.001f) || !(vy >= 0.001f) || !(vz >= 0.001f))   // VoxelGrid.h:69-71 bounds
    return setErr(ctx, B200REG_E_INVALID_PARAM, "VoxelGridDataPointsFilter: vSize must be >= 0.001");
  const int64_t n = inout->n;
  if (n == 0) return B200REG_OK;
  if (n >= (int64_t)0xffffffffu) return setErr(ctx, B200REG_E_INVALID_PARAM, "voxelgrid: too many points");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  double sum[3];
  float mn[3], mx[3];
  TRY(cloudStats(ctx, inout, sum, mn, mx));
  VgParams P;
  P.vx = vx; P.vy = vy; P.vz = vz;
  // VoxelGrid.cpp:100-121, fp32 host arithmetic in the reference's order
  const float minBx = mn[0] / vx, maxBx = mx[0] / vx, minBy = mn[1] / vy, maxBy = mx[1] / vy, minBz = mn[2] / vz, maxBz = mx[2] / vz;
  P.minBx = minBx; P.minBy = minBy; P.minBz = minBz;
  volatile float dx = 1 + maxBx, dy = 1 + maxBy, dz = 1 + maxBz;
  dx = dx - minBx; dy = dy - minBy; dz = dz - minBz;
  P.nx = (unsigned)dx; P.ny = (unsigned)dy; P.nz = (unsigned)dz;
  const unsigned numVox = P.nx * P.ny * P.nz;
  if (numVox == 0)   // VoxelGrid.cpp:127-130
    return setErr(ctx, B200REG_E_INVALID_PARAM, "VoxelGridDataPointsFilter: The number of voxel couldn't be computed. There might be NaNs in the feature matrix.");
  TRY(ensureBuf(ctx, ctx->keys0, sizeof(unsigned long long) * (size_t)n));
  TRY(ensureBuf(ctx, ctx->keys1, sizeof(unsigned long long) * (size_t)n));
  vgKeyKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(inout->xyz, n, P, (unsigned long long*)ctx->keys0.p);
  LAUNCH_CHECK(ctx);
  cub::DoubleBuffer<unsigned long long> dk((unsigned long long*)ctx->keys0.p, (unsigned long long*)ctx->keys1.p);
  size_t tmpSort = 0, tmpScan = 0;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortKeys(nullptr, tmpSort, dk, (int)n, 0, 64, ctx->stream));
  // tmp arrays: results stashed at the first point's slot, flags + offsets for the order-preserving compaction
  const size_t f4 = sizeof(float4) * (size_t)n, u4 = sizeof(unsigned) * (size_t)n;
  TRY(ensureBuf(ctx, ctx->misc, 2 * f4 + 2 * u4 + 64));
  float4* tmpXyz = (float4*)ctx->misc.p;
  float4* tmpNrm = tmpXyz + n;
  unsigned* flags = (unsigned*)(tmpNrm + n);
  unsigned* offs = flags + n;
  CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmpScan, flags, offs, (int)n, ctx->stream));
  TRY(ensureBuf(ctx, ctx->cubtmp, tmpSort > tmpScan ? tmpSort : tmpScan));
  size_t tb = ctx->cubtmp.cap;
  // only the bits that can be set: 32 index bits + the cell bits
  int cellBits = 1;
  while (cellBits < 32 && (1ull << cellBits) < (unsigned long long)numVox) ++cellBits;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortKeys(ctx->cubtmp.p, tb, dk, (int)n, 0, 32 + cellBits, ctx->stream));
  ctx->launches += (uint64_t)((32 + cellBits + 7) / 8 + 1);
  CUDA_TRY(ctx, cudaMemsetAsync(flags, 0, u4, ctx->stream));
  vgReduceKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(dk.Current(), n, inout->xyz, inout->nrm, P, use_centroid, average_descriptors, tmpXyz,
                                                         tmpNrm, flags);
  LAUNCH_CHECK(ctx);
  tb = ctx->cubtmp.cap;
  CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(ctx->cubtmp.p, tb, flags, offs, (int)n, ctx->stream));
  ctx->launches += 2;
  vgCompactKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(flags, offs, n, tmpXyz, tmpNrm, inout->xyz, inout->nrm);
  LAUNCH_CHECK(ctx);
  unsigned* h = (unsigned*)ctx->pinned;
  CUDA_TRY(ctx, cudaMemcpyAsync(h, offs + (n - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(h + 1, flags + (n - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  inout->n = (int64_t)h[0] + (int64_t)h[1];
  return B200REG_OK;
}
