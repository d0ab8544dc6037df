p='open3d_slam_advanced_rss_2024_public_b200/csrc/map.cu'
s=open(p).read()
s=s.replace("""namespace {

typedef unsigned long long u64;
""","""// device-resident mirror of an open3d::geometry::PointCloud (fp64 points + optional normals): what travels between the
// scan-processing steps of ScanToMapRegistration.cpp:36-69 without touching the host
struct b200reg_cloud64 {
  int64_t n = 0, cap = 0;
  double* xyz = nullptr;
  double* nrm = nullptr;   // allocated with xyz; valid iff hasNormals
  bool hasNormals = false;
};

namespace {

typedef unsigned long long u64;
""",1)
core = r'''
int cloud64Reserve(b200reg_ctx* ctx, b200reg_cloud64* c, int64_t want) {
  if (want <= c->cap) return B200REG_OK;
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  cudaFree(c->xyz); cudaFree(c->nrm);
  c->xyz = c->nrm = nullptr;
  c->cap = 0;
  const int64_t cap = want + want / 4 + 1024;
  cudaError_t e = cudaMalloc(&c->xyz, sizeof(double) * 3 * (size_t)cap);
  if (e == cudaSuccess) e = cudaMalloc(&c->nrm, sizeof(double) * 3 * (size_t)cap);
  if (e != cudaSuccess) { cudaFree(c->xyz); c->xyz = nullptr; return setErr(ctx, B200REG_E_OOM, "cloud64: cannot reserve %lld points: %s", (long long)cap, cudaGetErrorString(e)); }
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
  TRY(ensureBuf(ctx, ctx->partials, sizeof(double) * 3 * (size_t)mb));
  minBoundKernel<<<mb, 256, 0, ctx->stream>>>(dIn, n, (double*)ctx->partials.p);
  LAUNCH_CHECK(ctx);
  TRY(ensurePinned(ctx, sizeof(double) * 3 * (size_t)mb + 64));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pinned, ctx->partials.p, sizeof(double) * 3 * (size_t)mb, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  double mn[3] = {INFINITY, INFINITY, INFINITY};
  for (int b = 0; b < mb; ++b) for (int d = 0; d < 3; ++d) mn[d] = fmin(mn[d], ((const double*)ctx->pinned)[3 * b + d]);
  for (int d = 0; d < 3; ++d) mn[d] = mn[d] - voxel * 0.5;
  int* dErr = (int*)((char*)ctx->dstate + 32768 + 64);
  CUDA_TRY(ctx, cudaMemsetAsync(dErr, 0, sizeof(int), ctx->stream));
  o3dKeyKernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(dIn, n, mn[0], mn[1], mn[2], voxel, (u64*)ctx->keys0.p, (unsigned*)ctx->vals0.p, dErr);
  LAUNCH_CHECK(ctx);
  cub::DoubleBuffer<u64> dk((u64*)ctx->keys0.p, (u64*)ctx->keys1.p);
  cub::DoubleBuffer<unsigned> dv((unsigned*)ctx->vals0.p, (unsigned*)ctx->vals1.p);
  size_t tmp = 0;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp, dk, dv, (int)n, 0, 63, ctx->stream));
  TRY(ensureBuf(ctx, ctx->cubtmp, tmp));
  tmp = ctx->cubtmp.cap;
  CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(ctx->cubtmp.p, tmp, dk, dv, (int)n, 0, 63, ctx->stream));
  ctx->launches += 9;
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

}  // namespace

extern "C" {
'''
s=s.replace("""
}  // namespace

extern "C" {
""", core, 1)
open(p,'w').write(s)
print("ok")
