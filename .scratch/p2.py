p='open3d_slam_advanced_rss_2024_public_b200/csrc/map.cu'
s=open(p).read()
i0=s.index("int32_t b200reg_map_insert(")
i1=s.index("int32_t b200reg_map_crop(")
s=s[:i0]+r'''int32_t b200reg_map_insert(b200reg_ctx* ctx, b200reg_map* map, const double* scan_xyz, const double* scan_nrm, int64_t n,
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

'''+s[i1:]
i0=s.index("int32_t b200reg_crop_f64(")
s=s[:i0]+r'''int32_t b200reg_crop_f64(b200reg_ctx* ctx, const b200reg_crop_spec* volume, const double* xyz, const double* nrm, int64_t n,
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
// order (Open3D's is hash order — unpinned, SURVEY 8a-a17).  xyz_out / nrm_out hold n points; *n_out = voxels.
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

//@@CLOUD64@@
}  // extern "C"
'''
open(p,'w').write(s)
print("ok")
