p='open3d_slam_advanced_rss_2024_public_b200/csrc/map.cu'
s=open(p).read()
api=r'''// ---- the same steps on device-resident fp64 clouds (no host round trip between them) ------------------------------------
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

int64_t b200reg_cloud64_size(const b200reg_cloud64* c) { return c ? c->n : 0; }
int32_t b200reg_cloud64_has_normals(const b200reg_cloud64* c) { return (c && c->hasNormals) ? 1 : 0; }

void b200reg_cloud64_free(b200reg_ctx* ctx, b200reg_cloud64* c) {
  if (!c) return;
  if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
  cudaFree(c->xyz); cudaFree(c->nrm);
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
    cudaFree(c->xyz); cudaFree(c->nrm);
    c->xyz = c->nrm = nullptr;
    c->n = 0;
    c->cap = in->n + in->n / 4 + 1024;
    cudaError_t e = cudaMalloc(&c->xyz, sizeof(float4) * (size_t)c->cap);
    if (e == cudaSuccess) e = cudaMalloc(&c->nrm, sizeof(float4) * (size_t)c->cap);
    if (e != cudaSuccess) {
      cudaFree(c->xyz);
      c->xyz = nullptr;
      c->cap = 0;
      if (fresh) delete c;
      return setErr(ctx, B200REG_E_OOM, "cloud64_to_f32: %s", cudaGetErrorString(e));
    }
  }
  if (!in->hasNormals && c->nrm) {   // DataPoints without a normals descriptor
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(c->nrm);
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

// estimateNormalsOrCovariancesIfNeeded (CloudRegistration.cpp:62-75): fp32 copy of the points -> SurfaceNormal kernels (kNN,
// covariance, eigenvector, orientation toward `origin`) -> fp64 normals of the cloud.  Same arithmetic as
// b200reg_cloud_upload_f64 + b200reg_surface_normals + a host widening of the result.
int32_t b200reg_cloud64_estimate_normals(b200reg_ctx* ctx, b200reg_cloud64* inout, int32_t knn, double radius, const double origin[3],
                                         int32_t* degenerate_count) {
  if (!ctx || !inout) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud64_estimate_normals: bad arguments");
  if (degenerate_count) *degenerate_count = 0;
  if (inout->n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  b200reg_cloud* tmp = nullptr;
  inout->hasNormals = false;
  TRY(b200reg_cloud64_to_f32(ctx, inout, &tmp));
  float o[3] = {0.f, 0.f, 0.f};
  if (origin) for (int d = 0; d < 3; ++d) o[d] = (float)origin[d];
  int code = b200reg_surface_normals(ctx, tmp, knn, (float)radius, 0.f, origin ? o : nullptr, nullptr, nullptr, nullptr, degenerate_count);
  if (code == B200REG_OK) {
    normalsToF64Kernel<<<divUp(inout->n, 256), 256, 0, ctx->stream>>>(tmp->nrm, inout->n, inout->nrm);
    ctx->launches++;
    if (cudaGetLastError() != cudaSuccess) code = setErr(ctx, B200REG_E_CUDA, "cloud64_estimate_normals: launch failed");
    else inout->hasNormals = true;
  }
  b200reg_cloud_free(ctx, tmp);
  return code;
}
'''
assert "//@@CLOUD64@@" in s
s=s.replace("//@@CLOUD64@@\n", api)
open(p,'w').write(s)
print("ok")
