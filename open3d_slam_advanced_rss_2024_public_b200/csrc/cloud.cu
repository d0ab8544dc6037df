// csrc/cloud.cu — context, device clouds (DataPoints mirror), uploads/downloads, RigidTransformation.
#include <cstdlib>

#include "common.cuh"

int setErr(b200reg_ctx* ctx, int code, const char* fmt, ...) {
  if (ctx) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(ctx->err, sizeof(ctx->err), fmt, ap);
    va_end(ap);
  }
  return code;
}

int ensureBuf(b200reg_ctx* ctx, DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return B200REG_OK;
  if (b.p) {
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    CUDA_TRY(ctx, devFree(ctx, b.p));
    b.p = nullptr;
    b.cap = 0;
  }
  size_t want = bytes + bytes / 4 + 256;
  CUDA_TRY(ctx, devAlloc(ctx, &b.p, want));
  b.cap = want;
  return B200REG_OK;
}

// pinned 128-byte slots for the index headers (index.cu): handed out from 64-slot chunks that live as long as the ctx
IndexHeader* acquireHeaderSlot(b200reg_ctx* ctx) {
  static_assert(sizeof(IndexHeader) <= 128, "header slot size");
  if (ctx->hdrFree.empty()) {
    void* chunk = nullptr;
    if (cudaMallocHost(&chunk, 64 * 128) != cudaSuccess) return nullptr;
    ctx->hdrChunks.push_back(chunk);
    for (int i = 63; i >= 0; --i) ctx->hdrFree.push_back((IndexHeader*)((char*)chunk + 128 * i));
  }
  IndexHeader* s = ctx->hdrFree.back();
  ctx->hdrFree.pop_back();
  return s;
}
void releaseHeaderSlot(b200reg_ctx* ctx, IndexHeader* slot) {
  if (ctx && slot) ctx->hdrFree.push_back(slot);
}

void profBegin(b200reg_ctx* ctx) {
  if (ctx->profile && ctx->profCount < B200_PROF_MAX) cudaEventRecord(ctx->profEv[2 * ctx->profCount], ctx->stream);
}
void profEnd(b200reg_ctx* ctx) {
  if (ctx->profile && ctx->profCount < B200_PROF_MAX) { cudaEventRecord(ctx->profEv[2 * ctx->profCount + 1], ctx->stream); ++ctx->profCount; }
}

int ensurePinned(b200reg_ctx* ctx, size_t bytes) {
  if (bytes <= ctx->pinned_cap) return B200REG_OK;
  if (ctx->pinned) {
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    CUDA_TRY(ctx, cudaFreeHost(ctx->pinned));
    ctx->pinned = nullptr;
  }
  CUDA_TRY(ctx, cudaMallocHost(&ctx->pinned, bytes));
  ctx->pinned_cap = bytes;
  return B200REG_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------------------------------------------
__global__ void packF32Kernel(const float* __restrict__ feat, int stride, const float* __restrict__ nrm, int64_t n,
                              float4* __restrict__ xyz, float4* __restrict__ outN) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* p = feat + i * stride;
  xyz[i] = make_float4(p[0], p[1], p[2], stride >= 4 ? p[3] : 1.f);
  if (nrm) outN[i] = make_float4(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2], 0.f);
}

// open3dToPointmatcher (open3d_conversions.cpp:57-118): fp64 -> fp32 round-to-nearest, pad = 1
__global__ void packF64Kernel(const double* __restrict__ xyzIn, const double* __restrict__ nrm, int64_t n,
                              float4* __restrict__ xyz, float4* __restrict__ outN) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  xyz[i] = make_float4(__double2float_rn(xyzIn[3 * i]), __double2float_rn(xyzIn[3 * i + 1]), __double2float_rn(xyzIn[3 * i + 2]), 1.f);
  if (nrm) outN[i] = make_float4(__double2float_rn(nrm[3 * i]), __double2float_rn(nrm[3 * i + 1]), __double2float_rn(nrm[3 * i + 2]), 0.f);
}

__global__ void unpackF32Kernel(const float4* __restrict__ xyz, const float4* __restrict__ nrm, int64_t n, float* __restrict__ feat,
                                int stride, float* __restrict__ outN) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = xyz[i];
  float* o = feat + i * stride;
  o[0] = p.x; o[1] = p.y; o[2] = p.z;
  if (stride >= 4) o[3] = p.w;
  if (outN && nrm) {
    const float4 q = nrm[i];
    outN[3 * i] = q.x; outN[3 * i + 1] = q.y; outN[3 * i + 2] = q.z;
  }
}

// RigidTransformation::inPlaceCompute (TransformationsImpl.cpp:61-96)
__global__ void transformKernel(float4* __restrict__ xyz, float4* __restrict__ nrm, int64_t n, Rigid12 T, float t30, float t31,
                                float t32, float t33) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = xyz[i];
  const float3 r = xformPoint(T, p.x, p.y, p.z, p.w);
  const float w = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(t30, p.x), __fmul_rn(t31, p.y)), __fmul_rn(t32, p.z)), __fmul_rn(t33, p.w));
  xyz[i] = make_float4(r.x, r.y, r.z, w);
  if (nrm) {
    const float4 q = nrm[i];
    const float3 s = xformNormal(T, q.x, q.y, q.z);
    nrm[i] = make_float4(s.x, s.y, s.z, 0.f);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------
extern "C" {

int32_t b200reg_abi_version(void) { return B200REG_ABI_VERSION; }

int32_t b200reg_ctx_create(int32_t device, b200reg_ctx** out) {
  if (!out) return B200REG_E_INVALID_PARAM;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0 || device < 0 || device >= count) return B200REG_E_CUDA;   // no CPU fallback
  b200reg_ctx* ctx = new b200reg_ctx();
  ctx->device = device;
  if (cudaSetDevice(device) != cudaSuccess) { delete ctx; return B200REG_E_CUDA; }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) ctx->sm_count = prop.multiProcessorCount;
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return B200REG_E_CUDA; }
  cudaEventCreate(&ctx->ev0);
  cudaEventCreate(&ctx->ev1);
  cudaEventCreate(&ctx->tev0);
  cudaEventCreate(&ctx->tev1);
  {   // keep freed blocks in the pool instead of returning them to the driver at every synchronisation
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
      unsigned long long keep = ~0ull;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
  }
  if (devAlloc(ctx, &ctx->dstate, DSTATE_BYTES) != cudaSuccess) { delete ctx; return B200REG_E_OOM; }
  cudaMemsetAsync(ctx->dstate, 0, DSTATE_BYTES, ctx->stream);
  if (ensurePinned(ctx, 1 << 16) != B200REG_OK) { delete ctx; return B200REG_E_OOM; }
  const char* s3 = getenv("B200REG_SELECT_3PASS");
  ctx->select3 = (s3 && s3[0] == '1') ? 1 : 0;
  const char* hl = getenv("B200REG_HOST_LOOP");
  ctx->hostLoop = (hl && hl[0] == '1') ? 1 : 0;
  const char* dbg = getenv("B200REG_DEBUG_SYNC");
  ctx->debugSync = (dbg && dbg[0] == '1') ? 1 : 0;
  const char* seed = getenv("B200REG_KNN_SEED");
  ctx->knnSeed = (seed && seed[0] == '0') ? 0 : 1;
  { int l2 = 0; if (cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, device) == cudaSuccess && l2 > 0) ctx->l2Bytes = (size_t)l2; }
  const char* lanes = getenv("B200REG_KNN_LANES");
  if (lanes && (lanes[0] == '1' || lanes[0] == '4' || lanes[0] == '8')) ctx->knnLanes = lanes[0] - '0';
  *out = ctx;
  return B200REG_OK;
}

int32_t b200reg_ctx_set_profiling(b200reg_ctx* ctx, int32_t on) {
  if (!ctx) return B200REG_E_INVALID_PARAM;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (on && !ctx->profEv) {
    ctx->profEv = new cudaEvent_t[2 * B200_PROF_MAX];
    for (int i = 0; i < 2 * B200_PROF_MAX; ++i) CUDA_TRY(ctx, cudaEventCreate(&ctx->profEv[i]));
  }
  ctx->profile = on ? 1 : 0;
  ctx->profCount = 0;
  return B200REG_OK;
}

int32_t b200reg_ctx_profile_read(b200reg_ctx* ctx, float* ms_out, int32_t cap, int32_t* count) {
  if (!ctx || !count) return B200REG_E_INVALID_PARAM;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  int n = ctx->profCount < cap ? ctx->profCount : cap;
  for (int i = 0; i < n; ++i) CUDA_TRY(ctx, cudaEventElapsedTime(&ms_out[i], ctx->profEv[2 * i], ctx->profEv[2 * i + 1]));
  *count = n;
  ctx->profCount = 0;
  return B200REG_OK;
}

int32_t b200reg_ctx_set_trace(b200reg_ctx* ctx, int32_t on) {
  if (!ctx) return B200REG_E_INVALID_PARAM;
  ctx->trace = on ? 1 : 0;
  return B200REG_OK;
}

int32_t b200reg_ctx_trace_read(b200reg_ctx* ctx, uint64_t* out, int32_t cap, int32_t* count) {
  if (!ctx || !count || cap < 0) return B200REG_E_INVALID_PARAM;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  unsigned long long cnt = 0;
  CUDA_TRY(ctx, cudaMemcpy(&cnt, (char*)ctx->dstate + DSTATE_TRACE_OFF, sizeof(cnt), cudaMemcpyDeviceToHost));
  int n = (int)(cnt < (unsigned long long)TRACE_MAX ? cnt : TRACE_MAX);
  if (n > cap) n = cap;
  if (n > 0) CUDA_TRY(ctx, cudaMemcpy(out, (char*)ctx->dstate + DSTATE_TRACE_OFF + 8, sizeof(uint64_t) * (size_t)n, cudaMemcpyDeviceToHost));
  *count = n;
  return B200REG_OK;
}

void b200reg_ctx_destroy(b200reg_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->profEv) {
    for (int i = 0; i < 2 * B200_PROF_MAX; ++i) cudaEventDestroy(ctx->profEv[i]);
    delete[] ctx->profEv;
  }
  DevBuf* bufs[] = {&ctx->keys0, &ctx->keys1, &ctx->vals0, &ctx->vals1, &ctx->cubtmp, &ctx->misc, &ctx->partials,
                    &ctx->matchIds, &ctx->matchD2, &ctx->matchPos, &ctx->flags, &ctx->flush, &ctx->cand};
  for (DevBuf* b : bufs)
    if (b->p) devFree(ctx, b->p);
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  for (void* c : ctx->hdrChunks) cudaFreeHost(c);
  if (ctx->dstate) devFree(ctx, ctx->dstate);
  cudaStreamSynchronize(ctx->stream);
  cudaEventDestroy(ctx->ev0);
  cudaEventDestroy(ctx->ev1);
  cudaEventDestroy(ctx->tev0);
  cudaEventDestroy(ctx->tev1);
  if (ctx->copyStream) { cudaStreamSynchronize(ctx->copyStream); cudaStreamDestroy(ctx->copyStream); }
  if (ctx->copyFence) cudaEventDestroy(ctx->copyFence);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
}

int32_t b200reg_ctx_device(const b200reg_ctx* ctx) { return ctx ? ctx->device : -1; }

const char* b200reg_last_error(const b200reg_ctx* ctx) { return ctx ? ctx->err : "null context"; }

int32_t b200reg_ctx_sync(b200reg_ctx* ctx) {
  if (!ctx) return B200REG_E_INVALID_PARAM;
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return B200REG_OK;
}

uint64_t b200reg_ctx_launch_count(const b200reg_ctx* ctx) { return ctx ? ctx->launches : 0; }

int32_t b200reg_ctx_timer_start(b200reg_ctx* ctx) {
  if (!ctx) return B200REG_E_INVALID_PARAM;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaEventRecord(ctx->tev0, ctx->stream));
  return B200REG_OK;
}

int32_t b200reg_ctx_timer_stop(b200reg_ctx* ctx, float* elapsed_ms) {
  if (!ctx || !elapsed_ms) return B200REG_E_INVALID_PARAM;
  CUDA_TRY(ctx, cudaEventRecord(ctx->tev1, ctx->stream));
  CUDA_TRY(ctx, cudaEventSynchronize(ctx->tev1));
  CUDA_TRY(ctx, cudaEventElapsedTime(elapsed_ms, ctx->tev0, ctx->tev1));
  return B200REG_OK;
}

int32_t b200reg_ctx_flush_l2(b200reg_ctx* ctx) {
  if (!ctx) return B200REG_E_INVALID_PARAM;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const size_t bytes = (size_t)256 << 20;   // > 126 MB L2
  TRY(ensureBuf(ctx, ctx->flush, bytes));
  ctx->flushToggle ^= 1;
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->flush.p, ctx->flushToggle, bytes, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return B200REG_OK;
}

static int allocCloud(b200reg_ctx* ctx, int64_t n, bool normals, b200reg_cloud** out) {
  b200reg_cloud* c = new b200reg_cloud();
  c->n = n;
  c->cap = n > 0 ? n : 1;
  cudaError_t e = devAlloc(ctx, &c->xyz, sizeof(float4) * (size_t)c->cap);
  if (e == cudaSuccess && normals) e = devAlloc(ctx, &c->nrm, sizeof(float4) * (size_t)c->cap);
  if (e != cudaSuccess) {
    if (c->xyz) devFree(ctx, c->xyz);
    delete c;
    return setErr(ctx, B200REG_E_OOM, "cloud allocation of %lld points failed: %s", (long long)n, cudaGetErrorString(e));
  }
  *out = c;
  return B200REG_OK;
}

int32_t b200reg_cloud_upload_f32(b200reg_ctx* ctx, const float* feat, int32_t feat_stride, const float* normals, int64_t n,
                                 b200reg_cloud** out) {
  if (!ctx || !out || n < 0 || (n > 0 && !feat) || feat_stride < 3) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud_upload_f32: bad arguments");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  b200reg_cloud* c = nullptr;
  TRY(allocCloud(ctx, n, normals != nullptr, &c));
  if (n > 0) {
    const size_t fb = sizeof(float) * (size_t)n * feat_stride, nb = normals ? sizeof(float) * 3 * (size_t)n : 0;
    if (feat_stride == 4 && !normals) {
      CUDA_TRY(ctx, cudaMemcpyAsync(c->xyz, feat, fb, cudaMemcpyHostToDevice, ctx->stream));
    } else {
      TRY(ensureBuf(ctx, ctx->misc, fb + nb + 16));
      float* dfeat = (float*)ctx->misc.p;
      float* dn = normals ? (float*)((char*)ctx->misc.p + ((fb + 15) / 16) * 16) : nullptr;
      CUDA_TRY(ctx, cudaMemcpyAsync(dfeat, feat, fb, cudaMemcpyHostToDevice, ctx->stream));
      if (normals) CUDA_TRY(ctx, cudaMemcpyAsync(dn, normals, nb, cudaMemcpyHostToDevice, ctx->stream));
      packF32Kernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(dfeat, feat_stride, dn, n, c->xyz, c->nrm);
      LAUNCH_CHECK(ctx);
    }
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));   // host buffers may be reused by the caller on return
  }
  *out = c;
  return B200REG_OK;
}

// Re-fill an existing device cloud from host memory (the streaming case: one scan buffer reused for every scan).  Capacity
// grows when needed and never shrinks.  async != 0: the copy is only enqueued — the caller keeps `feat` / `normals` valid and
// unmodified until the next call on this ctx that synchronises (any call returning results, or b200reg_ctx_sync); with
// pinned host memory upload and registration then cost a single synchronisation.
int32_t b200reg_cloud_update_f32(b200reg_ctx* ctx, b200reg_cloud* c, const float* feat, int32_t feat_stride, const float* normals,
                                 int64_t n, int32_t async) {
  if (!ctx || !c || n < 0 || (n > 0 && !feat) || feat_stride < 3) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud_update_f32: bad arguments");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (n > c->cap || (normals && !c->nrm)) {
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    const int64_t cap = n > c->cap ? n : c->cap;
    float4 *nx = nullptr, *nn = nullptr;
    cudaError_t e = devAlloc(ctx, &nx, sizeof(float4) * (size_t)cap);
    if (e == cudaSuccess && normals) e = devAlloc(ctx, &nn, sizeof(float4) * (size_t)cap);
    if (e != cudaSuccess) {
      if (nx) devFree(ctx, nx);
      return setErr(ctx, B200REG_E_OOM, "cloud_update_f32: allocation of %lld points failed: %s", (long long)cap, cudaGetErrorString(e));
    }
    if (c->xyz) devFree(ctx, c->xyz);
    if (c->nrm) devFree(ctx, c->nrm);
    c->xyz = nx; c->nrm = nn; c->cap = cap;
  }
  c->n = n;
  if (!normals && c->nrm) {   // the new content has no 'normals' descriptor: the previous scan's must not survive (stream-ordered free)
    CUDA_TRY(ctx, devFree(ctx, c->nrm));
    c->nrm = nullptr;
  }
  if (n == 0) return B200REG_OK;
  const size_t fb = sizeof(float) * (size_t)n * feat_stride, nb = normals ? sizeof(float) * 3 * (size_t)n : 0;
  if (feat_stride == 4 && !normals) {
    CUDA_TRY(ctx, cudaMemcpyAsync(c->xyz, feat, fb, cudaMemcpyHostToDevice, ctx->stream));
  } else {
    TRY(ensureBuf(ctx, ctx->misc, fb + nb + 16));
    float* dfeat = (float*)ctx->misc.p;
    float* dn = normals ? (float*)((char*)ctx->misc.p + ((fb + 15) / 16) * 16) : nullptr;
    CUDA_TRY(ctx, cudaMemcpyAsync(dfeat, feat, fb, cudaMemcpyHostToDevice, ctx->stream));
    if (normals) CUDA_TRY(ctx, cudaMemcpyAsync(dn, normals, nb, cudaMemcpyHostToDevice, ctx->stream));
    packF32Kernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(dfeat, feat_stride, dn, n, c->xyz, normals ? c->nrm : nullptr);
    LAUNCH_CHECK(ctx);
  }
  if (!async) CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return B200REG_OK;
}

int32_t b200reg_cloud_upload_f64(b200reg_ctx* ctx, const double* xyz, const double* normals, int64_t n, b200reg_cloud** out) {
  if (!ctx || !out || n < 0 || (n > 0 && !xyz)) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud_upload_f64: bad arguments");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  b200reg_cloud* c = nullptr;
  TRY(allocCloud(ctx, n, normals != nullptr, &c));
  if (n > 0) {
    const size_t fb = sizeof(double) * 3 * (size_t)n;
    TRY(ensureBuf(ctx, ctx->misc, 2 * fb));
    double* dx = (double*)ctx->misc.p;
    double* dn = normals ? dx + 3 * (size_t)n : nullptr;
    CUDA_TRY(ctx, cudaMemcpyAsync(dx, xyz, fb, cudaMemcpyHostToDevice, ctx->stream));
    if (normals) CUDA_TRY(ctx, cudaMemcpyAsync(dn, normals, fb, cudaMemcpyHostToDevice, ctx->stream));
    packF64Kernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(dx, dn, n, c->xyz, c->nrm);
    LAUNCH_CHECK(ctx);
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  }
  *out = c;
  return B200REG_OK;
}

int32_t b200reg_cloud_download_f32(b200reg_ctx* ctx, const b200reg_cloud* cloud, float* feat, int32_t feat_stride, float* normals) {
  if (!ctx || !cloud || feat_stride < 3) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud_download_f32: bad arguments");
  if (normals && !cloud->nrm) return setErr(ctx, B200REG_E_INVALID_FIELD, "cloud has no 'normals' descriptor");
  const int64_t n = cloud->n;
  if (n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const size_t fb = sizeof(float) * (size_t)n * feat_stride, nb = normals ? sizeof(float) * 3 * (size_t)n : 0;
  TRY(ensureBuf(ctx, ctx->misc, fb + nb + 16));
  float* dfeat = (float*)ctx->misc.p;
  float* dn = normals ? (float*)((char*)ctx->misc.p + ((fb + 15) / 16) * 16) : nullptr;
  unpackF32Kernel<<<divUp(n, 256), 256, 0, ctx->stream>>>(cloud->xyz, cloud->nrm, n, dfeat, feat_stride, dn);
  LAUNCH_CHECK(ctx);
  if (feat) CUDA_TRY(ctx, cudaMemcpyAsync(feat, dfeat, fb, cudaMemcpyDeviceToHost, ctx->stream));
  if (normals) CUDA_TRY(ctx, cudaMemcpyAsync(normals, dn, nb, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return B200REG_OK;
}

int64_t b200reg_cloud_size(const b200reg_cloud* cloud) { return cloud ? cloud->n : 0; }
int32_t b200reg_cloud_has_normals(const b200reg_cloud* cloud) { return (cloud && cloud->nrm) ? 1 : 0; }

void b200reg_cloud_free(b200reg_ctx* ctx, b200reg_cloud* cloud) {
  if (!cloud) return;
  if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
  if (cloud->xyz) devFree(ctx, cloud->xyz);
  if (cloud->nrm) devFree(ctx, cloud->nrm);
  delete cloud;
}

int32_t b200reg_cloud_transform(b200reg_ctx* ctx, b200reg_cloud* inout, const float T[16]) {
  if (!ctx || !inout || !T) return setErr(ctx, B200REG_E_INVALID_PARAM, "cloud_transform: bad arguments");
  // checkParameters (TransformationsImpl.cpp:100-114): |1 - det R| > 1e-3 -> TransformationError
  const float det = T[0] * (T[5] * T[10] - T[6] * T[9]) - T[1] * (T[4] * T[10] - T[6] * T[8]) + T[2] * (T[4] * T[9] - T[5] * T[8]);
  if (fabsf(1.f - det) > 0.001f) return setErr(ctx, B200REG_E_TRANSFORM, "RigidTransformation: Error, rotation matrix is not orthogonal.");
  if (inout->n == 0) return B200REG_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Rigid12 R;
  memcpy(R.m, T, sizeof(R.m));
  transformKernel<<<divUp(inout->n, 256), 256, 0, ctx->stream>>>(inout->xyz, inout->nrm, inout->n, R, T[12], T[13], T[14], T[15]);
  LAUNCH_CHECK(ctx);
  return B200REG_OK;
}

}  // extern "C"
