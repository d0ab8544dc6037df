// csrc/common.cuh — internal types shared by the libb200reg.so translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <vector>

#include <math_constants.h>

#include "../../include/b200reg.h"

#define B200_PROF_MAX 256
#define B200_SM_COUNT 148   // B200: 2 dies x 74 SMs; grids are sized in multiples of this

struct DevBuf {   // growable device scratch (never shrinks; freed with the ctx)
  void* p = nullptr;
  size_t cap = 0;
};

struct IndexHeader;
struct b200reg_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, tev0 = nullptr, tev1 = nullptr;
  int sm_count = B200_SM_COUNT;
  char err[512] = {0};
  uint64_t launches = 0;
  // scratch
  DevBuf keys0, keys1, vals0, vals1, cubtmp, misc, partials, matchIds, matchD2, matchPos, flags, flush, cand;
  int flushToggle = 0;
  int profile = 0;           // b200reg_ctx_set_profiling
  int profCount = 0;
  cudaEvent_t* profEv = nullptr;   // 2 x B200_PROF_MAX events, created on first use
  int trace = 0;             // b200reg_ctx_set_trace: stamp kernel starts of the device-resident loop
  int select3 = 0;           // env B200REG_SELECT_3PASS=1: three-launch radix select inside the graph loop too (A/B)
  int hostLoop = 0;          // env B200REG_HOST_LOOP=1: keep the ICP iterate loop on the host (A/B against the graph loop)
  int debugSync = 0;         // env B200REG_DEBUG_SYNC=1: synchronise after every checked launch (fault localisation)
  int loopGridPerSm = 0;     // working blocks per SM of the graph loop's match kernel when the reference exceeds the L2 (0: all; the lanes of a batch: 2)
  size_t l2Bytes = 0;        // cudaDevAttrL2CacheSize
  int knnSeed = 1;           // env B200REG_KNN_SEED=0: do not seed the k = 1 match with the previous iteration's matches (A/B)
  int knnLanes = 0;          // lanes per query of the k = 1 match kernel: 0 automatic (8; the loop switches to 4 once far queries are rare),
                             // 8, 4, or 1 (the general kernel) forced by env B200REG_KNN_LANES for A/B runs
  void* pinned = nullptr;   // pinned host staging for small D2H results
  size_t pinned_cap = 0;
  void* dstate = nullptr;   // small device state block (select state, counters)
  // pinned slots the index builds copy their device-derived headers into (index.cu); chunks are freed with the ctx
  std::vector<void*> hdrChunks;
  std::vector<IndexHeader*> hdrFree;
  // next-scan prefetch (b200reg_cloud64_prefetch): host-to-device copies run here, ordered behind the compute stream through copyFence
  cudaStream_t copyStream = nullptr;
  cudaEvent_t copyFence = nullptr;
};

// Device memory comes from the device's stream-ordered pool (release threshold raised in b200reg_ctx_create so freed blocks stay
// cached): clouds, indices and growing scratch are allocated and freed on the ctx stream without a device-wide synchronisation.
// Objects belong to the ctx that created them.
template <class T>
inline cudaError_t devAlloc(b200reg_ctx* ctx, T** p, size_t bytes) {
  return cudaMallocAsync((void**)p, bytes > 0 ? bytes : 1, ctx->stream);
}
inline cudaError_t devFree(b200reg_ctx* ctx, void* p) {
  if (!p) return cudaSuccess;
  return ctx ? cudaFreeAsync(p, ctx->stream) : cudaFree(p);
}

struct b200reg_cloud {
  int64_t n = 0;
  int64_t cap = 0;
  float4* xyz = nullptr;   // x y z pad
  float4* nrm = nullptr;   // nx ny nz 0 (nullable)
};

struct GridDesc {
  float ox, oy, oz;        // origin (bbox min of the centred reference)
  float cell, inv_cell;
  float slack;             // conservative bound on fp32 cell-assignment / bound-evaluation error (see knn.cu)
  float ring2Safe;         // every point two or more cells away from the query's cell has d2 >= this
  int nx, ny, nz;
};

// device-resident mirror of an open3d::geometry::PointCloud (fp64 points + optional normals): what travels between the
// scan-processing steps of ScanToMapRegistration.cpp:36-69 without touching the host
struct b200reg_cloud64 {
  int64_t n = 0, cap = 0;
  double* xyz = nullptr;
  double* nrm = nullptr;   // allocated with xyz; valid iff hasNormals
  bool hasNormals = false;
  cudaEvent_t ready = nullptr;   // b200reg_cloud64_prefetch: completion of the copy-stream upload
  bool prefetched = false;       // an upload is in flight on the copy stream; b200reg_cloud64_acquire clears it
};

// What the index build derives ON THE DEVICE from the cloud (index.cu gridKernel): the kernels read it from HBM through
// b200reg_index::dHdr, the host gets an asynchronous copy (hHdr) it waits for only when it needs a value.
struct IndexHeader {
  GridDesc g;
  float mean[3];            // subtracted reference mean (0 if built un-centred)
  float bbMin[3], bbMax[3]; // bbox of the cloud as given
  double sum[3];            // fp64 coordinate sums
  long long m;
  unsigned numCells;
  int status;               // 0 ok, 1 non-finite coordinates in the cloud
};

struct b200reg_index {
  int64_t m = 0;
  int64_t mCap = 0;            // points the pts / nrm allocations hold (a recycled index keeps its storage)
  float4* pts = nullptr;       // sorted by cell: x y z (centred) + original index bits in .w
  float4* nrm = nullptr;       // sorted normals (nullable): nrmStore when the current content has normals
  float4* nrmStore = nullptr;  // the allocation behind nrm (kept when a recycled index is rebuilt over a cloud without normals)
  bool hasNrm = false;
  uint32_t* dirBase = nullptr; // allocation of the directory: 3 pad words, then cellStart[numCells + 1] (so that the counts, cellStart + 1, are 16-byte aligned)
  uint32_t* cellStart = nullptr;   // numCells + 1 entries (numCells <= cellCap)
  uint32_t cellCap = 0;
  IndexHeader* dHdr = nullptr;     // device
  IndexHeader* hHdr = nullptr;     // pinned host copy, valid once hdrReady has completed
  cudaEvent_t hdrReady = nullptr;
  bool hdrFetched = false;
};
// host view of the header (waits once for the build's asynchronous copy); nullptr + ctx error if the build failed on the device
const IndexHeader* indexHeader(b200reg_ctx* ctx, const b200reg_index* ix);
// enqueue the build of an index over `ref` on the ctx stream (centre != 0: after subtracting the cloud's mean); no host wait.
// recycle (nullable): an index of the SAME ctx that is no longer needed — its storage is reused when it is large enough (no
// allocator calls on the path of a periodic rebuild), otherwise it is freed; either way the caller must forget it.
int buildIndex(b200reg_ctx* ctx, const b200reg_cloud* ref, int centre, float cell, b200reg_index** out, b200reg_index* recycle = nullptr);
// Open3D EstimateNormals on an fp64 cloud (normals.cu); `rounded` = its fp32 copy, same order
int estimateNormals64(b200reg_ctx* ctx, b200reg_cloud64* cloud, const b200reg_cloud* rounded, int knn, double radius, const double* origin,
                      int32_t* degenerate_count);
IndexHeader* acquireHeaderSlot(b200reg_ctx* ctx);
void releaseHeaderSlot(b200reg_ctx* ctx, IndexHeader* slot);

// ---- device-resident registration loop (CUDA-graph while node, icp.cu): everything a kernel of the loop body needs is read
// from these two blocks in HBM, so ONE instantiated graph serves every reading size, reference and configuration.
#define CHK_HIST 64   // transforms remembered by the DifferentialTransformationChecker ring (smoothLength <= CHK_HIST - 1)
struct CheckerState {   // Counter / Differential / Bound checkers (TransformationCheckersImpl.cpp:58-225)
  int counter;
  int count;                 // transforms pushed so far
  float rot[CHK_HIST][4];    // quaternions (w x y z), entry i at i % CHK_HIST
  float tr[CHK_HIST][3];
  float q0[4], t0[3];
  int pad;
};
enum { ICPERR_NONE = 0, ICPERR_ROT_NAN, ICPERR_TRANS_NAN, ICPERR_BOUND, ICPERR_NO_QUANTILE, ICPERR_NO_POINTS, ICPERR_NOT_RIGID, ICPERR_ITER_CAP };
struct IcpDevParams {   // written by the host once per registration
  const float4* pts;
  const float4* nrm;
  const uint32_t* cellStart;
  const GridDesc* gp;     // the index header's grid (device-derived, index.cu)
  const float4* reading;
  const float4* readingNrm;
  long long n;
  int* ids;
  float* d2;
  int* pos;
  double* partials;
  float maxd2, trimRatio, cosAngle;
  int useTrim, useAngle, force2D, force4DOF;
  int maxIter, smoothLength;
  float minDiffRot, minDiffTrans, boundRot, boundTrans;
  int checkerOrder[3];
  int lanesForced;
  int matchBlocks;        // blocks of the match kernel's resident grid that take work (the rest return at once)
  int seedOn;             // 0: chosen per iteration from the far-query count
  // fused trimmed-distance selection (nullable = filter off): the match kernel histograms the top 11 bits of d2 into fhist0
  unsigned* fhist0;            // 2048 bins, zero between iterations
  unsigned* fhist1;            // 2048 bins
  unsigned* fcnt;              // [0] candidate count, [1] last-block ticket
  unsigned* cand;              // bit patterns of the distances in the selected 11-bit bucket (capacity n)
  unsigned long long* trace;   // nullable: [0] = entry count, [1 + i] = (kernel tag << 56) | %globaltimer at the start of the i-th kernel of the loop
};
struct IcpDevState {    // initialised by the host, advanced by icpTailKernel every iteration, read back once at the end
  float T[16];          // T_iter (row-major)
  int iterate, errKind, iterations, maxIterReached;
  float errA, errB, errC, errD;
  double kept, valid;
  float tau;
  int pad;
  unsigned long long visits;
  unsigned long long farPrev;   // far-query counter at the end of the previous iteration
  int lanes;                    // lanes per query of the next match launch
  int pad2;
  CheckerState chk;
};
#define ICP_STATE_HEAD_BYTES offsetof(IcpDevState, chk)

// resident blocks per SM of the k = 1 match kernels (knn.cu): 4 x 256 threads leave 64 registers per thread — no spills; 5, 6 and 8
// blocks (48 / 40 / 32 registers, 140-460 B of spills) and 2-3 blocks (80-119 registers) all measured slower (profiles/r1_notes.md)
#ifndef KNN1_LOOP_MINB
#define KNN1_LOOP_MINB 4
#endif

#define DSTATE_BYTES 131072      /* ctx->dstate; [32768, 49152) is scratch of normals.cu / map.cu */
#define DSTATE_TRACE_OFF 16384   /* 1 + 2047 x uint64 inside ctx->dstate */
#define DSTATE_FH0_OFF 65536     /* fused select: pass-0 histogram, 2048 x uint32 */
#define DSTATE_FH1_OFF 73728     /* fused select: pass-1 histogram, 2048 x uint32 */
#define DSTATE_FCNT_OFF 81920    /* fused select: candidate count, ticket */
#define TRACE_MAX 2047
#ifdef __CUDACC__
// device-side trace of the graph loop (ncu does not see inside a conditional node): one thread per kernel stamps its start
__device__ __forceinline__ void traceStamp(unsigned long long* trace, unsigned tag) {
  if (trace && blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    const unsigned long long i = atomicAdd(trace, 1ull);
    if (i < TRACE_MAX) trace[1 + i] = ((unsigned long long)tag << 56) | (t & 0x00ffffffffffffffull);
  }
}
#endif
int setErr(b200reg_ctx* ctx, int code, const char* fmt, ...);
// mean sums and bbox of a cloud on the HOST (one synchronisation; voxelgrid.cu — the index build derives its own on the device)
int cloudStats(b200reg_ctx* ctx, const b200reg_cloud* c, double sum[3], float mn[3], float mx[3]);
int ensureBuf(b200reg_ctx* ctx, DevBuf& b, size_t bytes);
int ensurePinned(b200reg_ctx* ctx, size_t bytes);
// event bracket around the match kernel while profiling is on (no-ops otherwise)
void profBegin(b200reg_ctx* ctx);
void profEnd(b200reg_ctx* ctx);

#define CUDA_TRY(ctx, expr)                                                                         \
  do {                                                                                              \
    cudaError_t _e = (expr);                                                                        \
    if (_e != cudaSuccess)                                                                          \
      return setErr((ctx), (_e == cudaErrorMemoryAllocation) ? B200REG_E_OOM : B200REG_E_CUDA,      \
                    "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__);    \
  } while (0)

#define LAUNCH_CHECK(ctx)                  \
  do {                                     \
    (ctx)->launches++;                     \
    CUDA_TRY((ctx), cudaGetLastError());   \
    if ((ctx)->debugSync) CUDA_TRY((ctx), cudaStreamSynchronize((ctx)->stream));   /* env B200REG_DEBUG_SYNC=1: localise a faulting kernel */ \
  } while (0)

#define TRY(expr)                 \
  do {                            \
    int _s = (expr);              \
    if (_s != B200REG_OK) return _s; \
  } while (0)

static inline int divUp(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// fp32 arithmetic with every operation individually rounded (no FMA contraction) — SURVEY 9.1 contract.
__device__ __forceinline__ float dist2_rn(float qx, float qy, float qz, float px, float py, float pz) {
  const float dx = __fsub_rn(qx, px), dy = __fsub_rn(qy, py), dz = __fsub_rn(qz, pz);
  return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

struct Rigid12 {   // rows 0..2 of a 4x4 (row-major)
  float m[12];
};

// features <- T * features, depth accumulated k = 0..3 without FMA (TransformationsImpl.cpp:79 under the SURVEY 9.1 contract)
__device__ __forceinline__ float3 xformPoint(const Rigid12& T, float x, float y, float z, float w) {
  float3 r;
  r.x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T.m[0], x), __fmul_rn(T.m[1], y)), __fmul_rn(T.m[2], z)), __fmul_rn(T.m[3], w));
  r.y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T.m[4], x), __fmul_rn(T.m[5], y)), __fmul_rn(T.m[6], z)), __fmul_rn(T.m[7], w));
  r.z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T.m[8], x), __fmul_rn(T.m[9], y)), __fmul_rn(T.m[10], z)), __fmul_rn(T.m[11], w));
  return r;
}
__device__ __forceinline__ float3 xformNormal(const Rigid12& T, float x, float y, float z) {
  float3 r;
  r.x = __fadd_rn(__fadd_rn(__fmul_rn(T.m[0], x), __fmul_rn(T.m[1], y)), __fmul_rn(T.m[2], z));
  r.y = __fadd_rn(__fadd_rn(__fmul_rn(T.m[4], x), __fmul_rn(T.m[5], y)), __fmul_rn(T.m[6], z));
  r.z = __fadd_rn(__fadd_rn(__fmul_rn(T.m[8], x), __fmul_rn(T.m[9], y)), __fmul_rn(T.m[10], z));
  return r;
}
