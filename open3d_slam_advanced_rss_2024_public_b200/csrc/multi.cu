// csrc/multi.cu — independent registrations sharded over the GPUs of one box from ONE process (the loop it replaces:
// open3d_slam/src/PlaceRecognition.cpp:71-174, independent iterations of a `for` over loop-closure candidates).
//
// A single registration does not shard (a chain of tiny global reductions over a working set that fits one GPU), so the only
// multi-GPU structure of this path is: pair i -> device i mod G, every device runs its pairs through its own b200reg_batch
// (lanes in flight, csrc/icp.cu) driven by its own host thread, and ONE ncclAllGather of fixed-size records over NVLink leaves every
// device — and the caller — with every pair's result.  There is no other collective and no data-path exchange.
//
// NCCL is resolved at run time (dlopen of libnccl.so.2, or the path in env B200REG_NCCL_LIB): libb200reg.so carries no link-time
// dependency on it, a process that already holds an NCCL (e.g. one that imported torch) shares that copy, and single-GPU users never
// load it.
#include <dlfcn.h>
#include <nccl.h>

#include <string>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {

struct NcclApi {
  void* lib = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  bool ok = false;
};

NcclApi& nccl() {
  static NcclApi api;
  if (api.lib) return api;
  // order: the library the user names (env B200REG_NCCL_LIB), an NCCL the process already holds (RTLD_NOLOAD), the system's.  A process
  // that will load a NEWER NCCL under the same soname later (e.g. `import torch` after this call) should load it first or name it here:
  // the dynamic loader keeps one libnccl.so.2 per process.
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  if (const char* user = getenv("B200REG_NCCL_LIB")) api.lib = dlopen(user, RTLD_NOW | RTLD_GLOBAL);
  for (const char* nm : names)
    if (!api.lib) api.lib = dlopen(nm, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
  for (const char* nm : names)
    if (!api.lib) api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
  if (!api.lib) return api;
#define LOAD(field, sym) api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.lib, sym))
  LOAD(GetErrorString, "ncclGetErrorString");
  LOAD(GetVersion, "ncclGetVersion");
  LOAD(CommInitAll, "ncclCommInitAll");
  LOAD(CommDestroy, "ncclCommDestroy");
  LOAD(AllGather, "ncclAllGather");
  LOAD(GroupStart, "ncclGroupStart");
  LOAD(GroupEnd, "ncclGroupEnd");
#undef LOAD
  api.ok = api.GetErrorString && api.CommInitAll && api.CommDestroy && api.AllGather && api.GroupStart && api.GroupEnd;
  return api;
}

// what travels in the gather: 24 floats = 96 bytes per pair
struct PairRecord {
  float T[16];
  float status, iterations, kept, tau;
  float owner, valid, pad0, pad1;
};
static_assert(sizeof(PairRecord) == 96, "record size");

}  // namespace

struct b200reg_multi {
  int G = 0;
  b200reg_icp_config cfg{};
  std::vector<int> devices;
  std::vector<b200reg_ctx*> ctx;
  std::vector<b200reg_batch*> batch;
  std::vector<ncclComm_t> comm;
  std::vector<PairRecord*> dSend, dRecv;   // device buffers, grown on demand
  std::vector<PairRecord*> hSend;           // pinned host staging per device
  PairRecord* hRecv = nullptr;              // pinned: device 0's gathered block
  size_t slotsCap = 0;
  int ncclVersion = 0;
  char err[512] = {0};
};

namespace {

int multiErr(b200reg_multi* M, int code, const std::string& msg) {
  if (M) snprintf(M->err, sizeof(M->err), "%s", msg.c_str());
  return code;
}

int ensureSlots(b200reg_multi* M, size_t slots) {
  if (slots <= M->slotsCap) return B200REG_OK;
  const size_t cap = slots + slots / 2 + 8;
  for (int d = 0; d < M->G; ++d) {
    cudaSetDevice(M->devices[d]);
    cudaStreamSynchronize(M->ctx[d]->stream);
    if (M->dSend[d]) cudaFree(M->dSend[d]);
    if (M->dRecv[d]) cudaFree(M->dRecv[d]);
    if (M->hSend[d]) cudaFreeHost(M->hSend[d]);
    M->dSend[d] = M->dRecv[d] = nullptr;
    M->hSend[d] = nullptr;
    if (cudaMalloc(&M->dSend[d], sizeof(PairRecord) * cap) != cudaSuccess || cudaMalloc(&M->dRecv[d], sizeof(PairRecord) * cap * M->G) != cudaSuccess ||
        cudaMallocHost(&M->hSend[d], sizeof(PairRecord) * cap) != cudaSuccess)
      return multiErr(M, B200REG_E_OOM, "multi: record buffers");
  }
  if (M->hRecv) cudaFreeHost(M->hRecv);
  M->hRecv = nullptr;
  if (cudaMallocHost(&M->hRecv, sizeof(PairRecord) * cap * M->G) != cudaSuccess) return multiErr(M, B200REG_E_OOM, "multi: record buffers");
  M->slotsCap = cap;
  return B200REG_OK;
}

}  // namespace

extern "C" {

int32_t b200reg_multi_create(const int32_t* devices, int32_t n_devices, const b200reg_icp_config* cfg, int32_t lanes, b200reg_multi** out) {
  if (!out) return B200REG_E_INVALID_PARAM;
  *out = nullptr;
  if (!cfg || n_devices < 0 || (n_devices > 0 && !devices)) return B200REG_E_INVALID_PARAM;
  int visible = 0;
  if (cudaGetDeviceCount(&visible) != cudaSuccess || visible <= 0) return B200REG_E_CUDA;   // no CPU fallback
  b200reg_multi* M = new b200reg_multi();
  M->cfg = *cfg;
  if (n_devices == 0) for (int d = 0; d < visible; ++d) M->devices.push_back(d);   // every visible GPU
  else M->devices.assign(devices, devices + n_devices);
  M->G = (int)M->devices.size();
  M->ctx.assign(M->G, nullptr);
  M->batch.assign(M->G, nullptr);
  M->comm.assign(M->G, nullptr);
  M->dSend.assign(M->G, nullptr);
  M->dRecv.assign(M->G, nullptr);
  M->hSend.assign(M->G, nullptr);
  *out = M;   // from here on the caller can read the error text and must destroy the object
  for (int d = 0; d < M->G; ++d) {
    int code = b200reg_ctx_create(M->devices[d], &M->ctx[d]);
    if (code != B200REG_OK) return multiErr(M, code, "multi_create: no usable CUDA device " + std::to_string(M->devices[d]));
    code = b200reg_batch_create(M->ctx[d], cfg, lanes, &M->batch[d]);
    if (code != B200REG_OK) return multiErr(M, code, std::string("multi_create: ") + b200reg_last_error(M->ctx[d]));
  }
  NcclApi& N = nccl();
  if (!N.ok) return multiErr(M, B200REG_E_CUDA, "multi_create: libnccl.so.2 could not be loaded (the result gather needs NCCL)");
  if (N.GetVersion) N.GetVersion(&M->ncclVersion);
  const ncclResult_t r = N.CommInitAll(M->comm.data(), M->G, M->devices.data());
  if (r != ncclSuccess) {
    for (auto& c : M->comm) c = nullptr;
    return multiErr(M, B200REG_E_CUDA, std::string("multi_create: ncclCommInitAll: ") + N.GetErrorString(r));
  }
  return B200REG_OK;
}

void b200reg_multi_destroy(b200reg_multi* M) {
  if (!M) return;
  NcclApi& N = nccl();
  for (int d = 0; d < M->G; ++d) {
    cudaSetDevice(M->devices[d]);
    if (M->ctx[d]) cudaStreamSynchronize(M->ctx[d]->stream);
    if (M->comm[d] && N.ok) N.CommDestroy(M->comm[d]);
    if (M->dSend[d]) cudaFree(M->dSend[d]);
    if (M->dRecv[d]) cudaFree(M->dRecv[d]);
    if (M->hSend[d]) cudaFreeHost(M->hSend[d]);
    if (M->batch[d]) b200reg_batch_destroy(M->batch[d]);
    if (M->ctx[d]) b200reg_ctx_destroy(M->ctx[d]);
  }
  if (M->hRecv) cudaFreeHost(M->hRecv);
  delete M;
}

int32_t b200reg_multi_devices(const b200reg_multi* M) { return M ? M->G : 0; }
int32_t b200reg_multi_nccl_version(const b200reg_multi* M) { return M ? M->ncclVersion : 0; }
const char* b200reg_multi_last_error(const b200reg_multi* M) { return M ? M->err : "null multi"; }
uint64_t b200reg_multi_launch_count(const b200reg_multi* M) {
  uint64_t n = 0;
  if (M) for (b200reg_ctx* c : M->ctx) n += c ? c->launches : 0;
  return n;
}

int32_t b200reg_multi_register_host(b200reg_multi* M, int32_t n, const b200reg_host_cloud* refs, const b200reg_host_cloud* readings,
                                    const float* T_init, float* T_out, b200reg_icp_stats* stats, int32_t* status, float* gather_ms) {
  if (!M) return B200REG_E_INVALID_PARAM;
  if (n < 0 || (n > 0 && (!refs || !readings || !T_init || !T_out))) return multiErr(M, B200REG_E_INVALID_PARAM, "multi_register_host: bad arguments");
  if (gather_ms) *gather_ms = 0.f;
  if (n == 0) return B200REG_OK;
  const int G = M->G;
  NcclApi& N = nccl();
  const size_t slots = ((size_t)n + G - 1) / G;
  TRY(ensureSlots(M, slots));
  // ---- static round-robin of SURVEY 8e: device d owns pairs d, d + G, d + 2G, ...; one host thread per device
  struct Shard {
    std::vector<b200reg_host_cloud> refs, readings;
    std::vector<float> Ti, To;
    std::vector<b200reg_icp_stats> st;
    std::vector<int32_t> code;
    int rc = B200REG_OK;
  };
  std::vector<Shard> sh(G);
  for (int d = 0; d < G; ++d) {
    for (int32_t i = d; i < n; i += G) {
      sh[d].refs.push_back(refs[i]);
      sh[d].readings.push_back(readings[i]);
      sh[d].Ti.insert(sh[d].Ti.end(), T_init + 16 * (size_t)i, T_init + 16 * (size_t)i + 16);
    }
    const size_t k = sh[d].refs.size();
    sh[d].To.assign(16 * k, 0.f);
    sh[d].st.assign(k, b200reg_icp_stats{});
    sh[d].code.assign(k, 0);
  }
  auto work = [&](int d) {
    Shard& s = sh[d];
    const int32_t k = (int32_t)s.refs.size();
    cudaSetDevice(M->devices[d]);
    if (k > 0)
      s.rc = b200reg_batch_register_host(M->batch[d], k, s.refs.data(), s.readings.data(), s.Ti.data(), s.To.data(), s.st.data(), s.code.data());
    // this device's block of the gather: slot j = pair d + j G; unused slots have valid = 0
    PairRecord* h = M->hSend[d];
    memset(h, 0, sizeof(PairRecord) * slots);
    for (int32_t j = 0; j < k; ++j) {
      memcpy(h[j].T, &s.To[16 * (size_t)j], sizeof(float) * 16);
      h[j].status = (float)s.code[j];
      h[j].iterations = (float)s.st[j].iterations;
      h[j].kept = (float)s.st[j].kept_last;
      h[j].tau = s.st[j].last_tau;
      h[j].owner = (float)d;
      h[j].valid = 1.f;
    }
    cudaMemcpyAsync(M->dSend[d], h, sizeof(PairRecord) * slots, cudaMemcpyHostToDevice, M->ctx[d]->stream);
  };
  std::vector<std::thread> threads;
  for (int d = 1; d < G; ++d) threads.emplace_back(work, d);
  work(0);
  for (auto& t : threads) t.join();
  for (int d = 0; d < G; ++d)
    if (sh[d].rc != B200REG_OK) return multiErr(M, sh[d].rc, std::string("multi_register_host: device ") + std::to_string(M->devices[d]) + ": " + b200reg_last_error(M->ctx[d]));
  // ---- the path's only exchange step: ONE all-gather of the records, on the devices' compute streams
  cudaSetDevice(M->devices[0]);
  cudaEventRecord(M->ctx[0]->ev0, M->ctx[0]->stream);
  ncclResult_t r = N.GroupStart();
  for (int d = 0; d < G && r == ncclSuccess; ++d)
    r = N.AllGather(M->dSend[d], M->dRecv[d], sizeof(PairRecord) * slots, ncclChar, M->comm[d], M->ctx[d]->stream);
  const ncclResult_t r2 = N.GroupEnd();
  if (r == ncclSuccess) r = r2;
  if (r != ncclSuccess) return multiErr(M, B200REG_E_CUDA, std::string("multi_register_host: ncclAllGather: ") + N.GetErrorString(r));
  cudaSetDevice(M->devices[0]);
  cudaEventRecord(M->ctx[0]->ev1, M->ctx[0]->stream);
  cudaMemcpyAsync(M->hRecv, M->dRecv[0], sizeof(PairRecord) * slots * G, cudaMemcpyDeviceToHost, M->ctx[0]->stream);
  for (int d = 0; d < G; ++d) {
    cudaSetDevice(M->devices[d]);
    const cudaError_t e = cudaStreamSynchronize(M->ctx[d]->stream);
    if (e != cudaSuccess) return multiErr(M, B200REG_E_CUDA, std::string("multi_register_host: ") + cudaGetErrorString(e));
  }
  if (gather_ms) cudaEventElapsedTime(gather_ms, M->ctx[0]->ev0, M->ctx[0]->ev1);
  // ---- results out of the GATHERED block (device 0's copy): pair i is slot i / G of device i mod G
  int first = B200REG_OK;
  for (int32_t i = 0; i < n; ++i) {
    const int d = i % G;
    const size_t j = (size_t)(i / G);
    const PairRecord& rec = M->hRecv[(size_t)d * slots + j];
    if (rec.valid != 1.f || (int)rec.owner != d) return multiErr(M, B200REG_E_CUDA, "multi_register_host: pair " + std::to_string(i) + " missing from the gather");
    memcpy(T_out + 16 * (size_t)i, rec.T, sizeof(float) * 16);
    const int code = (int)rec.status;
    if (status) status[i] = code;
    if (stats) stats[i] = sh[d].st[j];
    if (code != B200REG_OK && first == B200REG_OK) first = code;
  }
  return status ? B200REG_OK : first;
}

}  // extern "C"
