// ccicp_multi.cu -- the multi-GPU split of the batch behind the C ABI (include/ccicp2d_b200.h, "multi-device").
//
// Scan pairs are independent: device g of G registers the contiguous pair range
// ccicp_shard_range(P, G, g) with its own ccicp_ctx, driven by its own host thread inside ONE process; there is no data-path
// collective.  The only exchange is the gather of the 160-byte records:
//   host path   (ccicp_register_batch_multi)        every shard's D2H copy lands directly in the caller's out[]
//   device path (ccicp_register_batch_multi_device) every device ends up with ALL records in its own d_out: NCCL
//               (ncclBroadcast of each shard, grouped; libnccl.so.2 is loaded lazily with dlopen so the library has no
//               link-time dependency on it) or, when NCCL is unavailable / a device is listed twice, peer copies.
// The reference-side consumer of such a call is graph_slam's loop-closure candidate producer
// (graph_slam/src/graphSlamTools.cpp:72-106 graphSlamGetKNN, :406-425 graphSlamAddEdgeToX; the ICP call at :274).
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <dlfcn.h>
#include <functional>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/ccicp2d_b200.h"

namespace {

// ---- the slice of the NCCL API this file uses, bound at run time --------------------------------------------------
typedef struct ncclComm* ncclComm_t;
typedef int ncclResult_t;
struct Nccl {
  void* lib = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, int /*ncclDataType_t*/, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool load() {
    if (lib) return true;
    lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return false;
    CommInitAll = (decltype(CommInitAll))dlsym(lib, "ncclCommInitAll");
    CommDestroy = (decltype(CommDestroy))dlsym(lib, "ncclCommDestroy");
    GroupStart = (decltype(GroupStart))dlsym(lib, "ncclGroupStart");
    GroupEnd = (decltype(GroupEnd))dlsym(lib, "ncclGroupEnd");
    Broadcast = (decltype(Broadcast))dlsym(lib, "ncclBroadcast");
    GetErrorString = (decltype(GetErrorString))dlsym(lib, "ncclGetErrorString");
    return CommInitAll && CommDestroy && GroupStart && GroupEnd && Broadcast;
  }
};
constexpr int kNcclInt8 = 0;   // ncclInt8 / ncclChar (nccl.h: first enumerator of ncclDataType_t)

// one worker thread per device: runs the closures the calls hand it, in order
struct Worker {
  std::thread th;
  std::mutex mu;
  std::condition_variable cv;
  const std::function<int()>* job = nullptr;
  int rc = 0;
  bool done = true, quit = false;
  void loop() {
    std::unique_lock<std::mutex> lk(mu);
    for (;;) {
      cv.wait(lk, [&] { return quit || job; });
      if (quit) return;
      const std::function<int()>* j = job;
      lk.unlock();
      const int r = (*j)();
      lk.lock();
      rc = r; job = nullptr; done = true;
      cv.notify_all();
    }
  }
  void submit(const std::function<int()>* j) {
    std::lock_guard<std::mutex> lk(mu);
    job = j; done = false;
    cv.notify_all();
  }
  int wait() {
    std::unique_lock<std::mutex> lk(mu);
    cv.wait(lk, [&] { return done; });
    return rc;
  }
};

}  // namespace

struct ccicp_multi {
  int n = 0;
  std::vector<int> dev;
  std::vector<ccicp_ctx*> ctx;
  std::vector<Worker*> worker;
  std::vector<cudaStream_t> gstream;      // gather streams (device path)
  std::vector<cudaEvent_t> gev;
  Nccl nccl;
  std::vector<ncclComm_t> comm;
  bool nccl_ready = false, nccl_tried = false, distinct = true;
  std::string last_err;
  int gather_kind = 0;                    // 0 none yet, 1 nccl, 2 peer copies
};

extern "C" {

void ccicp_shard_range(int32_t n_pairs, int n_devices, int g, int32_t* begin, int32_t* end) {
  // contiguous, balanced: sizes differ by at most one, larger shards first (the rule of slam_b200/sharding.py)
  if (n_devices < 1) n_devices = 1;
  const int32_t base = n_pairs / n_devices, rem = n_pairs % n_devices;
  const int32_t b = g * base + (g < rem ? g : rem);
  if (begin) *begin = b;
  if (end) *end = b + base + (g < rem ? 1 : 0);
}

int ccicp_multi_create(const int* devices, int n_devices, ccicp_multi** out) {
  if (!out) return CCICP_E_INVALID;
  *out = nullptr;
  int avail = 0;
  if (cudaGetDeviceCount(&avail) != cudaSuccess || avail <= 0) { cudaGetLastError(); return CCICP_E_NO_DEVICE; }
  if (n_devices <= 0) { if (devices) return CCICP_E_INVALID; n_devices = avail; }   // NULL, 0: every visible device
  ccicp_multi* m = new (std::nothrow) ccicp_multi();
  if (!m) return CCICP_E_NOMEM;
  m->n = n_devices;
  for (int g = 0; g < n_devices; g++) {
    const int d = devices ? devices[g] : g;
    for (int k : m->dev) if (k == d) m->distinct = false;
    m->dev.push_back(d);
  }
  int rc = CCICP_OK;
  for (int g = 0; g < n_devices && rc == CCICP_OK; g++) {
    ccicp_ctx* c = nullptr;
    rc = ccicp_create(&c, m->dev[g]);
    if (rc == CCICP_OK) m->ctx.push_back(c);
  }
  for (int g = 0; g < n_devices && rc == CCICP_OK; g++) {
    cudaStream_t st = nullptr; cudaEvent_t ev = nullptr;
    if (cudaSetDevice(m->dev[g]) != cudaSuccess || cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); rc = CCICP_E_CUDA; break; }
    m->gstream.push_back(st); m->gev.push_back(ev);
    Worker* w = new (std::nothrow) Worker();
    if (!w) { rc = CCICP_E_NOMEM; break; }
    w->th = std::thread([w] { w->loop(); });
    m->worker.push_back(w);
  }
  if (rc != CCICP_OK) { ccicp_multi_destroy(m); return rc; }
  *out = m;
  return CCICP_OK;
}

int ccicp_multi_destroy(ccicp_multi* m) {
  if (!m) return CCICP_OK;
  for (Worker* w : m->worker) {
    { std::lock_guard<std::mutex> lk(w->mu); w->quit = true; w->cv.notify_all(); }
    if (w->th.joinable()) w->th.join();
    delete w;
  }
  if (m->nccl_ready) for (ncclComm_t c : m->comm) if (c) m->nccl.CommDestroy(c);
  for (size_t g = 0; g < m->gstream.size(); g++) {
    cudaSetDevice(m->dev[g]);
    if (m->gev.size() > g && m->gev[g]) cudaEventDestroy(m->gev[g]);
    if (m->gstream[g]) cudaStreamDestroy(m->gstream[g]);
  }
  for (ccicp_ctx* c : m->ctx) ccicp_destroy(c);
  delete m;
  return CCICP_OK;
}

int ccicp_multi_device_count(const ccicp_multi* m) { return m ? m->n : 0; }
ccicp_ctx* ccicp_multi_ctx(ccicp_multi* m, int g) { return (m && g >= 0 && g < m->n) ? m->ctx[g] : nullptr; }
const char* ccicp_multi_last_error(const ccicp_multi* m) { return m ? m->last_err.c_str() : ""; }
int ccicp_multi_gather_kind(const ccicp_multi* m) { return m ? m->gather_kind : 0; }

// ---- host buffers: same arguments as ccicp_register_batch, the batch is split over the devices ----------------------
int ccicp_register_batch_multi(ccicp_multi* m, int32_t n_pairs, int32_t n_seeds, const double* model_xy,
                               const int32_t* model_off, const double* scene_xy, const int32_t* scene_off,
                               const double* init_Rt, const ccicp_params* params, ccicp_result* out, int32_t* corr_idx) {
  if (!m || !params || n_pairs < 0 || n_seeds < 1) return CCICP_E_INVALID;
  if (n_pairs == 0) return CCICP_OK;
  if (!model_off || !scene_off || !init_Rt || !out) return CCICP_E_INVALID;
  if (corr_idx && n_seeds > 1) return CCICP_E_INVALID;      // seed-major index blocks are not contiguous per shard
  const bool best_of = (params->flags & CCICP_F_BEST_OF_SEEDS) && n_seeds > 1;
  std::vector<std::function<int()>> jobs((size_t)m->n);
  std::vector<std::vector<int32_t>> moff((size_t)m->n), soff((size_t)m->n);
  for (int g = 0; g < m->n; g++) {
    int32_t b, e;
    ccicp_shard_range(n_pairs, m->n, g, &b, &e);
    const int32_t np = e - b;
    if (np <= 0) { jobs[g] = [] { return (int)CCICP_OK; }; continue; }
    const int64_t m0 = model_off[2 * (size_t)b], s0 = scene_off[2 * (size_t)b];
    try {
      moff[g].resize(2 * (size_t)np + 1); soff[g].resize(2 * (size_t)np + 1);
    } catch (const std::bad_alloc&) { return CCICP_E_NOMEM; }
    for (int64_t k = 0; k <= 2 * (int64_t)np; k++) {        // the host call takes offsets that start at 0
      moff[g][k] = (int32_t)(model_off[2 * (size_t)b + k] - m0);
      soff[g][k] = (int32_t)(scene_off[2 * (size_t)b + k] - s0);
    }
    ccicp_ctx* c = m->ctx[g];
    const int32_t* mo = moff[g].data(); const int32_t* so = soff[g].data();
    const double* mx = model_xy ? model_xy + 2 * m0 : nullptr; const double* sx = scene_xy ? scene_xy + 2 * s0 : nullptr;
    const double* in = init_Rt + (size_t)b * n_seeds * 6;
    ccicp_result* o = out + (best_of ? (size_t)b : (size_t)b * n_seeds);
    int32_t* ci = corr_idx ? corr_idx + s0 : nullptr;
    jobs[g] = [=] { return ccicp_register_batch(c, np, n_seeds, mx, mo, sx, so, in, params, o, ci); };
  }
  for (int g = 0; g < m->n; g++) m->worker[g]->submit(&jobs[g]);
  int rc = CCICP_OK;
  for (int g = 0; g < m->n; g++) {
    const int r = m->worker[g]->wait();
    if (r != CCICP_OK && rc == CCICP_OK) { rc = r; m->last_err = std::string("device ") + std::to_string(m->dev[g]) + ": " + ccicp_last_cuda_error(m->ctx[g]); }
  }
  return rc;
}

// ---- device-resident shards -----------------------------------------------------------------------------------------
// Device g holds ITS shard (pairs [b_g, e_g) of the n_pairs, ccicp_shard_range) as a CSR of its own (offsets from 0) and
// a record buffer d_out[g] for ALL n_pairs (x n_seeds unless best-of) records.  Every device registers its shard into
// its slice of d_out[g]; the slices are then exchanged so that each d_out[g] holds every record.  Asynchronous: returns
// after enqueueing; ccicp_multi_sync waits for all devices.
int ccicp_register_batch_multi_device(ccicp_multi* m, int32_t n_pairs, int32_t n_seeds, const double* const* d_model_xy,
                                      const int32_t* const* d_model_off, const double* const* d_scene_xy,
                                      const int32_t* const* d_scene_off, const double* const* d_init_Rt,
                                      const ccicp_params* params, ccicp_result* const* d_out, const int32_t* max_model,
                                      const int32_t* max_scene, const int32_t* max_cells_hint) {
  if (!m || !params || n_pairs < 0 || n_seeds < 1 || !d_out || !max_model || !max_scene) return CCICP_E_INVALID;
  if (n_pairs == 0) return CCICP_OK;
  const bool best_of = (params->flags & CCICP_F_BEST_OF_SEEDS) && n_seeds > 1;
  const size_t per_pair = best_of ? 1 : (size_t)n_seeds;
  // gather transport, chosen once
  if (m->gather_kind == 0) {
    if (m->n > 1 && m->distinct && !m->nccl_tried) {
      m->nccl_tried = true;
      if (m->nccl.load()) {
        m->comm.assign((size_t)m->n, nullptr);
        const ncclResult_t r = m->nccl.CommInitAll(m->comm.data(), m->n, m->dev.data());
        if (r == 0) m->nccl_ready = true;
        else m->last_err = std::string("ncclCommInitAll: ") + (m->nccl.GetErrorString ? m->nccl.GetErrorString(r) : "failed");
      } else {
        m->last_err = "libnccl.so.2 not found: records are exchanged by peer copies";
      }
    }
    m->gather_kind = m->nccl_ready ? 1 : 2;
    if (m->gather_kind == 2 && m->distinct) {
      for (int g = 0; g < m->n; g++) {
        cudaSetDevice(m->dev[g]);
        for (int h = 0; h < m->n; h++) if (h != g) { int can = 0; cudaDeviceCanAccessPeer(&can, m->dev[g], m->dev[h]); if (can) cudaDeviceEnablePeerAccess(m->dev[h], 0); }
        cudaGetLastError();
      }
    }
  }
  // 1. every device registers its shard on its gather stream
  for (int g = 0; g < m->n; g++) {
    int32_t b, e;
    ccicp_shard_range(n_pairs, m->n, g, &b, &e);
    if (e > b) {
      const int rc = ccicp_register_batch_device_ex(m->ctx[g], e - b, n_seeds, d_model_xy[g], d_model_off[g], d_scene_xy[g],
                                                    d_scene_off[g], d_init_Rt[g], params, d_out[g] + (size_t)b * per_pair, nullptr,
                                                    max_model[g], max_scene[g], max_cells_hint ? max_cells_hint[g] : 0,
                                                    (void*)m->gstream[g]);
      if (rc != CCICP_OK) { m->last_err = std::string("device ") + std::to_string(m->dev[g]) + ": " + ccicp_last_cuda_error(m->ctx[g]); return rc; }
    }
    if (cudaSetDevice(m->dev[g]) != cudaSuccess || cudaEventRecord(m->gev[g], m->gstream[g]) != cudaSuccess) { cudaGetLastError(); return CCICP_E_CUDA; }
  }
  if (m->n == 1) return CCICP_OK;
  // 2. exchange the slices
  if (m->gather_kind == 1) {
    ncclResult_t r = m->nccl.GroupStart();
    for (int root = 0; root < m->n && r == 0; root++) {
      int32_t b, e;
      ccicp_shard_range(n_pairs, m->n, root, &b, &e);
      if (e <= b) continue;
      const size_t bytes = (size_t)(e - b) * per_pair * sizeof(ccicp_result);
      for (int g = 0; g < m->n && r == 0; g++) {
        void* p = (void*)(d_out[g] + (size_t)b * per_pair);
        r = m->nccl.Broadcast(p, p, bytes, kNcclInt8, root, m->comm[g], m->gstream[g]);
      }
    }
    const ncclResult_t r2 = m->nccl.GroupEnd();
    if (r != 0 || r2 != 0) { m->last_err = std::string("nccl broadcast: ") + (m->nccl.GetErrorString ? m->nccl.GetErrorString(r ? r : r2) : "failed"); return CCICP_E_CUDA; }
  } else {
    for (int g = 0; g < m->n; g++) {          // device g pulls every other shard once that shard's kernel is done
      if (cudaSetDevice(m->dev[g]) != cudaSuccess) { cudaGetLastError(); return CCICP_E_CUDA; }
      for (int h = 0; h < m->n; h++) {
        if (h == g) continue;
        int32_t b, e;
        ccicp_shard_range(n_pairs, m->n, h, &b, &e);
        if (e <= b) continue;
        const size_t bytes = (size_t)(e - b) * per_pair * sizeof(ccicp_result);
        cudaError_t er = cudaStreamWaitEvent(m->gstream[g], m->gev[h], 0);
        if (er == cudaSuccess)
          er = cudaMemcpyPeerAsync(d_out[g] + (size_t)b * per_pair, m->dev[g], d_out[h] + (size_t)b * per_pair, m->dev[h], bytes, m->gstream[g]);
        if (er != cudaSuccess) { m->last_err = std::string("peer copy: ") + cudaGetErrorString(er); cudaGetLastError(); return CCICP_E_CUDA; }
      }
    }
  }
  return CCICP_OK;
}

int ccicp_multi_sync(ccicp_multi* m) {
  if (!m) return CCICP_E_INVALID;
  for (int g = 0; g < m->n; g++) {
    if (cudaSetDevice(m->dev[g]) != cudaSuccess || cudaStreamSynchronize(m->gstream[g]) != cudaSuccess) {
      m->last_err = "cudaStreamSynchronize failed"; cudaGetLastError(); return CCICP_E_CUDA;
    }
  }
  return CCICP_OK;
}

}  // extern "C"
