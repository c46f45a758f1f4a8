// C ABI (include/orb_b200.h), matcher / vocabulary / frame-geometry part: brute-force kNN, DescriptorDistance batches, guided
// search (SearchByProjection, SearchByBoW), ComputeBoW, UndistortKeyPoints, ComputeDistinctiveDescriptors. Host logic only;
// per-device scratch behind a mutex + an event chain across streams (WsOrder), so these entry points are callable from any thread and on any
// stream. No CPU fallback.
#include <dlfcn.h>

#include "orb_host.h"

extern "C" {

// ---- matcher ------------------------------------------------------------------------------------------------
struct KnnWorkspace {   // per-device scratch, grown on demand; guarded by a mutex (calls on one device serialise their setup)
  std::mutex mu;
  DevBuf partial, q, db, out;
  DevBuf sh_local, sh_gathered;   // orb_hamming_knn_sharded: this rank's records [Q], all ranks' records [nranks][Q]
  // orb_search_by_projection: sorted keypoints + cell starts + per-query top-2, and the staging copies of host arguments
  DevBuf g_bq, g_bnq, g_stage[17];   // orb_search_by_bow: device-built queries, staging of host arguments
  DevBuf g_items, g_cells, g_lists, g_lcnt, g_kps, g_desc, g_ur, g_claim, g_nk, g_q, g_qd, g_nq, g_assigned, g_qmatch, g_nm;
  int num_sms = 0;
  cudaEvent_t ev_last = nullptr;   // recorded on the caller's stream at the end of every call that touched the scratch (see WsOrder)
};
static KnnWorkspace g_knn_ws[16];

// The scratch is shared by all streams of a device. Device-mode calls return while their kernels are still queued, so the mutex alone would
// let a call on ANOTHER stream overwrite lists / partials that are still being read (ORBmatcher is used by Tracking, LocalMapping and
// LoopClosing at once). Every entry point that touches the scratch therefore makes its stream wait for the previous user's event first and
// records its own event last — both while holding the mutex, so the chain of users is totally ordered on the GPU.
struct WsOrder {
  cudaEvent_t* ev;
  cudaStream_t st;
  WsOrder(cudaEvent_t* e, cudaStream_t s) : ev(e), st(s) {
    if (!*ev && cudaEventCreateWithFlags(ev, cudaEventDisableTiming) != cudaSuccess) { *ev = nullptr; cudaGetLastError(); }
    if (*ev) cudaStreamWaitEvent(st, *ev, 0);   // no-op for an event that was never recorded or was recorded on `st` itself
  }
  ~WsOrder() {
    if (*ev) cudaEventRecord(*ev, st);
  }
};

static int current_device_ws(KnnWorkspace** ws) {
  int dev = 0;
  CU(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 16) return fail(ORB_E_BADARG, "device index %d unsupported", dev);
  *ws = &g_knn_ws[dev];
  if ((*ws)->num_sms == 0) {
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, dev));
    (*ws)->num_sms = prop.multiProcessorCount;
  }
  return ORB_OK;
}

int orb_hamming_knn(const uint8_t* query, int q_mem, int Q, const uint8_t* db, int db_mem, int64_t D, int64_t index_base,
                    int sentinel, orb_knn2* out, int out_mem, void* stream) {
  if (!query || !out || Q <= 0 || D < 0 || (D > 0 && !db) || sentinel < 1) return fail(ORB_E_BADARG, "bad argument");
  if (index_base < 0 || index_base + D >= (1LL << 40)) return fail(ORB_E_BADARG, "index range exceeds 2^40");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  WsOrder order(&ws->ev_last, st);
  const uint8_t* dq = query;
  const uint8_t* ddb = db;
  orb_knn2* dout = out;
  const bool any_host = q_mem != ORB_MEM_DEVICE || db_mem != ORB_MEM_DEVICE || out_mem != ORB_MEM_DEVICE;
  if (q_mem != ORB_MEM_DEVICE || ((uintptr_t)query & 15)) {
    if ((rc = ws->q.ensure((size_t)Q * 32)) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->q.p, query, (size_t)Q * 32, q_mem == ORB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    dq = ws->q.as<uint8_t>();
  }
  if (D > 0 && (db_mem != ORB_MEM_DEVICE || ((uintptr_t)db & 15))) {
    if ((rc = ws->db.ensure((size_t)D * 32)) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->db.p, db, (size_t)D * 32, db_mem == ORB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    ddb = ws->db.as<uint8_t>();
  }
  if (out_mem != ORB_MEM_DEVICE) {
    if ((rc = ws->out.ensure((size_t)Q * sizeof(orb_knn2))) != ORB_OK) return rc;
    dout = ws->out.as<orb_knn2>();
  }
  const int nchunks = orb_knn_num_chunks(Q, D, ws->num_sms);
  if ((rc = ws->partial.ensure((size_t)nchunks * Q * 2 * sizeof(unsigned long long))) != ORB_OK) return rc;
  orb_launch_knn(dq, Q, ddb, D, index_base, sentinel, ws->partial.as<unsigned long long>(), nchunks, (OrbKnn2*)dout, st);
  CU(cudaGetLastError());
  if (out_mem != ORB_MEM_DEVICE) CU(cudaMemcpyAsync(out, dout, (size_t)Q * sizeof(orb_knn2), cudaMemcpyDeviceToHost, st));
  if (any_host) CU(cudaStreamSynchronize(st));
  return ORB_OK;
}

int orb_hamming_knn_merge(const orb_knn2* shards, int nshards, int Q, int sentinel, orb_knn2* out, void* stream) {
  if (!shards || !out || nshards < 1 || Q <= 0) return fail(ORB_E_BADARG, "bad argument");
  orb_launch_knn_merge_shards((const OrbKnn2*)shards, nshards, Q, sentinel, (OrbKnn2*)out, (cudaStream_t)stream);
  CU(cudaGetLastError());
  return ORB_OK;
}

// ---- NCCL (dlopen'ed: no header and no link-time dependency) ---------------------------------------------------------------------------
// dlopen("libnccl.so.2") returns the NCCL the process has ALREADY loaded when there is one (glibc matches loaded objects by SONAME), so a
// communicator created by the caller's NCCL - torch.distributed's included - is used with the library that made it.
struct OrbNccl {
  typedef struct ncclComm* comm_t;
  struct unique_id { char internal[128]; };   // ncclUniqueId (NCCL_UNIQUE_ID_BYTES = 128)
  int (*AllGather)(const void*, void*, size_t, int, comm_t, cudaStream_t) = nullptr;
  int (*CommCount)(comm_t, int*) = nullptr;
  int (*CommUserRank)(comm_t, int*) = nullptr;
  int (*GetUniqueId)(unique_id*) = nullptr;
  int (*CommInitRank)(comm_t*, int, unique_id, int) = nullptr;
  int (*CommDestroy)(comm_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool ok = false;
};
static OrbNccl* orb_nccl() {
  static OrbNccl N;
  static std::once_flag once;
  std::call_once(once, []() {
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return;
    *(void**)&N.AllGather = dlsym(h, "ncclAllGather");
    *(void**)&N.CommCount = dlsym(h, "ncclCommCount");
    *(void**)&N.CommUserRank = dlsym(h, "ncclCommUserRank");
    *(void**)&N.GetUniqueId = dlsym(h, "ncclGetUniqueId");
    *(void**)&N.CommInitRank = dlsym(h, "ncclCommInitRank");
    *(void**)&N.CommDestroy = dlsym(h, "ncclCommDestroy");
    *(void**)&N.GetErrorString = dlsym(h, "ncclGetErrorString");
    N.ok = N.AllGather && N.CommCount && N.CommUserRank && N.GetUniqueId && N.CommInitRank && N.CommDestroy && N.GetErrorString;
  });
  return N.ok ? &N : nullptr;
}
#define NC(call)                                                                                              \
  do {                                                                                                        \
    int r_ = (call);                                                                                          \
    if (r_ != 0) return fail(ORB_E_CUDA, "%s: NCCL error %d (%s)", #call, r_, N->GetErrorString(r_));         \
  } while (0)

int orb_nccl_unique_id(uint8_t* id128) {
  if (!id128) return fail(ORB_E_BADARG, "null argument");
  OrbNccl* N = orb_nccl();
  if (!N) return fail(ORB_E_CUDA, "libnccl.so.2 not found (dlopen)");
  OrbNccl::unique_id id;
  NC(N->GetUniqueId(&id));
  memcpy(id128, id.internal, 128);
  return ORB_OK;
}
int orb_nccl_comm_create(const uint8_t* id128, int nranks, int rank, void** comm) {
  if (!id128 || !comm || nranks < 1 || rank < 0 || rank >= nranks) return fail(ORB_E_BADARG, "bad argument");
  OrbNccl* N = orb_nccl();
  if (!N) return fail(ORB_E_CUDA, "libnccl.so.2 not found (dlopen)");
  OrbNccl::unique_id id;
  memcpy(id.internal, id128, 128);
  OrbNccl::comm_t c = nullptr;
  NC(N->CommInitRank(&c, nranks, id, rank));   // on the caller's current device
  *comm = c;
  return ORB_OK;
}
void orb_nccl_comm_destroy(void* comm) {
  OrbNccl* N = orb_nccl();
  if (N && comm) N->CommDestroy((OrbNccl::comm_t)comm);
}

int orb_hamming_knn_sharded(const uint8_t* query, int q_mem, int Q, const uint8_t* db_shard, int db_mem, int64_t D_shard, int64_t index_base,
                            int sentinel, orb_knn2* out, int out_mem, void* nccl_comm, void* stream) {
  if (!query || !out || !nccl_comm || Q <= 0 || D_shard < 0 || (D_shard > 0 && !db_shard) || sentinel < 1) return fail(ORB_E_BADARG, "bad argument");
  if (index_base < 0 || index_base + D_shard >= (1LL << 40)) return fail(ORB_E_BADARG, "index range exceeds 2^40");
  OrbNccl* N = orb_nccl();
  if (!N) return fail(ORB_E_CUDA, "libnccl.so.2 not found (dlopen)");
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  OrbNccl::comm_t comm = (OrbNccl::comm_t)nccl_comm;
  int nranks = 0;
  NC(N->CommCount(comm, &nranks));
  if (nranks < 1) return fail(ORB_E_BADARG, "communicator with %d ranks", nranks);
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  WsOrder order(&ws->ev_last, st);
  const uint8_t* dq = query;
  const uint8_t* ddb = db_shard;
  orb_knn2* dout = out;
  const bool any_host = q_mem != ORB_MEM_DEVICE || db_mem != ORB_MEM_DEVICE || out_mem != ORB_MEM_DEVICE;
  if (q_mem != ORB_MEM_DEVICE || ((uintptr_t)query & 15)) {
    if ((rc = ws->q.ensure((size_t)Q * 32)) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->q.p, query, (size_t)Q * 32, q_mem == ORB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    dq = ws->q.as<uint8_t>();
  }
  if (D_shard > 0 && (db_mem != ORB_MEM_DEVICE || ((uintptr_t)db_shard & 15))) {
    if ((rc = ws->db.ensure((size_t)D_shard * 32)) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->db.p, db_shard, (size_t)D_shard * 32, db_mem == ORB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    ddb = ws->db.as<uint8_t>();
  }
  const size_t rec = (size_t)Q * sizeof(orb_knn2);
  if (out_mem != ORB_MEM_DEVICE) {
    if ((rc = ws->out.ensure(rec)) != ORB_OK) return rc;
    dout = ws->out.as<orb_knn2>();
  }
  if ((rc = ws->sh_local.ensure(rec)) != ORB_OK) return rc;
  if ((rc = ws->sh_gathered.ensure(rec * nranks)) != ORB_OK) return rc;
  const int nchunks = orb_knn_num_chunks(Q, D_shard, ws->num_sms);
  if ((rc = ws->partial.ensure((size_t)nchunks * Q * 2 * sizeof(unsigned long long))) != ORB_OK) return rc;
  // 1. this rank's shard -> records with GLOBAL row indices; 2. one all-gather of Q x 24 bytes per rank (ascending rank = ascending rows);
  // 3. the two lexicographically smallest (distance, index) of all shards = the sequential scan over the concatenated database
  orb_launch_knn(dq, Q, ddb, D_shard, index_base, sentinel, ws->partial.as<unsigned long long>(), nchunks, ws->sh_local.as<OrbKnn2>(), st);
  CU(cudaGetLastError());
  NC(N->AllGather(ws->sh_local.p, ws->sh_gathered.p, rec, /*ncclInt8*/ 0, comm, st));
  orb_launch_knn_merge_shards(ws->sh_gathered.as<OrbKnn2>(), nranks, Q, sentinel, (OrbKnn2*)dout, st);
  CU(cudaGetLastError());
  if (out_mem != ORB_MEM_DEVICE) CU(cudaMemcpyAsync(out, dout, rec, cudaMemcpyDeviceToHost, st));
  if (any_host) CU(cudaStreamSynchronize(st));
  return ORB_OK;
}

// ---- sharded scan with peer stores (no NCCL) -------------------------------------------------------------------------------------------------
// Mailbox of a rank, one cudaMalloc (CUDA IPC exports whole allocations): [256 B: nranks epoch flags][256 B: ticket, error flag][2][nranks][max_q] x 16 B
struct orb_knn_peer {
  int device = 0, rank = 0, nranks = 0, max_q = 0;
  void* box_mem = nullptr;                              // this rank's mailbox
  void* peer_mem[ORB_KNN_PEER_MAX_RANKS] = {};          // every rank's mailbox as mapped in this process (own entry = box_mem)
  struct { unsigned long long* slots[ORB_KNN_PEER_MAX_RANKS]; unsigned* flags[ORB_KNN_PEER_MAX_RANKS]; } box = {};
  unsigned epoch = 0;
  bool connected = false;
};
static size_t peer_box_bytes(int nranks, int max_q) { return 512 + (size_t)2 * nranks * max_q * 16; }

int orb_knn_peer_create(int rank, int nranks, int max_q, orb_knn_peer** out) {
  if (!out || nranks < 1 || nranks > ORB_KNN_PEER_MAX_RANKS || rank < 0 || rank >= nranks || max_q < 1) return fail(ORB_E_BADARG, "bad argument");
  *out = nullptr;
  orb_knn_peer* p = new orb_knn_peer();
  p->rank = rank; p->nranks = nranks; p->max_q = max_q;
  if (cudaGetDevice(&p->device) != cudaSuccess) { delete p; cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  const size_t bytes = peer_box_bytes(nranks, max_q);
  cudaError_t e = cudaMalloc(&p->box_mem, bytes);
  if (e == cudaSuccess) e = cudaMemset(p->box_mem, 0, bytes);
  if (e != cudaSuccess) { if (p->box_mem) cudaFree(p->box_mem); delete p; return fail(ORB_E_CUDA, "mailbox: %s", cudaGetErrorString(e)); }
  p->peer_mem[rank] = p->box_mem;
  *out = p;
  return ORB_OK;
}

int orb_knn_peer_export(orb_knn_peer* p, uint8_t* handle64) {
  if (!p || !handle64) return fail(ORB_E_BADARG, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == ORB_KNN_PEER_HANDLE_BYTES, "cudaIpcMemHandle_t is 64 bytes");
  ORB_ON_DEVICE(p->device);
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, p->box_mem));
  memcpy(handle64, &h, sizeof(h));
  return ORB_OK;
}

int orb_knn_peer_connect(orb_knn_peer* p, const uint8_t* handles) {
  if (!p || (!handles && p->nranks > 1)) return fail(ORB_E_BADARG, "null argument");
  if (p->connected) return fail(ORB_E_BADARG, "already connected");
  ORB_ON_DEVICE(p->device);
  for (int r = 0; r < p->nranks; ++r) {
    if (r != p->rank) {
      cudaIpcMemHandle_t h;
      memcpy(&h, handles + (size_t)r * ORB_KNN_PEER_HANDLE_BYTES, sizeof(h));
      cudaError_t e = cudaIpcOpenMemHandle(&p->peer_mem[r], h, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) { cudaGetLastError(); return fail(ORB_E_CUDA, "cudaIpcOpenMemHandle(rank %d): %s (peer access between the GPUs is required)", r, cudaGetErrorString(e)); }
    }
    p->box.flags[r] = reinterpret_cast<unsigned*>(p->peer_mem[r]);
    p->box.slots[r] = reinterpret_cast<unsigned long long*>(static_cast<uint8_t*>(p->peer_mem[r]) + 512);
  }
  p->connected = true;
  return ORB_OK;
}

static int peer_error(orb_knn_peer* p, cudaStream_t st) {   // stream already synchronised
  int flag = 0;
  CU(cudaMemcpyAsync(&flag, static_cast<uint8_t*>(p->box_mem) + 256 + 4, sizeof(int), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (flag) {
    cudaMemsetAsync(static_cast<uint8_t*>(p->box_mem) + 256 + 4, 0, sizeof(int), st);
    return fail(ORB_E_CAPACITY, "orb_hamming_knn_peer: a peer's flag did not arrive (flags %d: 32 = wait timed out; did every rank make the call?)", flag);
  }
  return ORB_OK;
}

int orb_knn_peer_check(orb_knn_peer* p, void* stream) {
  if (!p) return fail(ORB_E_BADARG, "null argument");
  ORB_ON_DEVICE(p->device);
  CU(cudaStreamSynchronize((cudaStream_t)stream));
  return peer_error(p, (cudaStream_t)stream);
}

int orb_hamming_knn_peer(orb_knn_peer* p, const uint8_t* query, int q_mem, int Q, const uint8_t* db_shard, int db_mem, int64_t D_shard,
                         int64_t index_base, int sentinel, orb_knn2* out, int out_mem, void* stream) {
  if (!p || !query || !out || Q <= 0 || D_shard < 0 || (D_shard > 0 && !db_shard) || sentinel < 1) return fail(ORB_E_BADARG, "bad argument");
  if (!p->connected) return fail(ORB_E_BADARG, "orb_knn_peer_connect has not been called");
  if (Q > p->max_q) return fail(ORB_E_CAPACITY, "Q = %d exceeds the mailbox capacity max_q = %d", Q, p->max_q);
  if (index_base < 0 || index_base + D_shard >= (1LL << 40)) return fail(ORB_E_BADARG, "index range exceeds 2^40");
  ORB_ON_DEVICE(p->device);
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  WsOrder order(&ws->ev_last, st);
  const uint8_t* dq = query;
  const uint8_t* ddb = db_shard;
  orb_knn2* dout = out;
  const bool any_host = q_mem != ORB_MEM_DEVICE || db_mem != ORB_MEM_DEVICE || out_mem != ORB_MEM_DEVICE;
  if (q_mem != ORB_MEM_DEVICE || ((uintptr_t)query & 15)) {
    if ((rc = ws->q.ensure((size_t)Q * 32)) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->q.p, query, (size_t)Q * 32, q_mem == ORB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    dq = ws->q.as<uint8_t>();
  }
  if (D_shard > 0 && (db_mem != ORB_MEM_DEVICE || ((uintptr_t)db_shard & 15))) {
    if ((rc = ws->db.ensure((size_t)D_shard * 32)) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->db.p, db_shard, (size_t)D_shard * 32, db_mem == ORB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    ddb = ws->db.as<uint8_t>();
  }
  const size_t rec = (size_t)Q * sizeof(orb_knn2);
  if (out_mem != ORB_MEM_DEVICE) {
    if ((rc = ws->out.ensure(rec)) != ORB_OK) return rc;
    dout = ws->out.as<orb_knn2>();
  }
  const int nchunks = orb_knn_num_chunks(Q, D_shard, ws->num_sms);
  if ((rc = ws->partial.ensure((size_t)nchunks * Q * 2 * sizeof(unsigned long long))) != ORB_OK) return rc;
  ++p->epoch;   // the same sequence on every rank: the calls are collective
  uint8_t* mine = static_cast<uint8_t*>(p->box_mem);
  orb_launch_knn_peer(dq, Q, ddb, D_shard, index_base, sentinel, ws->partial.as<unsigned long long>(), nchunks, &p->box, p->rank, p->nranks, p->max_q,
                      p->epoch, reinterpret_cast<unsigned*>(mine + 256), (OrbKnn2*)dout, reinterpret_cast<int*>(mine + 256 + 4), st);
  CU(cudaGetLastError());
  if (out_mem != ORB_MEM_DEVICE) CU(cudaMemcpyAsync(out, dout, rec, cudaMemcpyDeviceToHost, st));
  if (any_host) {
    CU(cudaStreamSynchronize(st));
    return peer_error(p, st);
  }
  return ORB_OK;
}

void orb_knn_peer_destroy(orb_knn_peer* p) {
  if (!p) return;
  int prev = -1;
  cudaGetDevice(&prev);
  cudaSetDevice(p->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < p->nranks; ++r)
    if (r != p->rank && p->peer_mem[r]) cudaIpcCloseMemHandle(p->peer_mem[r]);
  if (p->box_mem) cudaFree(p->box_mem);
  if (prev >= 0) cudaSetDevice(prev);
  cudaGetLastError();
  delete p;
}

int orb_descriptor_distance_batch(const uint8_t* a, const uint8_t* b, int64_t n, int32_t* out, void* stream) {
  if (!a || !b || !out || n < 0) return fail(ORB_E_BADARG, "bad argument");
  if (((uintptr_t)a & 15) || ((uintptr_t)b & 15)) return fail(ORB_E_BADARG, "descriptor arrays must be 16-byte aligned");
  if (n == 0) return ORB_OK;
  orb_launch_distance_batch(a, b, n, out, (cudaStream_t)stream);
  CU(cudaGetLastError());
  return ORB_OK;
}

int orb_descriptor_distance_host(const uint8_t* a, const uint8_t* b, int64_t n, int32_t* out) {
  if (!a || !b || !out || n < 0) return fail(ORB_E_BADARG, "bad argument");
  if (n == 0) return ORB_OK;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  std::lock_guard<std::mutex> lk(ws->mu);
  WsOrder order(&ws->ev_last, (cudaStream_t)0);
  if ((rc = ws->q.ensure((size_t)n * 32)) != ORB_OK) return rc;
  if ((rc = ws->db.ensure((size_t)n * 32)) != ORB_OK) return rc;
  if ((rc = ws->out.ensure((size_t)n * sizeof(int))) != ORB_OK) return rc;
  CU(cudaMemcpyAsync(ws->q.p, a, (size_t)n * 32, cudaMemcpyHostToDevice, 0));
  CU(cudaMemcpyAsync(ws->db.p, b, (size_t)n * 32, cudaMemcpyHostToDevice, 0));
  orb_launch_distance_batch(ws->q.as<uint8_t>(), ws->db.as<uint8_t>(), n, ws->out.as<int>(), 0);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(out, ws->out.p, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, 0));
  CU(cudaStreamSynchronize(0));
  return ORB_OK;
}

int64_t orb_matcher_launch_count(void) { return orb_knn_launches(); }

// ---- vocabulary / ComputeBoW ------------------------------------------------------------------------------------------------

struct orb_vocabulary {
  int device = 0, k = 0, L = 0, n_nodes = 0, n_words = 0;
  DevBuf child_start, child_id, child_desc, is_leaf, word_of, word_weight;
  DevBuf s_desc, s_n, s_word, s_node, s_bw, s_bv, s_bc, s_fn, s_fs, s_ff, s_fc;   // staging for host-memory calls
  std::mutex mu;
  cudaEvent_t ev_last = nullptr;   // orders the users of the s_* scratch across streams (WsOrder)
};

int orb_vocabulary_create(int k, int L, int n_nodes, const int32_t* parent, const uint8_t* is_leaf, const uint8_t* desc,
                          const double* weight, int device, orb_vocabulary** out) {
  if (!out) return fail(ORB_E_BADARG, "null argument");
  *out = nullptr;
  if (!parent || !is_leaf || !desc || !weight || k < 1 || L < 1 || n_nodes < 2) return fail(ORB_E_BADARG, "bad vocabulary arguments");
  std::vector<int> cnt(n_nodes + 1, 0);
  for (int i = 1; i < n_nodes; ++i) {
    if (parent[i] < 0 || parent[i] >= n_nodes || parent[i] == i) return fail(ORB_E_BADARG, "node %d: parent %d out of range", i, parent[i]);
    if (is_leaf[parent[i]] && parent[i] != 0) return fail(ORB_E_BADARG, "node %d hangs below leaf %d", i, parent[i]);
    ++cnt[parent[i] + 1];
  }
  for (int i = 1; i < n_nodes; ++i)
    if (!is_leaf[i] && cnt[i + 1] == 0) return fail(ORB_E_BADARG, "inner node %d has no children", i);
  if (cnt[1] == 0) return fail(ORB_E_BADARG, "root has no children");
  std::vector<int> start(n_nodes + 1, 0);
  for (int i = 0; i < n_nodes; ++i) start[i + 1] = start[i] + cnt[i + 1];
  const int slots = start[n_nodes];   // == n_nodes - 1
  std::vector<int> cursor(start.begin(), start.end() - 1), child_id(slots), word_of(n_nodes, -1);
  std::vector<uint8_t> child_desc((size_t)slots * 32), leaf(n_nodes, 0);
  std::vector<double> word_weight;
  for (int i = 1; i < n_nodes; ++i) {   // file order: children in push_back order, words numbered as the leaves appear
    const int s = cursor[parent[i]]++;
    child_id[s] = i;
    memcpy(&child_desc[(size_t)s * 32], desc + (size_t)i * 32, 32);
    leaf[i] = is_leaf[i] ? 1 : 0;
    if (leaf[i]) { word_of[i] = (int)word_weight.size(); word_weight.push_back(weight[i]); }
  }
  // a walk deeper than n_nodes levels would mean a cycle; parents precede nothing in particular, so check by depth
  for (int i = 1; i < n_nodes; ++i) {
    int d = 0, p = i;
    while (p != 0 && d <= n_nodes) { p = parent[p]; ++d; }
    if (p != 0) return fail(ORB_E_BADARG, "node %d is not connected to the root", i);
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  ORB_ON_DEVICE(device);
  orb_vocabulary* v = new orb_vocabulary;
  v->device = device; v->k = k; v->L = L; v->n_nodes = n_nodes; v->n_words = (int)word_weight.size();
  struct Up { DevBuf* b; const void* src; size_t bytes; } ups[] = {
      {&v->child_start, start.data(), start.size() * sizeof(int)}, {&v->child_id, child_id.data(), child_id.size() * sizeof(int)},
      {&v->child_desc, child_desc.data(), child_desc.size()},      {&v->is_leaf, leaf.data(), leaf.size()},
      {&v->word_of, word_of.data(), word_of.size() * sizeof(int)}, {&v->word_weight, word_weight.data(), word_weight.size() * sizeof(double)}};
  for (const Up& u : ups) {
    int rc = u.b->ensure(std::max<size_t>(u.bytes, 16));
    if (rc == ORB_OK && cudaMemcpy(u.b->p, u.src, u.bytes, cudaMemcpyHostToDevice) != cudaSuccess) rc = fail(ORB_E_CUDA, "vocabulary upload failed");
    if (rc != ORB_OK) { orb_vocabulary_destroy(v); return rc; }
  }
  *out = v;
  return ORB_OK;
}

void orb_vocabulary_destroy(orb_vocabulary* v) {
  if (!v) return;
  OrbDeviceGuard guard(v->device);
  DevBuf* bufs[] = {&v->child_start, &v->child_id, &v->child_desc, &v->is_leaf, &v->word_of, &v->word_weight, &v->s_desc, &v->s_n, &v->s_word,
                    &v->s_node, &v->s_bw, &v->s_bv, &v->s_bc, &v->s_fn, &v->s_fs, &v->s_ff, &v->s_fc};
  for (DevBuf* b : bufs) b->release();
  if (v->ev_last) cudaEventDestroy(v->ev_last);
  delete v;
}

int orb_vocabulary_words(const orb_vocabulary* v) { return v ? v->n_words : 0; }

int orb_bow_transform(orb_vocabulary* v, int mem, const uint8_t* desc, const int32_t* n, int cap, int frames, int levelsup, int32_t* word,
                      int32_t* node, int32_t* bow_word, double* bow_value, int32_t* bow_count, int32_t* fv_node, int32_t* fv_start,
                      int32_t* fv_feature, int32_t* fv_count, void* stream) {
  if (!v || !desc || !n || !bow_word || !bow_value || !bow_count || !fv_node || !fv_start || !fv_feature || !fv_count)
    return fail(ORB_E_BADARG, "null argument");
  if (cap <= 0 || cap > 16384 || frames <= 0 || levelsup < 0) return fail(ORB_E_BADARG, "bad size (0 < cap <= 16384, frames > 0, levelsup >= 0)");
  std::lock_guard<std::mutex> lk(v->mu);
  ORB_ON_DEVICE(v->device);
  cudaStream_t st = (cudaStream_t)stream;
  WsOrder order(&v->ev_last, st);
  const size_t F = (size_t)frames;
  const bool host = mem != ORB_MEM_DEVICE;
  int rc;
  const uint8_t* d_desc = desc; const int* d_n = n;
  int *d_word = word, *d_node = node, *d_bw = bow_word, *d_bc = bow_count, *d_fn = fv_node, *d_fs = fv_start, *d_ff = fv_feature, *d_fc = fv_count;
  double* d_bv = bow_value;
  if (host) {
    struct S { DevBuf* b; size_t bytes; } st_bufs[] = {{&v->s_desc, F * cap * 32}, {&v->s_n, F * 4}, {&v->s_bw, F * cap * 4}, {&v->s_bv, F * cap * 8},
                                                      {&v->s_bc, F * 4}, {&v->s_fn, F * cap * 4}, {&v->s_fs, F * (cap + 1) * 4},
                                                      {&v->s_ff, F * cap * 4}, {&v->s_fc, F * 4}};
    for (const S& s : st_bufs) if ((rc = s.b->ensure(s.bytes)) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(v->s_desc.p, desc, F * cap * 32, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(v->s_n.p, n, F * 4, cudaMemcpyHostToDevice, st));
    d_desc = v->s_desc.as<uint8_t>(); d_n = v->s_n.as<int>();
    d_bw = v->s_bw.as<int>(); d_bv = v->s_bv.as<double>(); d_bc = v->s_bc.as<int>();
    d_fn = v->s_fn.as<int>(); d_fs = v->s_fs.as<int>(); d_ff = v->s_ff.as<int>(); d_fc = v->s_fc.as<int>();
  } else if ((uintptr_t)desc & 15) {
    return fail(ORB_E_BADARG, "descriptor array must be 16-byte aligned device memory");
  }
  if (host || !word) { if ((rc = v->s_word.ensure(F * cap * 4)) != ORB_OK) return rc; d_word = v->s_word.as<int>(); }
  if (host || !node) { if ((rc = v->s_node.ensure(F * cap * 4)) != ORB_OK) return rc; d_node = v->s_node.as<int>(); }
  BowVoc V{v->child_start.as<int>(), v->child_id.as<int>(), v->child_desc.as<uint4>(), v->is_leaf.as<uint8_t>(), v->word_of.as<int>(),
           v->word_weight.as<double>(), v->L};
  long long launched = 0;
  orb_launch_bow(V, levelsup, d_desc, d_n, cap, frames, d_word, d_node, d_bw, d_bv, d_bc, d_fn, d_fs, d_ff, d_fc, st, &launched);
  orb_knn_count_launches(launched);
  CU(cudaGetLastError());
  if (host) {
    if (word) CU(cudaMemcpyAsync(word, d_word, F * cap * 4, cudaMemcpyDeviceToHost, st));
    if (node) CU(cudaMemcpyAsync(node, d_node, F * cap * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(bow_word, d_bw, F * cap * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(bow_value, d_bv, F * cap * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(bow_count, d_bc, F * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(fv_node, d_fn, F * cap * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(fv_start, d_fs, F * (cap + 1) * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(fv_feature, d_ff, F * cap * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(fv_count, d_fc, F * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  }
  return ORB_OK;
}

int orb_search_for_initialization(int mem, const orb_keypoint* kps_un, const uint8_t* desc, const int32_t* n_kps, int cap,
                                  const orb_proj_query* queries, const uint8_t* qdesc, const int32_t* n_queries, int qcap, int frames,
                                  const orb_search_params* prm, int32_t* matches12, int32_t* matches21, int32_t* nmatches, void* stream) {
  if (!kps_un || !desc || !n_kps || !queries || !qdesc || !n_queries || !prm || !matches12 || !nmatches) return fail(ORB_E_BADARG, "null argument");
  if (cap <= 0 || cap > 32768 || qcap <= 0 || frames <= 0) return fail(ORB_E_BADARG, "bad size (0 < cap <= 32768, qcap > 0, frames > 0)");
  if (!(prm->max_x > prm->min_x) || !(prm->max_y > prm->min_y)) return fail(ORB_E_BADARG, "empty image bounds");
  if (prm->th_dist < 0 || prm->th_dist > 255 || !(prm->nn_ratio > 0.f)) return fail(ORB_E_BADARG, "th_dist in [0, 255] and nn_ratio > 0 required");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  WsOrder order(&ws->ev_last, st);
  const size_t F = (size_t)frames;
  const bool host = mem != ORB_MEM_DEVICE;
  const void* d_in[6] = {kps_un, desc, n_kps, queries, qdesc, n_queries};
  int* d_m12 = matches12;
  int* d_m21 = matches21;
  int* d_nm = nmatches;
  if (host) {
    DevBuf* bufs[6] = {&ws->g_kps, &ws->g_desc, &ws->g_nk, &ws->g_q, &ws->g_qd, &ws->g_nq};
    const size_t bytes[6] = {F * cap * sizeof(orb_keypoint), F * cap * 32, F * sizeof(int), F * qcap * sizeof(orb_proj_query), F * qcap * 32, F * sizeof(int)};
    for (int i = 0; i < 6; ++i) {
      if ((rc = bufs[i]->ensure(bytes[i])) != ORB_OK) return rc;
      CU(cudaMemcpyAsync(bufs[i]->p, d_in[i], bytes[i], cudaMemcpyHostToDevice, st));
      d_in[i] = bufs[i]->p;
    }
    if ((rc = ws->g_qmatch.ensure(F * qcap * sizeof(int))) != ORB_OK) return rc;
    if ((rc = ws->g_nm.ensure(F * sizeof(int))) != ORB_OK) return rc;
    d_m12 = ws->g_qmatch.as<int>();
    d_nm = ws->g_nm.as<int>();
  }
  if (host || !matches21) {
    if ((rc = ws->g_assigned.ensure(F * cap * sizeof(int))) != ORB_OK) return rc;
    d_m21 = ws->g_assigned.as<int>();
  }
  if ((rc = ws->g_items.ensure(F * cap * sizeof(uint4))) != ORB_OK) return rc;
  if ((rc = ws->g_cells.ensure(F * (64 * 48 + 1) * sizeof(int))) != ORB_OK) return rc;
  if ((rc = ws->g_lists.ensure(F * qcap * 32 * sizeof(uint2))) != ORB_OK) return rc;
  if ((rc = ws->g_lcnt.ensure(F * qcap * sizeof(int))) != ORB_OK) return rc;
  long long launched = 0;
  orb_launch_search_init((const OrbKeypoint*)d_in[0], (const uint8_t*)d_in[1], (const int*)d_in[2], cap, d_in[3], (const uint8_t*)d_in[4],
                         (const int*)d_in[5], qcap, frames, prm->min_x, prm->min_y, prm->max_x, prm->max_y, prm->th_dist, prm->nn_ratio,
                         prm->check_orientation, ws->g_items.as<uint4>(), ws->g_cells.as<int>(), ws->g_lists.as<uint2>(), ws->g_lcnt.as<int>(),
                         d_m21, d_m12, d_nm, st, &launched);
  orb_knn_count_launches(launched);
  CU(cudaGetLastError());
  if (host) {
    CU(cudaMemcpyAsync(matches12, d_m12, F * qcap * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (matches21) CU(cudaMemcpyAsync(matches21, d_m21, F * cap * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(nmatches, d_nm, F * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  }
  return ORB_OK;
}

}  // extern "C"

// SearchByBoW (epi == NULL) and SearchForTriangulation (epi != NULL: per-pair epipolar geometry, stereo flags of both keyframes,
// no ratio test, later candidate wins ties) share the query builder, the candidate lists and the sequential resolve.
static int search_by_bow_impl(int mem, const uint8_t* kf_desc, const orb_keypoint* kf_kps_un, const uint8_t* kf_valid, const int32_t* kf_fv_node,
                              const int32_t* kf_fv_start, const int32_t* kf_fv_feature, const int32_t* kf_fv_count, int qcap, const uint8_t* f_desc,
                              const orb_keypoint* f_kps, const uint8_t* f_blocked, const int32_t* f_fv_node, const int32_t* f_fv_start,
                              const int32_t* f_fv_feature, const int32_t* f_fv_count, int cap, int frames, int th_dist, float nn_ratio,
                              int check_orientation, int32_t* assigned, int32_t* kf_match, int32_t* nmatches, void* stream,
                              const orb_epipolar* epi, const float* kf_uright, const float* f_uright) {
  const void* in[17] = {kf_desc, kf_kps_un, kf_valid, kf_fv_node, kf_fv_start, kf_fv_feature, kf_fv_count,
                        f_desc,  f_kps,     f_fv_node, f_fv_start, f_fv_feature, f_fv_count, f_blocked /* optional */,
                        epi /* optional from here */, kf_uright, f_uright};
  for (int i = 0; i < 13; ++i)
    if (!in[i]) return fail(ORB_E_BADARG, "null argument");
  if (!nmatches || (!assigned && !epi) || (epi && !kf_match)) return fail(ORB_E_BADARG, "null argument");
  if (cap <= 0 || cap >= (1 << 18) || qcap <= 0 || qcap >= (1 << 18) || frames <= 0) return fail(ORB_E_BADARG, "bad size (0 < cap, qcap < 2^18, frames > 0)");
  if (th_dist < 0 || th_dist > 255 || (!epi && !(nn_ratio > 0.f))) return fail(ORB_E_BADARG, "th_dist in [0, 255] and nn_ratio > 0 required");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  WsOrder order(&ws->ev_last, st);
  const size_t F = (size_t)frames, Q = (size_t)qcap, N = (size_t)cap;
  const bool host = mem != ORB_MEM_DEVICE;
  int* d_assigned = assigned;
  int* d_nm = nmatches;
  int* d_qmatch = kf_match;
  if (host) {
    // FeatureVectors in host memory are checked before they index device arrays (device-memory callers pass orb_bow_transform output)
    auto csr_ok = [](const int32_t* start, const int32_t* feature, int count, int limit) {
      if (count < 0 || count > limit || (count > 0 && start[0] != 0)) return false;
      for (int j = 0; j < count; ++j)
        if (start[j + 1] < start[j] || start[j + 1] > limit) return false;
      for (int e = 0; e < (count > 0 ? start[count] : 0); ++e)
        if (feature[e] < 0 || feature[e] >= limit) return false;
      return true;
    };
    for (int f = 0; f < frames; ++f) {
      if (!csr_ok(kf_fv_start + (size_t)f * (Q + 1), kf_fv_feature + (size_t)f * Q, kf_fv_count[f], qcap) ||
          !csr_ok(f_fv_start + (size_t)f * (N + 1), f_fv_feature + (size_t)f * N, f_fv_count[f], cap))
        return fail(ORB_E_BADARG, "frame %d: malformed FeatureVector (counts, ascending starts, feature indices in range)", f);
    }
    const size_t bytes[17] = {F * Q * 32, F * Q * sizeof(orb_keypoint), F * Q, F * Q * 4, F * (Q + 1) * 4, F * Q * 4, F * 4,
                              F * N * 32, F * N * sizeof(orb_keypoint), F * N * 4, F * (N + 1) * 4, F * N * 4, F * 4, F * N,
                              F * sizeof(orb_epipolar), F * Q * 4, F * N * 4};
    for (int i = 0; i < 17; ++i) {
      if (!in[i]) continue;
      if ((rc = ws->g_stage[i].ensure(bytes[i])) != ORB_OK) return rc;
      CU(cudaMemcpyAsync(ws->g_stage[i].p, in[i], bytes[i], cudaMemcpyHostToDevice, st));
      in[i] = ws->g_stage[i].p;
    }
    if ((rc = ws->g_assigned.ensure(F * N * sizeof(int))) != ORB_OK) return rc;
    if ((rc = ws->g_nm.ensure(F * sizeof(int))) != ORB_OK) return rc;
    d_assigned = ws->g_assigned.as<int>();
    d_nm = ws->g_nm.as<int>();
  }
  if (host || !kf_match) {
    if ((rc = ws->g_qmatch.ensure(F * Q * sizeof(int))) != ORB_OK) return rc;
    d_qmatch = ws->g_qmatch.as<int>();
  }
  if (!host && !assigned) {   // SearchForTriangulation callers only want vMatches12
    if ((rc = ws->g_assigned.ensure(F * N * sizeof(int))) != ORB_OK) return rc;
    d_assigned = ws->g_assigned.as<int>();
  }
  if ((rc = ws->g_bq.ensure(F * Q * 32)) != ORB_OK) return rc;
  if ((rc = ws->g_bnq.ensure(F * sizeof(int))) != ORB_OK) return rc;
  if ((rc = ws->g_lists.ensure(F * Q * 32 * sizeof(uint2))) != ORB_OK) return rc;
  if ((rc = ws->g_lcnt.ensure(F * Q * sizeof(int))) != ORB_OK) return rc;
  long long launched = 0;
  orb_launch_search_by_bow((const uint8_t*)in[0], (const OrbKeypoint*)in[1], (const uint8_t*)in[2], (const int*)in[3], (const int*)in[4],
                           (const int*)in[5], (const int*)in[6], qcap, (const uint8_t*)in[7], (const OrbKeypoint*)in[8], (const uint8_t*)in[13],
                           (const int*)in[9], (const int*)in[10], (const int*)in[11], (const int*)in[12], cap, frames, th_dist, nn_ratio,
                           check_orientation, ws->g_bq.p, ws->g_bnq.as<int>(), ws->g_lists.as<uint2>(), ws->g_lcnt.as<int>(), d_assigned, d_qmatch,
                           d_nm, st, &launched, in[14], (const float*)in[15], (const float*)in[16]);
  orb_knn_count_launches(launched);
  CU(cudaGetLastError());
  if (host) {
    if (assigned) CU(cudaMemcpyAsync(assigned, d_assigned, F * N * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (kf_match) CU(cudaMemcpyAsync(kf_match, d_qmatch, F * Q * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(nmatches, d_nm, F * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  }
  return ORB_OK;
}

extern "C" {

int orb_search_by_bow(int mem, const uint8_t* kf_desc, const orb_keypoint* kf_kps_un, const uint8_t* kf_valid, const int32_t* kf_fv_node,
                      const int32_t* kf_fv_start, const int32_t* kf_fv_feature, const int32_t* kf_fv_count, int qcap, const uint8_t* f_desc,
                      const orb_keypoint* f_kps, const uint8_t* f_blocked, const int32_t* f_fv_node, const int32_t* f_fv_start,
                      const int32_t* f_fv_feature, const int32_t* f_fv_count, int cap, int frames, int th_dist, float nn_ratio,
                      int check_orientation, int32_t* assigned, int32_t* kf_match, int32_t* nmatches, void* stream) {
  return search_by_bow_impl(mem, kf_desc, kf_kps_un, kf_valid, kf_fv_node, kf_fv_start, kf_fv_feature, kf_fv_count, qcap, f_desc, f_kps, f_blocked,
                            f_fv_node, f_fv_start, f_fv_feature, f_fv_count, cap, frames, th_dist, nn_ratio, check_orientation, assigned, kf_match,
                            nmatches, stream, nullptr, nullptr, nullptr);
}

int orb_search_for_triangulation(int mem, const uint8_t* kf1_desc, const orb_keypoint* kf1_kps_un, const uint8_t* kf1_free, const float* kf1_uright,
                                 const int32_t* kf1_fv_node, const int32_t* kf1_fv_start, const int32_t* kf1_fv_feature, const int32_t* kf1_fv_count,
                                 int qcap, const uint8_t* kf2_desc, const orb_keypoint* kf2_kps_un, const uint8_t* kf2_blocked, const float* kf2_uright,
                                 const int32_t* kf2_fv_node, const int32_t* kf2_fv_start, const int32_t* kf2_fv_feature, const int32_t* kf2_fv_count,
                                 int cap, int pairs, const orb_epipolar* epi, int th_dist, int check_orientation, int32_t* matches12,
                                 int32_t* nmatches, void* stream) {
  if (!epi || !kf1_uright || !kf2_uright || !kf2_blocked || !matches12) return fail(ORB_E_BADARG, "null argument");
  return search_by_bow_impl(mem, kf1_desc, kf1_kps_un, kf1_free, kf1_fv_node, kf1_fv_start, kf1_fv_feature, kf1_fv_count, qcap, kf2_desc, kf2_kps_un,
                            kf2_blocked, kf2_fv_node, kf2_fv_start, kf2_fv_feature, kf2_fv_count, cap, pairs, th_dist, 0.f, check_orientation, nullptr,
                            matches12, nmatches, stream, epi, kf1_uright, kf2_uright);
}

int orb_undistort_keypoints(int mem, const orb_camera* cam, const orb_keypoint* kps, const int32_t* n, int cap, int frames,
                            orb_keypoint* kps_un, void* stream) {
  if (!cam || !kps || !n || !kps_un || cap <= 0 || frames <= 0) return fail(ORB_E_BADARG, "bad argument");
  if (cam->n_dist != 4 && cam->n_dist != 5 && cam->n_dist != 8) return fail(ORB_E_BADARG, "n_dist must be 4, 5 or 8");
  if (!(cam->fx != 0.f) || !(cam->fy != 0.f)) return fail(ORB_E_BADARG, "fx / fy must be non-zero");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  WsOrder order(&ws->ev_last, st);
  const size_t bytes = (size_t)frames * cap * sizeof(orb_keypoint);
  if (mem == ORB_MEM_DEVICE) {
    orb_launch_undistort(cam, (const OrbKeypoint*)kps, n, cap, frames, (OrbKeypoint*)kps_un, st);
    orb_knn_count_launches(1);
    CU(cudaGetLastError());
    return ORB_OK;
  }
  if ((rc = ws->g_kps.ensure(bytes)) != ORB_OK) return rc;
  if ((rc = ws->g_nk.ensure((size_t)frames * sizeof(int))) != ORB_OK) return rc;
  CU(cudaMemcpyAsync(ws->g_kps.p, kps, bytes, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(ws->g_nk.p, n, (size_t)frames * sizeof(int), cudaMemcpyHostToDevice, st));
  orb_launch_undistort(cam, ws->g_kps.as<OrbKeypoint>(), ws->g_nk.as<int>(), cap, frames, ws->g_kps.as<OrbKeypoint>(), st);
  orb_knn_count_launches(1);
  CU(cudaGetLastError());
  if (kps_un != kps) memcpy(kps_un, kps, bytes);   // rows >= n[f] keep the caller's content (the kernel only rewrites rows < n[f])
  for (int f = 0; f < frames; ++f) {
    const int m = std::min(std::max(n[f], 0), cap);
    if (m > 0)
      CU(cudaMemcpyAsync(kps_un + (size_t)f * cap, ws->g_kps.as<orb_keypoint>() + (size_t)f * cap, (size_t)m * sizeof(orb_keypoint),
                         cudaMemcpyDeviceToHost, st));
  }
  CU(cudaStreamSynchronize(st));
  return ORB_OK;
}

int orb_image_bounds(const orb_camera* cam, int cols, int rows, float* bounds) {
  if (!cam || !bounds || cols <= 0 || rows <= 0) return fail(ORB_E_BADARG, "bad argument");
  if (cam->dist[0] == 0.0f) {   // src/Frame.cc:861-867
    bounds[0] = 0.f; bounds[1] = 0.f; bounds[2] = (float)cols; bounds[3] = (float)rows;
    return ORB_OK;
  }
  orb_keypoint c[4];
  memset(c, 0, sizeof(c));
  c[1].x = (float)cols; c[2].y = (float)rows; c[3].x = (float)cols; c[3].y = (float)rows;   // :838-845
  const int32_t four = 4;
  const int rc = orb_undistort_keypoints(ORB_MEM_HOST, cam, c, &four, 4, 1, c, nullptr);
  if (rc != ORB_OK) return rc;
  bounds[0] = std::min(c[0].x, c[2].x);   // mnMinX (:855)
  bounds[2] = std::max(c[1].x, c[3].x);   // mnMaxX
  bounds[1] = std::min(c[0].y, c[1].y);   // mnMinY
  bounds[3] = std::max(c[2].y, c[3].y);   // mnMaxY
  return ORB_OK;
}

int orb_distinctive_descriptors(int mem, const uint8_t* desc, const int64_t* offsets, int points, int32_t* best_idx,
                                int32_t* best_median, void* stream) {
  if (!offsets || !best_idx || points < 0) return fail(ORB_E_BADARG, "bad argument");
  if (points == 0) return ORB_OK;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  WsOrder order(&ws->ev_last, st);
  const bool host = mem != ORB_MEM_DEVICE;
  if (!host) {
    if (!desc || ((uintptr_t)desc & 15)) return fail(ORB_E_BADARG, "descriptor array must be 16-byte aligned device memory");
    orb_launch_distinctive(desc, (const long long*)offsets, points, best_idx, best_median, st);
    CU(cudaGetLastError());
    return ORB_OK;
  }
  const int64_t total = offsets[points];
  if (offsets[0] != 0 || total < 0 || (total > 0 && !desc)) return fail(ORB_E_BADARG, "offsets must start at 0 and be ascending");
  for (int p = 0; p < points; ++p)
    if (offsets[p + 1] < offsets[p] || offsets[p + 1] - offsets[p] >= (1 << 22)) return fail(ORB_E_BADARG, "offsets not ascending / more than 2^22 observations");
  if ((rc = ws->db.ensure((size_t)std::max<int64_t>(total, 1) * 32)) != ORB_OK) return rc;
  if ((rc = ws->q.ensure((size_t)(points + 1) * sizeof(int64_t))) != ORB_OK) return rc;
  if ((rc = ws->out.ensure((size_t)points * 2 * sizeof(int))) != ORB_OK) return rc;
  if (total > 0) CU(cudaMemcpyAsync(ws->db.p, desc, (size_t)total * 32, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(ws->q.p, offsets, (size_t)(points + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  int* d_idx = ws->out.as<int>();
  int* d_med = d_idx + points;
  orb_launch_distinctive(ws->db.as<uint8_t>(), ws->q.as<long long>(), points, d_idx, d_med, st);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(best_idx, d_idx, (size_t)points * sizeof(int), cudaMemcpyDeviceToHost, st));
  if (best_median) CU(cudaMemcpyAsync(best_median, d_med, (size_t)points * sizeof(int), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return ORB_OK;
}

int orb_search_by_projection(int mem, const orb_keypoint* kps_un, const uint8_t* desc, const float* uright, const uint8_t* claimed,
                             const int32_t* n_kps, int cap, const orb_proj_query* queries, const uint8_t* qdesc,
                             const int32_t* n_queries, int qcap, int frames, const orb_search_params* prm, int32_t* assigned,
                             int32_t* query_match, int32_t* nmatches, void* stream) {
  static_assert(sizeof(orb_proj_query) == 32 && sizeof(orb_search_params) == 32, "ABI struct layout");
  if (!kps_un || !desc || !n_kps || !queries || !qdesc || !n_queries || !prm || !assigned || !nmatches)
    return fail(ORB_E_BADARG, "null argument");
  if (cap <= 0 || cap >= (1 << 18) || qcap <= 0 || frames <= 0) return fail(ORB_E_BADARG, "bad size (0 < cap < 2^18, qcap > 0, frames > 0)");
  if (!(prm->max_x > prm->min_x) || !(prm->max_y > prm->min_y)) return fail(ORB_E_BADARG, "empty image bounds");
  if (prm->th_dist < 0 || prm->th_dist > 255) return fail(ORB_E_BADARG, "th_dist must be in [0, 255]");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  WsOrder order(&ws->ev_last, st);
  const size_t F = (size_t)frames;
  const bool host = mem != ORB_MEM_DEVICE;
  // device views of the arguments (staged when the caller's arrays are host memory)
  const void* d_in[7] = {kps_un, desc, uright, claimed, n_kps, queries, qdesc};
  const void* d_nq = n_queries;
  int* d_assigned = assigned;
  int* d_qmatch = query_match;
  int* d_nm = nmatches;
  if (host) {
    DevBuf* bufs[7] = {&ws->g_kps, &ws->g_desc, &ws->g_ur, &ws->g_claim, &ws->g_nk, &ws->g_q, &ws->g_qd};
    const size_t bytes[7] = {F * cap * sizeof(orb_keypoint), F * cap * 32, F * cap * sizeof(float), F * cap, F * sizeof(int),
                             F * qcap * sizeof(orb_proj_query), F * qcap * 32};
    for (int i = 0; i < 7; ++i) {
      if (!d_in[i]) continue;
      if ((rc = bufs[i]->ensure(bytes[i])) != ORB_OK) return rc;
      CU(cudaMemcpyAsync(bufs[i]->p, d_in[i], bytes[i], cudaMemcpyHostToDevice, st));
      d_in[i] = bufs[i]->p;
    }
    if ((rc = ws->g_nq.ensure(F * sizeof(int))) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->g_nq.p, n_queries, F * sizeof(int), cudaMemcpyHostToDevice, st));
    d_nq = ws->g_nq.p;
    if ((rc = ws->g_assigned.ensure(F * cap * sizeof(int))) != ORB_OK) return rc;
    if ((rc = ws->g_nm.ensure(F * sizeof(int))) != ORB_OK) return rc;
    d_assigned = ws->g_assigned.as<int>();
    d_nm = ws->g_nm.as<int>();
  }
  if (host || !query_match) {   // the resolve kernel always keeps the per-query record (rotation filter)
    if ((rc = ws->g_qmatch.ensure(F * qcap * sizeof(int))) != ORB_OK) return rc;
    d_qmatch = ws->g_qmatch.as<int>();
  }
  if ((rc = ws->g_items.ensure(F * cap * sizeof(uint4))) != ORB_OK) return rc;
  if ((rc = ws->g_cells.ensure(F * (64 * 48 + 1) * sizeof(int))) != ORB_OK) return rc;
  if ((rc = ws->g_lists.ensure(F * qcap * 32 * sizeof(uint2))) != ORB_OK) return rc;
  if ((rc = ws->g_lcnt.ensure(F * qcap * sizeof(int))) != ORB_OK) return rc;
  long long launched = 0;
  orb_launch_guided((const OrbKeypoint*)d_in[0], (const uint8_t*)d_in[1], (const float*)d_in[2], (const uint8_t*)d_in[3], (const int*)d_in[4],
                    cap, d_in[5], (const uint8_t*)d_in[6], (const int*)d_nq, qcap, frames, prm->min_x, prm->min_y, prm->max_x, prm->max_y,
                    prm->th_dist, prm->nn_ratio, prm->check_orientation, ws->g_items.as<uint4>(), ws->g_cells.as<int>(),
                    ws->g_lists.as<uint2>(), ws->g_lcnt.as<int>(), d_assigned, d_qmatch, d_nm, st, &launched);
  orb_knn_count_launches(launched);
  CU(cudaGetLastError());
  if (host) {
    CU(cudaMemcpyAsync(assigned, d_assigned, F * cap * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (query_match) CU(cudaMemcpyAsync(query_match, d_qmatch, F * qcap * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(nmatches, d_nm, F * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  }
  return ORB_OK;
}

int orb_fuse_search(int mem, const orb_keypoint* kps_un, const uint8_t* desc, const float* uright, const int32_t* n_kps, int cap,
                    const orb_proj_query* queries, const uint8_t* qdesc, const int32_t* n_queries, int qcap, int frames,
                    const orb_search_params* prm, const float* inv_level_sigma2, int32_t* best_idx, int32_t* best_dist, void* stream) {
  if (!kps_un || !desc || !n_kps || !queries || !qdesc || !n_queries || !prm || !best_idx) return fail(ORB_E_BADARG, "null argument");
  if (cap <= 0 || cap >= (1 << 18) || qcap <= 0 || frames <= 0) return fail(ORB_E_BADARG, "bad size (0 < cap < 2^18, qcap > 0, frames > 0)");
  if (!(prm->max_x > prm->min_x) || !(prm->max_y > prm->min_y)) return fail(ORB_E_BADARG, "empty image bounds");
  if (prm->th_dist < 0 || prm->th_dist > 255) return fail(ORB_E_BADARG, "th_dist must be in [0, 255]");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  WsOrder order(&ws->ev_last, st);
  const size_t F = (size_t)frames;
  const bool host = mem != ORB_MEM_DEVICE;
  const void* d_in[7] = {kps_un, desc, uright, nullptr, n_kps, queries, qdesc};
  const void* d_nq = n_queries;
  int* d_bi = best_idx;
  int* d_bd = best_dist;
  if (host) {
    DevBuf* bufs[7] = {&ws->g_kps, &ws->g_desc, &ws->g_ur, &ws->g_claim, &ws->g_nk, &ws->g_q, &ws->g_qd};
    const size_t bytes[7] = {F * cap * sizeof(orb_keypoint), F * cap * 32, F * cap * sizeof(float), 0, F * sizeof(int),
                             F * qcap * sizeof(orb_proj_query), F * qcap * 32};
    for (int i = 0; i < 7; ++i) {
      if (!d_in[i]) continue;
      if ((rc = bufs[i]->ensure(bytes[i])) != ORB_OK) return rc;
      CU(cudaMemcpyAsync(bufs[i]->p, d_in[i], bytes[i], cudaMemcpyHostToDevice, st));
      d_in[i] = bufs[i]->p;
    }
    if ((rc = ws->g_nq.ensure(F * sizeof(int))) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->g_nq.p, n_queries, F * sizeof(int), cudaMemcpyHostToDevice, st));
    d_nq = ws->g_nq.p;
    if ((rc = ws->g_qmatch.ensure(F * qcap * sizeof(int))) != ORB_OK) return rc;
    d_bi = ws->g_qmatch.as<int>();
  }
  if (host || !best_dist) {
    if ((rc = ws->g_lcnt.ensure(F * qcap * sizeof(int))) != ORB_OK) return rc;
    d_bd = ws->g_lcnt.as<int>();
  }
  const float* d_sig = nullptr;
  if (inv_level_sigma2) {   // always a host table of 16 floats (mvInvLevelSigma2, zero-padded), like prm
    if ((rc = ws->g_nm.ensure(16 * sizeof(float))) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->g_nm.p, inv_level_sigma2, 16 * sizeof(float), cudaMemcpyHostToDevice, st));
    d_sig = ws->g_nm.as<float>();
  }
  if ((rc = ws->g_items.ensure(F * cap * sizeof(uint4))) != ORB_OK) return rc;
  if ((rc = ws->g_cells.ensure(F * (64 * 48 + 1) * sizeof(int))) != ORB_OK) return rc;
  long long launched = 0;
  orb_launch_fuse_search((const OrbKeypoint*)d_in[0], (const uint8_t*)d_in[1], (const float*)d_in[2], (const int*)d_in[4], cap, d_in[5],
                         (const uint8_t*)d_in[6], (const int*)d_nq, qcap, frames, prm->min_x, prm->min_y, prm->max_x, prm->max_y, d_sig, prm->th_dist,
                         ws->g_items.as<uint4>(), ws->g_cells.as<int>(), d_bi, d_bd, st, &launched);
  orb_knn_count_launches(launched);
  CU(cudaGetLastError());
  if (host) {
    CU(cudaMemcpyAsync(best_idx, d_bi, F * qcap * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (best_dist) CU(cudaMemcpyAsync(best_dist, d_bd, F * qcap * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  } else if (inv_level_sigma2) {
    CU(cudaStreamSynchronize(st));   // the staged table must be consumed before the next call overwrites it
  }
  return ORB_OK;
}

double orb_measure_popc_peak(int device) {
  OrbDeviceGuard guard(device);
  if (guard.err != cudaSuccess) { cudaGetLastError(); return 0.0; }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return 0.0;
  return orb_popc_peak(prop.multiProcessorCount);
}

}  // extern "C"
