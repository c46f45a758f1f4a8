// Brute-force 256-bit Hamming k=2: ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2073-2093) inside the
// best/second-best scan of src/ORBmatcher.cc:304-335.
//
// The reference's sequential rule (strict <, earliest index wins, a duplicate of the minimum becomes `second`)
// returns exactly the two lexicographically smallest (distance, index) pairs with distance < sentinel, so the
// scan is order-independent: every thread keeps the two smallest packed keys (d << 22 | local row) of its own
// queries with three integer min/max per pair, database chunks are scanned by different CTAs, and a merge takes
// the two smallest of all partial keys. INT pipe (popc) bound; the database streams through shared memory as
// uint4 pairs read by broadcast, queries live in registers.
#include <atomic>
#include <mutex>
#include <cuda_pipeline.h>

#include "orb_host.h"

#define KNN_THREADS 128
#define KNN_QPT 2
#define KNN_TILE 256   // database rows per smem stage (8 KB)

struct OrbKnn2 {   // == orb_knn2
  int d1, d2;
  long long i1, i2;
};

__device__ __forceinline__ unsigned knn_dist(const uint4& qa, const uint4& qb, const uint4& a, const uint4& b) {
  return __popc(qa.x ^ a.x) + __popc(qa.y ^ a.y) + __popc(qa.z ^ a.z) + __popc(qa.w ^ a.w) + __popc(qb.x ^ b.x) +
         __popc(qb.y ^ b.y) + __popc(qb.z ^ b.z) + __popc(qb.w ^ b.w);
}

__global__ void __launch_bounds__(KNN_THREADS)
orb_knn_kernel(const uint4* __restrict__ query, int Q, const uint4* __restrict__ db, long long D, long long rows_per_chunk,
               long long index_base, unsigned long long* __restrict__ partial /*[nchunks][Q][2]*/) {
  __shared__ uint4 s_db[2][KNN_TILE * 2];
  const int tid = threadIdx.x;
  uint4 qa[KNN_QPT], qb[KNN_QPT];
  unsigned k1[KNN_QPT], k2[KNN_QPT];
#pragma unroll
  for (int i = 0; i < KNN_QPT; ++i) {
    const int qi = blockIdx.x * (KNN_THREADS * KNN_QPT) + i * KNN_THREADS + tid;
    qa[i] = qi < Q ? query[2 * qi] : make_uint4(0, 0, 0, 0);
    qb[i] = qi < Q ? query[2 * qi + 1] : make_uint4(0, 0, 0, 0);
    k1[i] = 0xffffffffu;
    k2[i] = 0xffffffffu;
  }
  const long long c0 = (long long)blockIdx.y * rows_per_chunk;
  const long long c1 = min(c0 + rows_per_chunk, D);
  const int ntiles = c1 > c0 ? (int)((c1 - c0 + KNN_TILE - 1) / KNN_TILE) : 0;
  auto stage = [&](int t, int buf) {
    const long long r0 = c0 + (long long)t * KNN_TILE;
    const int rows = (int)min((long long)KNN_TILE, c1 - r0);
    const uint4* src = db + 2 * r0;
    for (int i = tid; i < rows * 2; i += KNN_THREADS) __pipeline_memcpy_async(&s_db[buf][i], src + i, 16);
    __pipeline_commit();
  };
  if (ntiles > 0) stage(0, 0);
  for (int t = 0; t < ntiles; ++t) {
    const int buf = t & 1;
    if (t + 1 < ntiles) { stage(t + 1, buf ^ 1); __pipeline_wait_prior(1); }
    else __pipeline_wait_prior(0);
    __syncthreads();
    const int rows = (int)min((long long)KNN_TILE, c1 - (c0 + (long long)t * KNN_TILE));
    const unsigned jbase = (unsigned)(t * KNN_TILE);
#pragma unroll 4
    for (int r = 0; r < rows; ++r) {
      const uint4 a = s_db[buf][2 * r], b = s_db[buf][2 * r + 1];
#pragma unroll
      for (int i = 0; i < KNN_QPT; ++i) {
        const unsigned key = (knn_dist(qa[i], qb[i], a, b) << 22) | (jbase + r);
        k2[i] = min(k2[i], max(k1[i], key));
        k1[i] = min(k1[i], key);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < KNN_QPT; ++i) {
    const int qi = blockIdx.x * (KNN_THREADS * KNN_QPT) + i * KNN_THREADS + tid;
    if (qi < Q) {
      unsigned long long* out = partial + ((long long)blockIdx.y * Q + qi) * 2;
      out[0] = k1[i] == 0xffffffffu ? ~0ull : (((unsigned long long)(k1[i] >> 22) << 40) | (unsigned long long)(index_base + c0 + (k1[i] & 0x3fffffu)));
      out[1] = k2[i] == 0xffffffffu ? ~0ull : (((unsigned long long)(k2[i] >> 22) << 40) | (unsigned long long)(index_base + c0 + (k2[i] & 0x3fffffu)));
    }
  }
}

__device__ __forceinline__ void knn_emit(unsigned long long a, unsigned long long b, int sentinel, OrbKnn2* o) {
  OrbKnn2 r;
  const long long da = a == ~0ull ? (long long)sentinel : (long long)(a >> 40);
  const long long dbb = b == ~0ull ? (long long)sentinel : (long long)(b >> 40);
  // a distance that is not < sentinel is never recorded by the reference loop
  if (da < sentinel) { r.d1 = (int)da; r.i1 = (long long)(a & 0xffffffffffull); } else { r.d1 = sentinel; r.i1 = -1; }
  if (da < sentinel && dbb < sentinel) { r.d2 = (int)dbb; r.i2 = (long long)(b & 0xffffffffffull); } else { r.d2 = sentinel; r.i2 = -1; }
  *o = r;
}

__global__ void orb_knn_merge_partial_kernel(const unsigned long long* __restrict__ partial, int nchunks, int Q, int sentinel,
                                             OrbKnn2* __restrict__ out) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= Q) return;
  unsigned long long a = ~0ull, b = ~0ull;
  for (int c = 0; c < nchunks; ++c) {
    const unsigned long long* p = partial + ((long long)c * Q + qi) * 2;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const unsigned long long k = p[j];
      b = min(b, max(a, k));
      a = min(a, k);
    }
  }
  knn_emit(a, b, sentinel, out + qi);
}

// Merge per-shard orb_knn2 results (e.g. all-gathered over NCCL): two lexicographically smallest (d, i).
__global__ void orb_knn_merge_shards_kernel(const OrbKnn2* __restrict__ shards, int nshards, int Q, int sentinel,
                                            OrbKnn2* __restrict__ out) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= Q) return;
  unsigned long long a = ~0ull, b = ~0ull;
  for (int s = 0; s < nshards; ++s) {
    const OrbKnn2 e = shards[(long long)s * Q + qi];
    const unsigned long long ka = (e.i1 < 0 || e.d1 >= sentinel) ? ~0ull : (((unsigned long long)e.d1 << 40) | (unsigned long long)e.i1);
    const unsigned long long kb = (e.i2 < 0 || e.d2 >= sentinel) ? ~0ull : (((unsigned long long)e.d2 << 40) | (unsigned long long)e.i2);
    b = min(b, max(a, ka)); a = min(a, ka);
    b = min(b, max(a, kb)); a = min(a, kb);
  }
  knn_emit(a, b, sentinel, out + qi);
}

__global__ void orb_distance_batch_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, long long n, int* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = (int)knn_dist(a[2 * i], a[2 * i + 1], b[2 * i], b[2 * i + 1]);
}

// Register-resident xor+popc+add loop: 8 independent chains per thread.
__global__ void orb_popc_peak_kernel(unsigned* out, int iters, unsigned seed) {
  unsigned x[8], acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { x[i] = seed * (threadIdx.x + 1) + i * 0x9e3779b9u; acc[i] = 0; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) { acc[i] += __popc(x[i] ^ acc[i]); }
  }
  unsigned s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += acc[i];
  if (s == 0xdeadbeefu) out[0] = s;
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:424-517): one warp per map point. For row i the lanes stride over
// the N descriptors, histogram the distances (0..256) in shared memory, and the median = element (int)(0.5 * (N - 1)) of the
// sorted row is read off the histogram's prefix sums; best row = min over (median << 22 | i).
#define DISTINCT_WARPS 8
__global__ void __launch_bounds__(DISTINCT_WARPS * 32)
orb_distinctive_kernel(const uint4* __restrict__ desc, const long long* __restrict__ offsets, int points, int* __restrict__ best_idx,
                       int* __restrict__ best_median) {
  __shared__ int s_hist[DISTINCT_WARPS][288];   // 257 bins, padded to 9 per lane
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int p = blockIdx.x * DISTINCT_WARPS + w;
  if (p >= points) return;
  const long long o0 = offsets[p];
  const int n = (int)(offsets[p + 1] - o0);
  const uint4* d = desc + 2 * o0;
  int* hist = s_hist[w];
  const int k = (int)(0.5 * (n - 1));   // vDists[0.5 * (N - 1)]
  unsigned best = 0xffffffffu;
  for (int i = 0; i < n; ++i) {
#pragma unroll
    for (int b = 0; b < 9; ++b) hist[lane * 9 + b] = 0;
    __syncwarp();
    const uint4 a0 = d[2 * i], a1 = d[2 * i + 1];
    for (int j = lane; j < n; j += 32) {
      const uint4 b0 = d[2 * j], b1 = d[2 * j + 1];
      const int dist = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                       __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
      atomicAdd(&hist[dist], 1);
    }
    __syncwarp();
    int c[9], sum = 0;
#pragma unroll
    for (int b = 0; b < 9; ++b) { c[b] = hist[lane * 9 + b]; sum += c[b]; }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    // the lane whose bins contain rank k (first lane with inclusive count > k)
    const unsigned has = __ballot_sync(0xffffffffu, incl > k);
    const int src = __ffs(has) - 1;
    int median = 0;
    if (lane == src) {
      int run = incl - sum;
#pragma unroll
      for (int b = 0; b < 9; ++b) {
        run += c[b];
        if (run > k) { median = lane * 9 + b; break; }
      }
    }
    median = __shfl_sync(0xffffffffu, median, src);
    best = min(best, ((unsigned)median << 22) | (unsigned)i);   // strict <: the first row with the smallest median wins
    __syncwarp();
  }
  if (lane == 0) {
    best_idx[p] = n > 0 ? (int)(best & 0x3fffffu) : -1;
    if (best_median) best_median[p] = n > 0 ? (int)(best >> 22) : -1;
  }
}

static std::atomic<long long> g_knn_launches{0};   // the kNN entry points are callable from any thread
long long orb_knn_launches() { return g_knn_launches.load(std::memory_order_relaxed); }
void orb_knn_count_launches(long long n) { g_knn_launches += n; }

int orb_knn_num_chunks(int Q, long long D, int num_sms) {
  const int qtiles = (Q + KNN_THREADS * KNN_QPT - 1) / (KNN_THREADS * KNN_QPT);
  // ~4 CTAs per SM across the grid, chunks of >= 1 tile and < 2^22 rows
  long long want = (4LL * num_sms + qtiles - 1) / qtiles;
  const long long max_chunks = (D + KNN_TILE - 1) / KNN_TILE;
  if (want > max_chunks) want = max_chunks;
  const long long min_chunks = (D + (1LL << 22) - 1) >> 22;
  if (want < min_chunks) want = min_chunks;
  if (want < 1) want = 1;
  return (int)want;
}

void orb_launch_knn(const uint8_t* q, int Q, const uint8_t* db, long long D, long long index_base, int sentinel,
                    unsigned long long* partial, int nchunks, OrbKnn2* out, cudaStream_t st) {
  const int qtiles = (Q + KNN_THREADS * KNN_QPT - 1) / (KNN_THREADS * KNN_QPT);
  long long rows = (D + nchunks - 1) / nchunks;
  rows = (rows + 7) & ~7LL;   // chunks need not end on a tile: the kernel's last tile is partial. Equal chunks keep all CTAs busy on small shards
  orb_knn_kernel<<<dim3(qtiles, nchunks), KNN_THREADS, 0, st>>>((const uint4*)q, Q, (const uint4*)db, D, rows, index_base, partial);
  orb_knn_merge_partial_kernel<<<(Q + 127) / 128, 128, 0, st>>>(partial, nchunks, Q, sentinel, out);
  g_knn_launches += 2;
}

// ---- sharded scan with the exchange inside the kernel (orb_hamming_knn_peer) -------------------------------------------------------------------
// ONE kernel = reduction of this rank's per-chunk keys + all-gather + merge, the gather done by plain stores into every peer's mailbox over
// NVLink (the mailboxes are cudaMalloc'ed by each rank and opened in the others with CUDA IPC):
//   1. thread = query: the two smallest packed keys (distance << 40 | GLOBAL row) of this rank's shard, written as one 16-byte store into slot
//      [parity][my rank][query] of EVERY rank's mailbox (its own included);
//   2. __threadfence_system(), a ticket per CTA; the CTA that draws the last ticket publishes: flag[my rank] = epoch in every mailbox
//      (st.release.sys);
//   3. every CTA waits until all flags of its OWN mailbox have reached the epoch (ld.acquire.sys on local memory; bounded: a missing peer becomes
//      an error flag, not a hang), then merges the nranks slots of each query (L1-bypassing loads) and emits the record.
// The grid is (Q + 127) / 128 CTAs, far below one wave, so every CTA is resident and the in-kernel wait cannot starve a publisher. Slots are
// double-buffered by the parity of the epoch: a rank can only be one call ahead of the slowest peer (it needs that peer's flag of the current
// call), so the buffer it overwrites two calls later has been read by everybody.
struct OrbPeerBox {
  unsigned long long* slots[ORB_KNN_PEER_MAX_RANKS];   // mailbox of rank r: [2][nranks][max_q] x 2 keys
  unsigned* flags[ORB_KNN_PEER_MAX_RANKS];              // mailbox of rank r: [nranks] epochs
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(128)
orb_knn_peer_kernel(const unsigned long long* __restrict__ partial, int nchunks, int Q, int sentinel, const __grid_constant__ OrbPeerBox box,
                    int my_rank, int nranks, int max_q, unsigned epoch, unsigned* __restrict__ ticket, OrbKnn2* __restrict__ out,
                    int* __restrict__ err_flag) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned par = epoch & 1u;
  // 1. this rank's two best keys -> every mailbox
  if (qi < Q) {
    unsigned long long a = ~0ull, b = ~0ull;
    for (int c = 0; c < nchunks; ++c) {
      const unsigned long long* p = partial + ((long long)c * Q + qi) * 2;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const unsigned long long k = p[j];
        b = min(b, max(a, k));
        a = min(a, k);
      }
    }
    const size_t slot = (((size_t)par * nranks + my_rank) * max_q + qi) * 2;
    for (int r = 0; r < nranks; ++r) *reinterpret_cast<ulonglong2*>(box.slots[r] + slot) = make_ulonglong2(a, b);
  }
  // 2. publish once every CTA of this rank has written
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    if (atomicAdd(ticket, 1u) == gridDim.x - 1) {
      *ticket = 0u;                     // every CTA has drawn: ready for the next call
      __threadfence_system();
      for (int r = 0; r < nranks; ++r) st_release_sys(box.flags[r] + my_rank, epoch);
    }
  }
  // 3. wait for all ranks' flags in MY mailbox, merge
  if (threadIdx.x < nranks) {
    const unsigned* f = box.flags[my_rank] + threadIdx.x;
    long long spins = 0;
    while ((int)(ld_acquire_sys(f) - epoch) < 0) {
      if (++spins > (1ll << 26)) { atomicOr(err_flag, 32); break; }   // seconds: a peer that never called
      __nanosleep(64);
    }
  }
  __syncthreads();
  if (qi >= Q) return;
  unsigned long long a = ~0ull, b = ~0ull;
  const unsigned long long* mine = box.slots[my_rank] + (size_t)par * nranks * max_q * 2;
  for (int r = 0; r < nranks; ++r) {
    const ulonglong2 e = __ldcg(reinterpret_cast<const ulonglong2*>(mine + ((size_t)r * max_q + qi) * 2));   // written by a peer: not through L1
    b = min(b, max(a, e.x)); a = min(a, e.x);
    b = min(b, max(a, e.y)); a = min(a, e.y);
  }
  knn_emit(a, b, sentinel, out + qi);
}

void orb_launch_knn_peer(const uint8_t* q, int Q, const uint8_t* db, long long D, long long index_base, int sentinel, unsigned long long* partial,
                         int nchunks, const void* box, int my_rank, int nranks, int max_q, unsigned epoch, unsigned* ticket, OrbKnn2* out,
                         int* err_flag, cudaStream_t st) {
  const int qtiles = (Q + KNN_THREADS * KNN_QPT - 1) / (KNN_THREADS * KNN_QPT);
  long long rows = (D + nchunks - 1) / nchunks;
  rows = (rows + 7) & ~7LL;   // chunks need not end on a tile: the kernel's last tile is partial. Equal chunks keep all CTAs busy on small shards
  orb_knn_kernel<<<dim3(qtiles, nchunks), KNN_THREADS, 0, st>>>((const uint4*)q, Q, (const uint4*)db, D, rows, index_base, partial);
  orb_knn_peer_kernel<<<(Q + 127) / 128, 128, 0, st>>>(partial, nchunks, Q, sentinel, *reinterpret_cast<const OrbPeerBox*>(box), my_rank, nranks,
                                                      max_q, epoch, ticket, out, err_flag);
  g_knn_launches += 2;
}

void orb_launch_knn_merge_shards(const OrbKnn2* shards, int nshards, int Q, int sentinel, OrbKnn2* out, cudaStream_t st) {
  orb_knn_merge_shards_kernel<<<(Q + 127) / 128, 128, 0, st>>>(shards, nshards, Q, sentinel, out);
  ++g_knn_launches;
}

void orb_launch_distance_batch(const uint8_t* a, const uint8_t* b, long long n, int* out, cudaStream_t st) {
  orb_distance_batch_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const uint4*)a, (const uint4*)b, n, out);
  ++g_knn_launches;
}

void orb_launch_distinctive(const uint8_t* desc, const long long* offsets, int points, int* best_idx, int* best_median, cudaStream_t st) {
  orb_distinctive_kernel<<<(points + DISTINCT_WARPS - 1) / DISTINCT_WARPS, DISTINCT_WARPS * 32, 0, st>>>((const uint4*)desc, offsets, points,
                                                                                                   best_idx, best_median);
  ++g_knn_launches;
}

double orb_popc_peak(int num_sms) {
  unsigned* d;
  if (cudaMalloc(&d, 4) != cudaSuccess) return 0.0;
  const int iters = 1 << 16, threads = 512, blocks = num_sms * 4;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  orb_popc_peak_kernel<<<blocks, threads>>>(d, 1024, 1u);
  double best = 0.0;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    orb_popc_peak_kernel<<<blocks, threads>>>(d, iters, 12345u + rep);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double ops = (double)blocks * threads * (double)iters * 8.0;
    if (ms > 0.f) best = fmax(best, ops / (ms * 1e-3));
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  return best;
}

// ---- allocation-free variant (no counterpart in the reference) -----------------------------
// orb_hamming_knn keeps its per-chunk partial keys in a grow-only cudaMalloc'ed workspace, which callers under CUDA graph capture, tight
// serving loops or harnesses that forbid allocation cannot use. This entry point runs the same two kernels with the partials in a
// file-scope device array of fixed size: every GPU that loads the library holds these 4 MiB until the process exits.
//   * In host code `g_knn_noalloc_scratch` is only a host-side stand-in for the symbol, NOT a device address: the address is resolved once
//     per device (that device current) with cudaGetSymbolAddress and that pointer is what the kernels receive.
//   * All calls on a device share the array and it keeps the previous call's contents. Nothing has to start at zero: the scan kernel writes
//     every word [chunk][query][2] that the merge kernel of the same call reads (empty chunks write the "no candidate" key), so stale
//     contents are never observed. Calls that OVERLAP on the GPU (two streams, two threads) race on it: issue them on one stream.
static __device__ __align__(256) unsigned long long g_knn_noalloc_scratch[ORB_KNN_NOALLOC_SCRATCH_BYTES / 8];

struct KnnNoAllocDev {
  std::once_flag once;
  cudaError_t err = cudaSuccess;
  unsigned long long* scratch = nullptr;   // device address of g_knn_noalloc_scratch on this device
  int num_sms = 0;
};
static KnnNoAllocDev g_knn_noalloc_dev[64];

extern "C" int orb_hamming_knn_noalloc(const uint8_t* query, int Q, const uint8_t* db, int64_t D, int64_t index_base, int sentinel,
                                       orb_knn2* out, void* stream) {
  if (!query || !out || Q <= 0 || D < 0 || (D > 0 && !db) || sentinel < 1) return fail(ORB_E_BADARG, "bad argument");
  if (index_base < 0 || index_base + D >= (1LL << 40)) return fail(ORB_E_BADARG, "index range exceeds 2^40");
  if (((uintptr_t)query & 15) || ((uintptr_t)db & 15) || ((uintptr_t)out & 7))
    return fail(ORB_E_BADARG, "orb_hamming_knn_noalloc: query / db must be 16-byte aligned DEVICE memory, out 8-byte aligned (no staging copy is possible without allocating)");
  if (Q > ORB_KNN_NOALLOC_MAX_Q)
    return fail(ORB_E_CAPACITY, "orb_hamming_knn_noalloc: Q = %d exceeds the fixed scratch (max %d queries per call; use orb_hamming_knn)", Q, ORB_KNN_NOALLOC_MAX_Q);
  const long long words = ORB_KNN_NOALLOC_SCRATCH_BYTES / 8;
  const long long fit_chunks = words / (2LL * Q);                       // chunks whose [Q][2] partial keys fit the array (>= 64)
  const long long min_chunks = (D + (1LL << 22) - 1) >> 22;             // a chunk holds < 2^22 rows (packed local row index)
  if (min_chunks > fit_chunks)
    return fail(ORB_E_CAPACITY, "orb_hamming_knn_noalloc: D = %lld rows x Q = %d needs %lld chunks, the fixed scratch holds %lld (use orb_hamming_knn)",
                (long long)D, Q, min_chunks, fit_chunks);
  int dev = 0;
  cudaError_t ce = cudaGetDevice(&dev);
  if (ce != cudaSuccess) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback): %s", cudaGetErrorString(ce)); }
  if (dev < 0 || dev >= 64) return fail(ORB_E_BADARG, "device %d out of range", dev);
  KnnNoAllocDev& S = g_knn_noalloc_dev[dev];
  // set-up, once per device, with that device current (the caller's current device is the one the buffers live on)
  std::call_once(S.once, [&S, dev]() {
    void* p = nullptr;
    S.err = cudaGetSymbolAddress(&p, g_knn_noalloc_scratch);
    if (S.err == cudaSuccess) S.err = cudaDeviceGetAttribute(&S.num_sms, cudaDevAttrMultiProcessorCount, dev);
    S.scratch = reinterpret_cast<unsigned long long*>(p);
  });
  if (S.err != cudaSuccess || !S.scratch)
    return fail(ORB_E_CUDA, "orb_hamming_knn_noalloc: cudaGetSymbolAddress(g_knn_noalloc_scratch) on device %d: %s", dev, cudaGetErrorString(S.err));
  long long nchunks = orb_knn_num_chunks(Q, D, S.num_sms);
  if (nchunks > fit_chunks) nchunks = fit_chunks;   // >= min_chunks, checked above
  orb_launch_knn(query, Q, db, D, index_base, sentinel, S.scratch, (int)nchunks, reinterpret_cast<OrbKnn2*>(out), (cudaStream_t)stream);
  CU(cudaGetLastError());
  return ORB_OK;
}
