// Frame::ComputeBoW (src/Frame.cc:750-765) = DBoW2 TemplatedVocabulary<FORB>::transform(features, BowVector, FeatureVector, levelsup)
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1136-1195, 1240-1280; FORB::distance FORB.cpp:81-101; BowVector.cpp addWeight /
// normalize; FeatureVector.cpp addFeature) for the ORB vocabulary's configuration: TF_IDF weighting, L1_NORM scoring.
//
//   1. orb_bow_walk_kernel  (16 lanes per descriptor): tree walk. The children of a node are stored contiguously, lane c takes
//      child c; min over dist << 20 | c reproduces "d < best_d" (strict: the first child wins ties). Emits word id and the
//      ancestor at level L - levelsup per feature.
//   2. orb_bow_reduce_kernel (CTA per frame): bitonic sort of (word, feature) in shared memory; per word the weights are
//      summed in feature order (double, sequential like BowVector::addWeight); L1 norm summed in ascending word order (the
//      std::map walk of BowVector::normalize) and divided; then the same sort on (node, feature) gives the FeatureVector as CSR.
#include <atomic>

#include "orb_common.cuh"

struct BowVoc {
  const int* child_start;    // [n_nodes + 1] into the child arrays
  const int* child_id;       // node id of child slot
  const uint4* child_desc;   // [slots][2]
  const uint8_t* is_leaf;    // [n_nodes]
  const int* word_of;        // [n_nodes] word id of a leaf
  const double* word_weight; // [n_words]
  int L;
};

__global__ void __launch_bounds__(256)
orb_bow_walk_kernel(BowVoc V, int levelsup, const uint8_t* __restrict__ desc, const int* __restrict__ n_feat, int cap, int* __restrict__ word,
                    int* __restrict__ node) {
  const int f = blockIdx.y;
  const int i = blockIdx.x * 16 + (threadIdx.x >> 4);
  const int sub = threadIdx.x & 15;
  const unsigned gmask = 0xffffu << (threadIdx.x & 16);   // the 16 lanes of this descriptor
  const int n = min(n_feat[f], cap);
  if (i >= n) return;
  const uint4* d = reinterpret_cast<const uint4*>(desc + ((size_t)f * cap + i) * 32);
  const uint4 a0 = d[0], a1 = d[1];
  const int nid_level = V.L - levelsup;
  int nid = 0;          // "if (nid_level <= 0) *nid = 0" (TemplatedVocabulary.h:1250)
  int final_id = 0, level = 0;
  do {
    ++level;
    const int c0 = V.child_start[final_id], c1 = V.child_start[final_id + 1];
    unsigned best = 0xffffffffu;
    for (int c = c0 + sub; c < c1; c += 16) {
      const uint4 b0 = V.child_desc[2 * c], b1 = V.child_desc[2 * c + 1];
      const unsigned dist = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                            __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
      best = min(best, (dist << 20) | (unsigned)(c - c0));
    }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(gmask, best, o));
    final_id = V.child_id[c0 + (int)(best & 0xfffffu)];
    if (level == nid_level) nid = final_id;
  } while (!V.is_leaf[final_id]);
  if (sub == 0) {
    word[(size_t)f * cap + i] = V.word_of[final_id];
    node[(size_t)f * cap + i] = nid;
  }
}

__device__ __forceinline__ void bow_bitonic_sort(unsigned long long* key, int np2, int tid, int nthreads) {
  for (int k = 2; k <= np2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = tid; t < np2; t += nthreads) {
        const int p = t ^ j;
        if (p > t) {
          const unsigned long long a = key[t], b = key[p];
          const bool up = (t & k) == 0;
          if ((a > b) == up) { key[t] = b; key[p] = a; }
        }
      }
      __syncthreads();
    }
  }
}

// exclusive block scan of one int per (thread, chunk) is avoided: head flags are counted with a running per-thread base
__global__ void __launch_bounds__(256)
orb_bow_reduce_kernel(BowVoc V, const int* __restrict__ n_feat, int cap, int np2_cap, const int* __restrict__ word,
                      const int* __restrict__ node, int* __restrict__ bow_word, double* __restrict__ bow_value, int* __restrict__ bow_count,
                      int* __restrict__ fv_node, int* __restrict__ fv_start, int* __restrict__ fv_feature, int* __restrict__ fv_count) {
  extern __shared__ unsigned long long s_key[];   // [np2_cap]
  __shared__ int s_scan[256];
  __shared__ int s_total;
  __shared__ double s_norm;
  const int f = blockIdx.x, tid = threadIdx.x;
  const int n = min(n_feat[f], cap);
  int np2 = 1;
  while (np2 < n) np2 <<= 1;
  word += (size_t)f * cap; node += (size_t)f * cap;
  bow_word += (size_t)f * cap; bow_value += (size_t)f * cap;
  fv_node += (size_t)f * cap; fv_start += (size_t)f * (cap + 1); fv_feature += (size_t)f * cap;

  for (int pass = 0; pass < 2; ++pass) {   // 0: BowVector over (word, feature); 1: FeatureVector over (node, feature)
    const int* src = pass == 0 ? word : node;
    for (int t = tid; t < np2; t += 256) {
      unsigned long long k = ~0ull;
      if (t < n && V.word_weight[word[t]] > 0.0) k = ((unsigned long long)(unsigned)src[t] << 32) | (unsigned)t;   // "if (w > 0) // not stopped"
      s_key[t] = k;
    }
    __syncthreads();
    bow_bitonic_sort(s_key, np2, tid, 256);
    // heads of the runs of equal high words; each thread owns a contiguous slice so that output positions are ordered
    const int per = (np2 + 255) / 256;
    const int t0 = min(tid * per, np2), t1 = min(t0 + per, np2);
    int heads = 0;
    for (int t = t0; t < t1; ++t) {
      const unsigned long long k = s_key[t];
      if (k != ~0ull && (t == 0 || (unsigned)(s_key[t - 1] >> 32) != (unsigned)(k >> 32))) ++heads;
    }
    s_scan[tid] = heads;
    __syncthreads();
    if (tid == 0) {
      int run = 0;
      for (int i = 0; i < 256; ++i) { const int h = s_scan[i]; s_scan[i] = run; run += h; }
      s_total = run;
    }
    __syncthreads();
    int pos = s_scan[tid];
    for (int t = t0; t < t1; ++t) {
      const unsigned long long k = s_key[t];
      if (k == ~0ull) continue;
      const unsigned hi = (unsigned)(k >> 32);
      if (pass == 1) fv_feature[t] = (int)(unsigned)k;   // feature indices ascending inside a node (push_back order)
      if (t == 0 || (unsigned)(s_key[t - 1] >> 32) != hi) {
        if (pass == 0) {
          // BowVector::addWeight in feature order: the first feature inserts w, every further one does "second += w"
          const double w = V.word_weight[hi];
          double v = w;
          for (int u = t + 1; u < np2 && (unsigned)(s_key[u] >> 32) == hi && s_key[u] != ~0ull; ++u) v += w;
          bow_word[pos] = (int)hi;
          bow_value[pos] = v;
        } else {
          fv_node[pos] = (int)hi;
          fv_start[pos] = t;
        }
        ++pos;
      }
    }
    __syncthreads();
    const int total = s_total;
    if (pass == 0) {
      // BowVector::normalize(L1): norm += fabs(second) over the map in ascending word order, then second /= norm
      if (tid < 32) {
        double norm = 0.0;
        for (int base = 0; base < total; base += 32) {
          const double v = base + tid < total ? fabs(bow_value[base + tid]) : 0.0;
          for (int j = 0; j < 32; ++j) norm += __shfl_sync(0xffffffffu, v, j);   // + 0.0 past the end leaves norm unchanged
        }
        if (tid == 0) s_norm = norm;
      }
      __syncthreads();
      const double norm = s_norm;
      if (norm > 0.0)
        for (int t = tid; t < total; t += 256) bow_value[t] = bow_value[t] / norm;
      if (tid == 0) bow_count[f] = total;
    } else {
      if (tid == 0) {
        fv_count[f] = total;
        int valid = 0;   // number of non-stopped features = end of the last node's range
        int lo = 0, hi = np2;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (s_key[mid] != ~0ull) lo = mid + 1; else hi = mid; }
        valid = lo;
        fv_start[total] = valid;
      }
    }
    __syncthreads();
  }
}

void orb_launch_bow(const BowVoc& V, int levelsup, const uint8_t* desc, const int* n_feat, int cap, int frames, int* word, int* node,
                    int* bow_word, double* bow_value, int* bow_count, int* fv_node, int* fv_start, int* fv_feature, int* fv_count,
                    cudaStream_t st, long long* launches) {
  orb_bow_walk_kernel<<<dim3((cap + 15) / 16, frames), 256, 0, st>>>(V, levelsup, desc, n_feat, cap, word, node);
  int np2 = 1;
  while (np2 < cap) np2 <<= 1;
  const size_t smem = (size_t)np2 * sizeof(unsigned long long);
  static std::atomic<bool> configured[64];
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (smem > 48 * 1024 && !configured[dev].load(std::memory_order_acquire)) {
    cudaFuncSetAttribute(orb_bow_reduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    configured[dev].store(true, std::memory_order_release);
  }
  orb_bow_reduce_kernel<<<frames, 256, smem, st>>>(V, n_feat, cap, np2, word, node, bow_word, bow_value, bow_count, fv_node, fv_start,
                                                   fv_feature, fv_count);
  *launches += 2;
}
