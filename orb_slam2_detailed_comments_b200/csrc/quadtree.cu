// ORBextractor::DistributeOctTree (src/ORBextractor.cc:1050-1417) + ExtractorNode::DivideNode (:963-1035)
// as a data-parallel round algorithm, one CTA per (frame, level).
//
// What the reference's std::list manipulation computes, restated:
//  * A key never changes relative order inside a node (push_back in scan order), so the per-node winner
//    "first key with the maximum response" (:1394-1409) is max(response), ties -> lowest raw index.
//    Keys therefore only carry a node id; no per-node vectors are materialised.
//  * One "round" splits an ordered parent sequence P. Children are push_front'ed in order n1..n4 and the
//    parent erased, so new_list = reverse(concat_r nonempty_children(P[r])) ++ (old_list minus split parents).
//  * Phase 1 (:1166-1268): P = every node with > 1 keys in list order (= reverse creation order of the
//    previous round); stop if size >= N or nothing grew; enter phase 2 when size + 3*nToExpand > N.
//  * Phase 2 (:1284-1374): P = previous round's multi-key children sorted by (size, creation seq) descending
//    (canonical replacement of the reference's (size, heap pointer) order), cut at the first
//    parent after which size >= N. The cut is a prefix-sum over P of (nonempty children - 1).
// Child slots: the first non-empty child reuses its parent's slot, so live slots == list size <= max(N+3, 4*nIni).
#include <cstdlib>

#include "orb_common.cuh"

// CTA size NT: 192 threads for the usual levels (the kernel shares the SMs with the blur of the same pass and the FAST kernel of the
// other pass stream: 512-thread CTAs cost the KITTI step 3.5 %, 192 instead of 256 gains 0.8 % / TUM 1.7 %), 512 for large levels
// (4K: +11 % images/s) and for small passes (per-call latency), chosen by orb_launch_quadtree.
#define QT_SORT_SMEM 2048
#define QT_ILP 4       // keys in flight per thread in the per-round key passes
#define QT_SMALL 192   // CTA size for the usual levels (register caps below the natural 56 were measured too: no gain)

template <int NT, typename T>
__device__ T qt_block_scan_excl(T* a, int n, T* s_warp) {
  // in-place exclusive scan over a[0..n); every thread of the block calls it. Returns the total.
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  T carry = 0;
  for (int base = 0; base < n; base += NT) {
    const int i = base + threadIdx.x;
    const T v = i < n ? a[i] : (T)0;
    T x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const T y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
      T w = lane < NT / 32 ? s_warp[lane] : (T)0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const T y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      s_warp[lane] = w;
    }
    __syncthreads();
    const T wo = warp ? s_warp[warp - 1] : (T)0;
    if (i < n) a[i] = carry + wo + x - v;
    carry += s_warp[NT / 32 - 1];
    __syncthreads();
  }
  return carry;
}

// bitonic sort, descending, n2 = power of two >= n (padding zeros sort last)
template <int NT>
__device__ void qt_sort_desc(unsigned long long* a, int n2) {
  for (int k = 2; k <= n2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n2; i += NT) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long x = a[i], y = a[ixj];
          const bool desc = (i & k) == 0;
          if (desc ? (x < y) : (x > y)) { a[i] = y; a[ixj] = x; }
        }
      }
      __syncthreads();
    }
  }
}

template <int NT>
__global__ void __launch_bounds__(NT)
orb_quadtree_kernel(const int* __restrict__ cell_cnt, const uint32_t* __restrict__ cell_kp, QtScratch S,
                    uint32_t* __restrict__ kept /*[B][node_total] packed key*/, int* __restrict__ kept_cnt /*[B][nlevels]*/,
                    int* __restrict__ err_flag, const __grid_constant__ OrbGeom g) {
  __shared__ unsigned long long s_warp64[32];
  __shared__ int s_warp32[32];
  __shared__ unsigned long long s_sort[QT_SORT_SMEM];
  __shared__ int s_nalloc, s_cut, s_ncand;

  const int frame = blockIdx.x / g.nlevels, level = blockIdx.x - frame * g.nlevels;
  const OrbLevel& L = g.lv[level];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int N = L.quota;

  const long long ro = (long long)frame * g.raw_total + L.raw_off;
  const long long no = (long long)frame * g.node_total + L.node_off;
  uint32_t* key = S.key + ro;
  int* knode = S.knode + ro;
  uint8_t* kchild = S.kchild + ro;
  int4* bbox = S.bbox + no;
  int* ncount = S.ncount + no;
  int* pidx = S.pidx + no;
  unsigned* best = S.best + no;
  int* ordOld = S.ordA + no;
  int* ordNew = S.ordB + no;
  int* par = S.par + no;
  unsigned long long* cand = S.cand + 2 * no;
  int* ccnt = S.ccnt + 4 * no;
  int* cslot = S.cslot + 4 * no;
  unsigned long long* scan = S.scan + no;
  uint32_t* kept_out = kept + no;

  // ---- gather this level's raw keypoints in (cell row, cell col, row-major in cell) order: vToDistributeKeys
  const int ncell = L.nRows * L.nCols;
  const int* ccount = cell_cnt + (long long)frame * g.ncells_total + L.cell_base;
  const uint32_t* ckp = cell_kp + ((long long)frame * g.ncells_total + L.cell_base) * g.cell_cap;
  // scan cell counts in chunks, reusing `scan` storage is too small for 4K levels, so do a running carry
  int n = 0;
  for (int base = 0; base < ncell; base += NT) {
    const int c = base + tid;
    const int v = c < ncell ? ccount[c] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp32[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int w = lane < NT / 32 ? s_warp32[lane] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      s_warp32[lane] = w;
    }
    __syncthreads();
    const int off = n + (warp ? s_warp32[warp - 1] : 0) + x - v;
    // each thread copies its own cell's entries (<= cell_cap, typically a handful)
    if (c < ncell) {
      const uint32_t* src = ckp + (long long)c * g.cell_cap;
      for (int e = 0; e < v; ++e)
        if (off + e < L.raw_cap) key[off + e] = src[e];
    }
    n += s_warp32[NT / 32 - 1];
    __syncthreads();
  }
  if (tid == 0) {
    S.raw_n[frame * g.nlevels + level] = n;
    if (n > L.raw_cap) atomicOr(err_flag, 2);
  }
  n = min(n, L.raw_cap);
  if (n == 0) {
    if (tid == 0) kept_cnt[frame * g.nlevels + level] = 0;
    return;
  }

  // ---- roots (:1060-1129)
  const int nIni = L.nIni;
  const float hX = L.hX;
  for (int i = tid; i < L.node_cap; i += NT) { ncount[i] = 0; pidx[i] = -1; }
  __syncthreads();
  for (int i = tid; i < nIni; i += NT)
    bbox[i] = make_int4((int)__fmul_rn(hX, (float)i), 0, (int)__fmul_rn(hX, (float)(i + 1)), L.maxBY - ORB_MIN_BORDER);
  {   // n keys land on nIni (1..4, more only for aspect > 4.5) counters: per-thread counts, one warp reduction, one atomic per warp and root
    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    for (int i = tid; i < n; i += NT) {
      const int r = (int)__fdiv_rn((float)(key[i] & 0xfffu), hX);
      knode[i] = r;
      if (r == 0) ++c0; else if (r == 1) ++c1; else if (r == 2) ++c2; else if (r == 3) ++c3; else atomicAdd(&ncount[r], 1);
    }
    c0 = __reduce_add_sync(0xffffffffu, c0); c1 = __reduce_add_sync(0xffffffffu, c1);
    c2 = __reduce_add_sync(0xffffffffu, c2); c3 = __reduce_add_sync(0xffffffffu, c3);
    if (lane == 0) {
      if (c0) atomicAdd(&ncount[0], c0);
      if (c1) atomicAdd(&ncount[1], c1);
      if (c2) atomicAdd(&ncount[2], c2);
      if (c3) atomicAdd(&ncount[3], c3);
    }
  }
  if (tid == 0) s_nalloc = nIni;
  __syncthreads();
  int m = 0, k = 0;   // list size, parents this round (uniform across the block)
  if (tid == 0) {     // nIni is 1..4: serial
    int mm = 0, kk = 0;
    for (int i = 0; i < nIni; ++i) {
      if (ncount[i] > 0) ordOld[mm++] = i;
      if (ncount[i] > 1) par[kk++] = i;
    }
    s_cut = mm;
    s_ncand = kk;
  }
  __syncthreads();
  m = s_cut;
  k = s_ncand;
  __syncthreads();

  bool phase2 = false;
  while (k > 0) {
    // A: register parents
    for (int r = tid; r < k; r += NT) {
      pidx[par[r]] = r;
      ccnt[4 * r] = 0; ccnt[4 * r + 1] = 0; ccnt[4 * r + 2] = 0; ccnt[4 * r + 3] = 0;
    }
    if (tid == 0) { s_cut = k; s_ncand = 0; }
    __syncthreads();
    // B: classify keys of parents (DivideNode :968-1022). Every key costs a chain of dependent loads (node -> parent slot -> box) that hit L2:
    //    QT_ILP keys per thread are in flight together, each stage's loads issued before the previous stage's results are used.
    const bool few_parents = 4 * k <= n / 8;   // on average >= 8 keys per child counter: aggregation pays
    for (int i0 = tid; i0 - lane < n; i0 += NT * QT_ILP) {   // warp-uniform trip count (MATCH.ANY needs the full warp)
      int nd[QT_ILP], r[QT_ILP];
      uint32_t q[QT_ILP];
#pragma unroll
      for (int u = 0; u < QT_ILP; ++u) {
        const int i = i0 + u * NT;
        nd[u] = i < n ? knode[i] : -1;
        q[u] = i < n ? key[i] : 0u;
      }
#pragma unroll
      for (int u = 0; u < QT_ILP; ++u) r[u] = nd[u] >= 0 ? pidx[nd[u]] : -1;
      int4 b[QT_ILP];
#pragma unroll
      for (int u = 0; u < QT_ILP; ++u) b[u] = r[u] >= 0 ? bbox[nd[u]] : make_int4(0, 0, 0, 0);
      // the early rounds have a handful of parents: thousands of keys count into the same few child counters. Lanes that hit the same counter
      // are found with one MATCH.ANY and their leader adds the group's size (the loop is warp-uniform, so all lanes take part)
#pragma unroll
      for (int u = 0; u < QT_ILP; ++u) {
        int slot = -1;
        if (r[u] >= 0) {
          const int hx = (b[u].z - b[u].x + 1) >> 1, hy = (b[u].w - b[u].y + 1) >> 1;   // ceil(w / 2.f), w >= 0
          const int x = q[u] & 0xfffu, y = (q[u] >> 12) & 0xfffu;
          const int c = (x < b[u].x + hx ? 0 : 1) + (y < b[u].y + hy ? 0 : 2);
          kchild[i0 + u * NT] = (uint8_t)c;
          slot = 4 * r[u] + c;
        }
        if (few_parents) {
          const unsigned same = __match_any_sync(0xffffffffu, slot);
          if (slot >= 0 && (int)(__ffs(same) - 1) == lane) atomicAdd(&ccnt[slot], __popc(same));
        } else if (slot >= 0) {
          atomicAdd(&ccnt[slot], 1);
        }
      }
    }
    __syncthreads();
    // C/D: scan nonempty / multi-key child counts over P, find the phase-2 cut
    for (int r = tid; r < k; r += NT) {
      unsigned ne = 0, nc = 0;
#pragma unroll
      for (int c = 0; c < 4; ++c) { const int v = ccnt[4 * r + c]; ne += v > 0; nc += v > 1; }
      scan[r] = (unsigned long long)ne | ((unsigned long long)nc << 32);
    }
    __syncthreads();
    const unsigned long long tot = qt_block_scan_excl<NT, unsigned long long>(scan, k, s_warp64);
    int kc = k;
    if (phase2) {
      for (int r = tid; r < k; r += NT) {
        unsigned ne = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) ne += ccnt[4 * r + c] > 0;
        const int after = m + (int)(unsigned)(scan[r] & 0xffffffffu) + (int)ne - (r + 1);   // list size after splitting P[0..r]
        if (after >= N) atomicMin(&s_cut, r + 1);
      }
      __syncthreads();
      kc = s_cut;
    }
    // totals over the committed prefix
    int T, ncand;
    if (kc == k) { T = (int)(unsigned)(tot & 0xffffffffu); ncand = (int)(tot >> 32); }
    else { const unsigned long long e = scan[kc]; T = (int)(unsigned)(e & 0xffffffffu); ncand = (int)(e >> 32); }
    // E: create children
    for (int r = tid; r < kc; r += NT) {
      const int p = par[r];
      const int4 b = bbox[p];
      const int hx = (b.z - b.x + 1) >> 1, hy = (b.w - b.y + 1) >> 1;
      const int mx = b.x + hx, my = b.y + hy;
      const unsigned long long e = scan[r];
      int o = (int)(unsigned)(e & 0xffffffffu), q = (int)(e >> 32);
      bool first = true;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int cnt = ccnt[4 * r + c];
        if (cnt == 0) continue;
        int slot;
        if (first) { slot = p; first = false; }
        else slot = atomicAdd(&s_nalloc, 1);
        if (slot < L.node_cap) {
          bbox[slot] = make_int4((c & 1) ? mx : b.x, (c & 2) ? my : b.y, (c & 1) ? b.z : mx, (c & 2) ? b.w : my);
          ncount[slot] = cnt;
          ordNew[T - 1 - o] = slot;
          if (cnt > 1) cand[q++] = ((unsigned long long)cnt << 40) | ((unsigned long long)o << 20) | (unsigned)slot;
        }
        cslot[4 * r + c] = slot;
        ++o;
      }
    }
    __syncthreads();
    if (s_nalloc > L.node_cap) {   // cannot happen by the size bound; guard against corrupt state
      if (tid == 0) { atomicOr(err_flag, 4); kept_cnt[frame * g.nlevels + level] = 0; }
      return;
    }
    // F: move keys to their child nodes (same software pipelining as B)
    for (int i0 = tid; i0 < n; i0 += NT * QT_ILP) {
      int nd[QT_ILP], r[QT_ILP], ch[QT_ILP];
#pragma unroll
      for (int u = 0; u < QT_ILP; ++u) {
        const int i = i0 + u * NT;
        nd[u] = i < n ? knode[i] : -1;
        ch[u] = i < n ? (int)kchild[i] : 0;
      }
#pragma unroll
      for (int u = 0; u < QT_ILP; ++u) r[u] = nd[u] >= 0 ? pidx[nd[u]] : -1;
      int sl[QT_ILP];
#pragma unroll
      for (int u = 0; u < QT_ILP; ++u) sl[u] = (r[u] >= 0 && r[u] < kc) ? cslot[4 * r[u] + ch[u]] : -1;
#pragma unroll
      for (int u = 0; u < QT_ILP; ++u)
        if (sl[u] >= 0) knode[i0 + u * NT] = sl[u];
    }
    // G: survivors keep their relative order behind the new children
    unsigned long long* sflag = scan;   // reuse (E is done with it after the barrier below)
    __syncthreads();
    for (int q = tid; q < m; q += NT) {
      const int r = pidx[ordOld[q]];
      sflag[q] = (r >= 0 && r < kc) ? 0ull : 1ull;
    }
    __syncthreads();
    {
      // survivor flags need the pre-scan value after the scan: keep it in a register per element
      // (each element is touched by the same thread before and after: index q = base + tid)
    }
    // scan in place; recover flag as (next - cur) is not possible for the last one, so re-derive from pidx
    qt_block_scan_excl<NT, unsigned long long>(sflag, m, s_warp64);
    for (int q = tid; q < m; q += NT) {
      const int nd = ordOld[q];
      const int r = pidx[nd];
      if (!(r >= 0 && r < kc)) ordNew[T + (int)sflag[q]] = nd;
    }
    __syncthreads();
    // H: unregister parents
    for (int r = tid; r < k; r += NT) pidx[par[r]] = -1;
    const int m_new = T + (m - kc);
    __syncthreads();
    { int* t = ordOld; ordOld = ordNew; ordNew = t; }
    const int m_prev = m;
    m = m_new;
    // ---- loop control (:1260-1268, :1360-1370)
    if (m >= N || m == m_prev) break;
    if (phase2 && kc < k) break;   // unreachable (cut implies m >= N), kept for safety
    if (!phase2 && m + 3 * ncand > N) phase2 = true;
    k = ncand;
    if (k == 0) break;
    if (phase2) {
      int n2 = 1;
      while (n2 < k) n2 <<= 1;
      unsigned long long* buf = n2 <= QT_SORT_SMEM ? s_sort : cand;
      if (buf == s_sort) {
        for (int i = tid; i < n2; i += NT) s_sort[i] = i < k ? cand[i] : 0ull;
      } else {
        for (int i = k + tid; i < n2; i += NT) cand[i] = 0ull;
      }
      __syncthreads();
      qt_sort_desc<NT>(buf, n2);
      for (int r = tid; r < k; r += NT) par[r] = (int)(buf[r] & 0xfffffu);
    } else {
      for (int r = tid; r < k; r += NT) par[r] = (int)(cand[k - 1 - r] & 0xfffffu);   // list order = reverse creation
    }
    __syncthreads();
  }

  // ---- per node keep the best key (:1387-1413), output in list order
  for (int q = tid; q < m; q += NT) best[ordOld[q]] = 0u;
  __syncthreads();
  for (int i0 = tid; i0 < n; i0 += NT * QT_ILP) {
    int nd[QT_ILP];
    uint32_t q[QT_ILP];
#pragma unroll
    for (int u = 0; u < QT_ILP; ++u) {
      const int i = i0 + u * NT;
      nd[u] = i < n ? knode[i] : -1;
      q[u] = i < n ? key[i] : 0u;
    }
#pragma unroll
    for (int u = 0; u < QT_ILP; ++u)
      if (nd[u] >= 0) atomicMax(&best[nd[u]], ((q[u] >> 24) << 24) | (0xffffffu - (unsigned)(i0 + u * NT)));
  }
  __syncthreads();
  for (int q = tid; q < m; q += NT) {
    const unsigned b = best[ordOld[q]];
    kept_out[q] = key[0xffffffu - (b & 0xffffffu)];
  }
  if (tid == 0) kept_cnt[frame * g.nlevels + level] = m;
}

void orb_launch_quadtree(const int* cell_cnt, const uint32_t* cell_kp, const QtScratch& S, uint32_t* kept, int* kept_cnt,
                         int* err_flag, const OrbGeom& g, int batch, cudaStream_t st, long long* launches) {
  // large levels (4K: 9017 cells, quota 1737 at level 0) are a long chain of rounds over tens of thousands of keys: twice the threads
  // ...and so are the few CTAs of a small pass (a Frame's one or two images: latency matters and the SMs are idle anyway)
  const bool huge = g.lv[0].nCols * g.lv[0].nRows >= 2048 || g.lv[0].quota >= 1024;   // 4K: tens of thousands of keys per level
  const bool big = batch <= 8;
  static const int force = getenv("ORB_B200_QT_THREADS") ? atoi(getenv("ORB_B200_QT_THREADS")) : 0;   // A/B switch
  const int nt = force ? force : huge ? 1024 : big ? 512 : QT_SMALL;
  if (nt >= 1024) orb_quadtree_kernel<1024><<<batch * g.nlevels, 1024, 0, st>>>(cell_cnt, cell_kp, S, kept, kept_cnt, err_flag, g);
  else if (nt >= 512) orb_quadtree_kernel<512><<<batch * g.nlevels, 512, 0, st>>>(cell_cnt, cell_kp, S, kept, kept_cnt, err_flag, g);
  else orb_quadtree_kernel<QT_SMALL><<<batch * g.nlevels, QT_SMALL, 0, st>>>(cell_cnt, cell_kp, S, kept, kept_cnt, err_flag, g);
  ++*launches;
}
