// Development probe (GPU box): which TMA box shapes / destinations load correctly on this driver. One case per process:
//   tma_probe BOX_W BOX_H DST_OFFSET NLOADS   -> prints OK / MISMATCH / the CUDA error
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../orb_slam2_detailed_comments_b200/csrc/orb_tma.cuh"
int orb_build_level_maps(const uint8_t* base, int frames, const OrbGeom&, int box_w, int box_h, void* maps_out);

__global__ void probe(const __grid_constant__ OrbLevelMaps maps, uint8_t* out, int bw, int bh, int dst_off, int nloads, int x0, int y0, int inc) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) unsigned long long bar;
  const unsigned b = orb_smem_u32(&bar);
  if (threadIdx.x == 0) orb_mbar_init(b, 1);
  __syncthreads();
  for (int i = 0; i < nloads; ++i) {
    __syncwarp();
    if (threadIdx.x == 0) {
      orb_mbar_expect_tx(b, bw * bh);
      orb_tma_load_3d(orb_smem_u32(smem + dst_off), &maps.m[0], x0 + i * inc, y0 + i, 0, b);
    }
    if (!orb_mbar_wait(b, i & 1)) { if (threadIdx.x == 0) printf("wait timed out\n"); return; }
    for (int k = threadIdx.x; k < bw * bh; k += blockDim.x) out[(size_t)i * bw * bh + k] = smem[dst_off + k];
  }
}

int main(int argc, char** argv) {
  const int bw = atoi(argv[1]), bh = atoi(argv[2]), off = atoi(argv[3]), nl = atoi(argv[4]);
  const int x0 = argc > 5 ? atoi(argv[5]) : 37, y0 = 21, inc = argc > 6 ? atoi(argv[6]) : 1;
  OrbGeom g; memset(&g, 0, sizeof(g));
  g.nlevels = 1;
  OrbLevel& L = g.lv[0];
  L.cols = 600; L.rows = 400; L.pitch = 640; L.hCell = 32;
  L.plane_off = (size_t)ORB_APRON_ROWS * L.pitch + 16;
  g.frame_stride = (size_t)L.pitch * (L.rows + 2 * ORB_APRON_ROWS);
  std::vector<uint8_t> h(g.frame_stride);
  for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)(i * 2654435761u >> 13);
  uint8_t *d, *o;
  cudaMalloc(&d, h.size()); cudaMalloc(&o, (size_t)nl * bw * bh);
  cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  alignas(64) static OrbLevelMaps maps;
  if (orb_build_level_maps(d, 1, g, bw, bh, &maps)) { printf("encode failed %dx%d\n", bw, bh); return 2; }
  probe<<<1, 32, off + bw * bh + 256>>>(maps, o, bw, bh, off, nl, x0, y0, inc);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("box %dx%d off %d loads %d x0 %d inc %d: %s\n", bw, bh, off, nl, x0, inc, cudaGetErrorString(e)); return 1; }
  std::vector<uint8_t> r((size_t)nl * bw * bh);
  cudaMemcpy(r.data(), o, r.size(), cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int i = 0; i < nl; ++i)
    for (int y = 0; y < bh; ++y)
      for (int x = 0; x < bw; ++x)
        bad += r[((size_t)i * bh + y) * bw + x] != h[(size_t)(y0 + i + y) * L.pitch + x0 + i * inc + x];
  printf("box %dx%d off %d loads %d x0 %d inc %d: %s (%d bad bytes)\n", bw, bh, off, nl, x0, inc, bad ? "MISMATCH" : "OK", bad);
  return bad != 0;
}
