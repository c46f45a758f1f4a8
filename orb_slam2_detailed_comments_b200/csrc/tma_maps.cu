// Host side of orb_tma.cuh: per-level TMA descriptors (cuTensorMapEncodeTiled, resolved through the runtime's driver entry point table:
// no link-time dependency on libcuda).
#include <atomic>
#include <cstring>

#include "orb_tma.cuh"

typedef CUresult (*OrbEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static OrbEncodeTiled orb_encode_tiled() {
  static std::atomic<OrbEncodeTiled> fn{nullptr};
  OrbEncodeTiled f = fn.load(std::memory_order_acquire);
  if (!f) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
      f = reinterpret_cast<OrbEncodeTiled>(p);
    fn.store(f, std::memory_order_release);
  }
  return f;
}

// One descriptor per level over the plane buffer `base` (pyramid or blurred) of a pass workspace holding `frames` frames: tensor
// [frames][plane rows][pitch bytes] u8 starting at the plane allocation, box = box_w bytes x box_h rows (box_h == 0: the level's FAST window
// height hCell + 6) x 1 frame. box_w must be a multiple of 16 and <= 256, box_h <= 256. Returns 0 on success.
int orb_build_level_maps(const uint8_t* base, int frames, const OrbGeom& g, int box_w, int box_h, void* maps_out) {
  static_assert(sizeof(OrbLevelMaps) == ORB_LEVEL_MAPS_BYTES, "CUtensorMap is 128 bytes");
  OrbLevelMaps* M = reinterpret_cast<OrbLevelMaps*>(maps_out);
  memset(M, 0, sizeof(*M));
  OrbEncodeTiled enc = orb_encode_tiled();
  if (!enc || !base) return -1;
  for (int l = 0; l < g.nlevels; ++l) {
    const OrbLevel& L = g.lv[l];
    void* p = const_cast<uint8_t*>(base) + L.plane_off - (size_t)ORB_APRON_ROWS * L.pitch - 16;
    const cuuint64_t dims[3] = {(cuuint64_t)L.pitch, (cuuint64_t)(L.rows + 2 * ORB_APRON_ROWS), (cuuint64_t)frames};
    const cuuint64_t strides[2] = {(cuuint64_t)L.pitch, (cuuint64_t)g.frame_stride};
    const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)(box_h ? box_h : L.hCell + 6), 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    if (box[1] > 256) return -1;
    if (enc(&M->m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, p, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return -1;
  }
  return 0;
}
