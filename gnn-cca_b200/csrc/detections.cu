// Detection tables -> device tensors: what libs/datasets.py:99-134 does with pandas per
// sequence and libs/utils.py:211-225 per detection, for a whole sequence at once.
//   foot_to_world_kernel : ground-plane coordinates of every detection: the bounding box's bottom-centre
//                          (xmin + round(width / 2), ymax) (datasets.py:101-102) through the camera's image->world homography
//                          (utils.py:211-225: prj = H [x, y, 1]^T, (prj0 / prj2, prj1 / prj2)), float64 like numpy.
//   frame_valid_kernel   : the frames the reference trains on (datasets.py:126-134): detections from MORE than one camera and at
//                          least one identity seen more than once.  One warp per frame over the frame-sorted table.
// The frame-sorted order (stable: camera-major inside a frame, as `data_det[data_det['frame'] == f]` returns it) and the
// per-frame offsets come from gnncca_csr_sort on the frame numbers.
#include "common.cuh"

namespace gnncca {
namespace det {

__global__ void __launch_bounds__(256) foot_to_world_kernel(const double* __restrict__ xmin, const double* __restrict__ xmax,
                                                            const double* __restrict__ ymax, const int32_t* __restrict__ id_cam,
                                                            const double* __restrict__ H /* [n_cams][9] */, int n_cams, int64_t n,
                                                            double* __restrict__ xw, double* __restrict__ yw, int32_t* __restrict__ flags) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = id_cam[i];
  if (c < 0 || c >= n_cams) { atomicOr(flags, 1); return; }
  const double* h = H + (size_t)c * 9;
  const double width = xmax[i] - xmin[i];                         // datasets.py:95 (xmin already shifted by the caller, :94)
  const double x = xmin[i] + rint(width / 2.0);                   // np.round: half to even
  const double y = ymax[i];
  // left-to-right sums without contraction, as a plain dot product evaluates them
  const double p0 = __dadd_rn(__dadd_rn(__dmul_rn(h[0], x), __dmul_rn(h[1], y)), h[2]);
  const double p1 = __dadd_rn(__dadd_rn(__dmul_rn(h[3], x), __dmul_rn(h[4], y)), h[5]);
  const double p2 = __dadd_rn(__dadd_rn(__dmul_rn(h[6], x), __dmul_rn(h[7], y)), h[8]);
  xw[i] = p0 / p2;
  yw[i] = p1 / p2;
}

// frame f owns rows frame_ptr[f] .. frame_ptr[f+1] of the frame-sorted table (perm maps them to the caller's rows)
__global__ void __launch_bounds__(256) frame_valid_kernel(const int32_t* __restrict__ frame_ptr, const int32_t* __restrict__ perm,
                                                          const int32_t* __restrict__ id_cam, const int64_t* __restrict__ ids,
                                                          int64_t n_frames, uint8_t* __restrict__ valid) {
  const int lane = threadIdx.x & 31;
  const int64_t f = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (f >= n_frames) return;
  const int a = frame_ptr[f], b = frame_ptr[f + 1];
  bool two_cams = false, repeated = false;
  if (b - a >= 2) {
    const int cam0 = id_cam[perm[a]];
    for (int i = a + lane; i < b; i += 32) {
      const int r = perm[i];
      if (id_cam[r] != cam0) two_cams = true;
      const int64_t id = ids[r];
      for (int j = i + 1; j < b; ++j) if (ids[perm[j]] == id) { repeated = true; break; }   // np.bincount(id).max() > 1
    }
  }
  two_cams = __any_sync(0xffffffffu, two_cams);
  repeated = __any_sync(0xffffffffu, repeated);
  if (lane == 0) valid[f] = (two_cams && repeated) ? 1 : 0;
}

}  // namespace det
}  // namespace gnncca

using namespace gnncca;

extern "C" int gnncca_foot_to_world(const double* xmin, const double* xmax, const double* ymax, const int32_t* id_cam, const double* H,
                                    int32_t n_cams, int64_t n_dets, double* xw, double* yw, int32_t* flags, void* stream) {
  if (n_dets < 0 || n_cams < 1 || !flags) return GNNCCA_ERR_BAD_ARG;
  if (n_dets == 0) return GNNCCA_OK;
  if (!xmin || !xmax || !ymax || !id_cam || !H || !xw || !yw) return GNNCCA_ERR_BAD_ARG;
  det::foot_to_world_kernel<<<(unsigned)ceil_div(n_dets, 256), 256, 0, (cudaStream_t)stream>>>(xmin, xmax, ymax, id_cam, H, n_cams, n_dets, xw, yw, flags);
  return check_launch();
}

extern "C" int gnncca_frame_valid(const int32_t* frame_ptr, const int32_t* perm, const int32_t* id_cam, const int64_t* ids,
                                  int64_t n_frames, uint8_t* valid, void* stream) {
  if (n_frames < 0) return GNNCCA_ERR_BAD_ARG;
  if (n_frames == 0) return GNNCCA_OK;
  if (!frame_ptr || !perm || !id_cam || !ids || !valid) return GNNCCA_ERR_BAD_ARG;
  det::frame_valid_kernel<<<(unsigned)ceil_div(n_frames, 8), 256, 0, (cudaStream_t)stream>>>(frame_ptr, perm, id_cam, ids, n_frames, valid);
  return check_launch();
}
