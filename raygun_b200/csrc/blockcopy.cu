// Block transfers of blockwise prediction (predict/worker.py:100 `source.to_ndarray(block.read_roi)` ... :160
// `destination[block.write_roi] = out`): a (d,h,w) sub-block of a uint8 (D,H,W) volume moves between HOST memory and a
// dense device buffer as ONE strided 3-D DMA (cudaMemcpy3DAsync) on the caller's copy stream -- no host-side gather /
// scatter, no staging ring, nothing on the CPU's critical path.  With page-locked host memory the copy is asynchronous and
// overlaps the convolutions of the neighbouring blocks; the same entry points serve a device-resident volume
// (device-to-device strided copy).
#include "common.cuh"

using namespace rg;

static int block_copy(const uint8_t* src, const int64_t* src_dims, const int64_t* src_org, uint8_t* dst,
                      const int64_t* dst_dims, const int64_t* dst_org, const int64_t* ext, cudaMemcpyKind kind, void* stream,
                      const char* who) {
  RG_CHECK_ARG(src && dst && src_dims && dst_dims && src_org && dst_org && ext, "%s: null argument", who);
  for (int i = 0; i < 3; ++i) {
    RG_CHECK_ARG(ext[i] >= 0 && src_org[i] >= 0 && dst_org[i] >= 0 && src_org[i] + ext[i] <= src_dims[i] &&
                 dst_org[i] + ext[i] <= dst_dims[i], "%s: block outside the volume (axis %d)", who, i);
  }
  if (ext[0] == 0 || ext[1] == 0 || ext[2] == 0) return RG_OK;
  cudaMemcpy3DParms p;
  memset(&p, 0, sizeof(p));
  // uint8 volumes, W innermost: pitch = W bytes, "height" = H rows, depth = D slices
  p.srcPtr = make_cudaPitchedPtr((void*)src, (size_t)src_dims[2], (size_t)src_dims[2], (size_t)src_dims[1]);
  p.dstPtr = make_cudaPitchedPtr((void*)dst, (size_t)dst_dims[2], (size_t)dst_dims[2], (size_t)dst_dims[1]);
  p.srcPos = make_cudaPos((size_t)src_org[2], (size_t)src_org[1], (size_t)src_org[0]);
  p.dstPos = make_cudaPos((size_t)dst_org[2], (size_t)dst_org[1], (size_t)dst_org[0]);
  p.extent = make_cudaExtent((size_t)ext[2], (size_t)ext[1], (size_t)ext[0]);
  p.kind = kind;
  cudaError_t e = cudaMemcpy3DAsync(&p, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, who);
  return RG_OK;
}

extern "C" {

int rg_h2d_block(const uint8_t* h_vol, const int64_t vol_dhw[3], const int64_t origin[3], const int64_t ext[3],
                 uint8_t* d_block, void* stream) {
  const int64_t zero[3] = {0, 0, 0};
  return block_copy(h_vol, vol_dhw, origin, d_block, ext, zero, ext, cudaMemcpyDefault, stream, "rg_h2d_block");
}

int rg_d2h_block(const uint8_t* d_block, const int64_t blk_dhw[3], const int64_t blk_origin[3], uint8_t* h_vol,
                 const int64_t vol_dhw[3], const int64_t origin[3], const int64_t ext[3], void* stream) {
  return block_copy(d_block, blk_dhw, blk_origin, h_vol, vol_dhw, origin, ext, cudaMemcpyDefault, stream, "rg_d2h_block");
}

}  // extern "C"
