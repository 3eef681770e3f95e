// Test infrastructure: C-ABI shim around the REFERENCE's own rotated-IoU CPU code
// (utils/export_models/OpenPCDet/pcdet/ops/iou3d_nms/src/iou3d_cpu.cpp, compiled where it lies by
// oracle/build_ref.py into oracle/_ref/libiou3d_ref.so). Nothing of the reference is copied here.
#include <torch/torch.h>
#include "iou3d_cpu.h"

extern "C" int ref_boxes_iou_bev_cpu(const float* a, int n, const float* b, int m, float* out) {
    auto opts = torch::TensorOptions().dtype(torch::kFloat32);
    at::Tensor ta = torch::from_blob(const_cast<float*>(a), {n, 7}, opts);
    at::Tensor tb = torch::from_blob(const_cast<float*>(b), {m, 7}, opts);
    at::Tensor to = torch::from_blob(out, {n, m}, opts);
    return boxes_iou_bev_cpu(ta, tb, to);
}
