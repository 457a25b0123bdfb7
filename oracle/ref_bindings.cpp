// ORACLE — TEST INFRASTRUCTURE ONLY.
// Minimal pybind11 module exposing the reference's four hot-path host functions (and the two of the DefTet
// sparse renderer, SURVEY.md section 8f row 4, kaolin/csrc/bindings.cpp:81-82), compiled from the
// reference sources where they lie under /root/reference (never copied into this repo). The
// declarations come from the reference's own headers; the four .def lines correspond to
// kaolin/csrc/bindings.cpp:75-80 (submodule render.mesh).
#include <torch/extension.h>
#include "kaolin/csrc/render/mesh/rasterization.h"
#include "kaolin/csrc/render/mesh/dibr_soft_mask.h"
#include "kaolin/csrc/render/mesh/deftet.h"

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  namespace py = pybind11;
  py::module render = m.def_submodule("render");
  py::module mesh = render.def_submodule("mesh");
  mesh.def("packed_rasterize_forward_cuda", &kaolin::packed_rasterize_forward_cuda);
  mesh.def("rasterize_backward_cuda", &kaolin::rasterize_backward_cuda);
  mesh.def("dibr_soft_mask_forward_cuda", &kaolin::dibr_soft_mask_forward_cuda);
  mesh.def("dibr_soft_mask_backward_cuda", &kaolin::dibr_soft_mask_backward_cuda);
  mesh.def("deftet_sparse_render_forward_cuda", &kaolin::deftet_sparse_render_forward_cuda);
  mesh.def("deftet_sparse_render_backward_cuda", &kaolin::deftet_sparse_render_backward_cuda);
}
