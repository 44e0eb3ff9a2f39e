// Host-side fast path of the gradient hand-over (no CUDA in here): the per-parameter loop that takes the gradients
// autograd produced, checks that the flatten kernel can read them as they are, writes their addresses into the
// dgs_seg table the kernel is launched with, and clears p.grad -- done natively instead of in Python, where it costs
// about 2 us per parameter.  torch's own DistributedDataParallel does the equivalent part of
// its reducer in C++ (reference harness/determined/pytorch/_pytorch_context.py:295-296 wraps the model in it).
//
// Built as a small torch extension next to libdgs.so.  The C ABI of libdgs.so is unchanged: this module only fills
// the host array that dgs_flatten_accumulate() is given.
#include <torch/extension.h>

#include <cstdint>
#include <vector>

namespace {

struct Seg {  // == dgs_seg of include/dgs.h
  const void* src;
  int64_t dst_off;
  int64_t numel;
};
static_assert(sizeof(Seg) == 24, "layout of dgs_seg");

at::ScalarType dtype_of(int code) {  // DGS_F32 / DGS_BF16 / DGS_F16
  switch (code) {
    case 0: return at::kFloat;
    case 1: return at::kBFloat16;
    case 2: return at::kHalf;
    default: TORCH_CHECK(false, "dgs_host: unknown dtype code ", code);
  }
}

// All-or-nothing.  If every parameter has a dense, contiguous gradient of the given dtype: write the gradients'
// addresses into table[k].src (k = position in `params`; dst_off and numel were filled when the table was built),
// set every p.grad to None and return the gradient tensors, which the caller keeps alive until the kernel that
// reads them has been enqueued.  Otherwise touch nothing and return None (the caller takes its general path).
py::object take_grads(const std::vector<at::Tensor>& params, int64_t table_addr, int dtype_code) {
  const at::ScalarType want = dtype_of(dtype_code);
  for (const at::Tensor& p : params) {
    const at::Tensor& g = p.grad();
    if (!g.defined() || g.scalar_type() != want || g.layout() != at::kStrided || !g.is_contiguous()) return py::none();
  }
  Seg* table = reinterpret_cast<Seg*>(static_cast<intptr_t>(table_addr));
  py::list taken(params.size());
  for (size_t k = 0; k < params.size(); ++k) {
    at::Tensor& slot = params[k].mutable_grad();
    at::Tensor g = std::move(slot);
    slot = at::Tensor();  // p.grad = None: autograd must not accumulate into a tensor the kernel consumes
    table[k].src = g.data_ptr();
    taken[k] = py::cast(std::move(g));
  }
  return std::move(taken);
}

// True if every parameter currently has a gradient (the sentinel hook's "is the bucket complete" test).
bool all_have_grads(const std::vector<at::Tensor>& params) {
  for (const at::Tensor& p : params)
    if (!p.grad().defined()) return false;
  return true;
}

// True if any parameter has a gradient (step_optimizer's check for gradients produced outside context.backward()).
bool any_has_grad(const std::vector<at::Tensor>& params) {
  for (const at::Tensor& p : params)
    if (p.grad().defined()) return true;
  return false;
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "host-side fast path of the grad-sync engine (fills dgs_seg tables from parameter gradients)";
  m.def("take_grads", &take_grads, py::arg("params"), py::arg("table_addr"), py::arg("dtype_code"));
  m.def("all_have_grads", &all_have_grads, py::arg("params"));
  m.def("any_has_grad", &any_has_grad, py::arg("params"));
}
