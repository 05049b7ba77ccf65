// PyTorch extension binding: registers `torch.ops.seqalign.*` so the solver can be called on
// torch tensors, on the current torch CUDA stream, from eager code or a captured graph region.
// (The reference has no Python entry point at all; its only interface is ./final on stdin.)
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include "seqalign.h"

namespace {

const sa_blob_header* header_of(const at::Tensor& header_cpu) {
    TORCH_CHECK(header_cpu.device().is_cpu() && header_cpu.scalar_type() == at::kByte && header_cpu.is_contiguous() &&
                    header_cpu.numel() >= (int64_t)sizeof(sa_blob_header),
                "header must be a contiguous CPU uint8 tensor holding the blob's first sizeof(sa_blob_header) = 1024 bytes");
    const auto* h = reinterpret_cast<const sa_blob_header*>(header_cpu.data_ptr());
    TORCH_CHECK(h->magic == SA_BLOB_MAGIC && h->version == SA_BLOB_VERSION, "not a packed problem of this version");
    return h;
}

// blob: CUDA uint8 tensor with the packed problem; scratch: CUDA uint8 tensor of
// scratch_bytes(nseq), zero on first use. Returns int32[nseq, 4] = (score, n, k, 0).
at::Tensor solve_blob(const at::Tensor& blob, const at::Tensor& header_cpu, at::Tensor scratch) {
    const sa_blob_header* h = header_of(header_cpu);
    TORCH_CHECK(blob.is_cuda() && blob.scalar_type() == at::kByte && blob.is_contiguous(), "blob must be a contiguous CUDA uint8 tensor");
    TORCH_CHECK(blob.numel() >= (int64_t)h->total_bytes, "blob tensor shorter than the packed problem");
    TORCH_CHECK(scratch.is_cuda() && scratch.device() == blob.device() && scratch.scalar_type() == at::kByte &&
                    scratch.numel() >= (int64_t)sa_cuda_scratch_bytes(h->nseq),
                "scratch must be a CUDA uint8 tensor of at least scratch_bytes(nseq) on the blob's device");
    c10::cuda::CUDAGuard guard(blob.device());
    at::Tensor out = at::empty({(int64_t)h->nseq, 4}, blob.options().dtype(at::kInt));
    if (h->nseq == 0) return out;
    const int rc = sa_cuda_solve_blob(blob.data_ptr(), h, out.data_ptr<int>(), scratch.data_ptr(),
                                      at::cuda::getCurrentCUDAStream().stream());
    TORCH_CHECK(rc == SA_OK, "seqalign: ", sa_last_error());
    return out;
}

int64_t scratch_bytes(int64_t nseq) { return (int64_t)sa_cuda_scratch_bytes((uint32_t)nseq); }

// text in the reference's stdin grammar -> CPU int32[nseq, 3] via the native session path.
at::Tensor solve_text(const std::string& text, int64_t device) {
    sa_problem* p = nullptr;
    TORCH_CHECK(sa_parse(text.data(), text.size(), &p) == SA_OK, "seqalign: ", sa_last_error());
    const int64_t n = sa_problem_nseq(p);
    at::Tensor out = at::empty({n, 3}, at::TensorOptions().dtype(at::kInt));
    const int rc = n ? sa_cuda_solve_problem(p, out.data_ptr<int>(), (int)device) : SA_OK;
    sa_problem_free(p);
    TORCH_CHECK(rc == SA_OK, "seqalign: ", sa_last_error());
    return out;
}

}  // namespace

TORCH_LIBRARY(seqalign, m) {
    m.def("solve_blob(Tensor blob, Tensor header, Tensor(a!) scratch) -> Tensor");
    m.def("scratch_bytes(int nseq) -> int", &scratch_bytes);
    m.def("solve_text(str text, int device) -> Tensor", &solve_text);
}

TORCH_LIBRARY_IMPL(seqalign, CUDA, m) { m.impl("solve_blob", &solve_blob); }

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) { m.doc() = "seqalign torch ops (see torch.ops.seqalign)"; }
