// Fast-variant instantiations of the likelihood kernels (placeholder: forwards to the faithful kernels until the
// restructured evaluator lands).
#include "common.cuh"

namespace wfpt {
cudaError_t launch_like_batch_fast(const BatchArgs& a) { return launch_like_batch_faithful(a); }
cudaError_t launch_pdf_array_fast(const PdfArrayArgs& a) { return launch_pdf_array_faithful(a); }
cudaError_t launch_like_multi_fast(const MultiArgs& a) { return launch_like_multi_faithful(a); }
}  // namespace wfpt
