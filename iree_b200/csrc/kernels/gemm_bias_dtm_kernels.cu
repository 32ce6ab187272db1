// gemm_bias_dtm_kernels.cu — `<name>_dtm` twins of the fused bias / activation GEMM exports
// (gemm_bias_exports.inc): tensor maps built on the device, launchable by an unmodified IREE
// CUDA driver. Same source as gemm_dtm_kernels.cu, its own cubin (and its own slots: an SM
// holds one CTA of either family at a time).
#define B200_DTM_BIAS 1
#include "gemm_dtm_kernels.cu"
