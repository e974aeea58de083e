// parsim_b200.cu -- single translation unit of libparsim_b200.so (kernels are defined in
// headers, so the host-side files are compiled together).
#include "parsim_rank.cu"
#include "parsim_capi.cu"
#include "parsim_noalloc.cu"
