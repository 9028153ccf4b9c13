// empty on purpose: shadows torch's CUDA-only THC/THCAtomics.cuh when the reference's
// correlation_kernel.cu is compiled as host code (see ../cuda_host_emul.h)
