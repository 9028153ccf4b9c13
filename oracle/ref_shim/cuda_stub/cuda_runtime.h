// empty on purpose (host build of the reference; see ../cuda_host_emul.h)
