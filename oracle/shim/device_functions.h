/* TEST INFRASTRUCTURE: forwards the CUDA header name the reference includes to the host shim. */
#include "cuda_shim.h"
