#define GB_T 256
#include "launch_impl.cuh"
