#define GB_T 64
#include "launch_impl.cuh"
