#define GB_T 32
#include "launch_impl.cuh"
