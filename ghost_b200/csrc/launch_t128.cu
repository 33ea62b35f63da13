#define GB_T 128
#include "launch_impl.cuh"
