// noalloc_internal.h — the two private hooks noalloc.cu needs from capi.cu (handles stay opaque).
#pragma once
#include "lowered.h"

const ghost_b200::LoweredModel* ghost_b200_internal_lowered(const GhostModel* model);
void ghost_b200_internal_set_error(int code, const char* message);
