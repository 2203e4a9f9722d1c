// KernelToChunks.h -- same header name as src/conv2d_memory_fusion/headers/KernelToChunks.h; the class lives in conv_udfs.h
#pragma once
#include "conv_udfs.h"
