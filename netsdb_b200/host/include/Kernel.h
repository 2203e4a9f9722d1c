// Kernel.h -- same header name as src/conv2d_memory_fusion/headers/Kernel.h; the class lives in conv_udfs.h
#pragma once
#include "conv_udfs.h"
