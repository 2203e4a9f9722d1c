// ConvChunksToImage.h -- same header name as src/conv2d_memory_fusion/headers/ConvChunksToImage.h; the class lives in conv_udfs.h
#pragma once
#include "conv_udfs.h"
