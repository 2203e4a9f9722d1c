// FFTransposeMult.h -- same header name as src/FF/headers/FFTransposeMult.h; the class lives in ff_udfs.h
#pragma once
#include "ff_udfs.h"
