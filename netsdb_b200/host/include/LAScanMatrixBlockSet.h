// LAScanMatrixBlockSet.h -- same header name as sharedLibraries/headers/LAScanMatrixBlockSet.h; the class lives in la_udfs.h
#pragma once
#include "la_udfs.h"
