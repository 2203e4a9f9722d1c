// LASillyMultiply2Aggregate.h -- same header name as sharedLibraries/headers/LASillyMultiply2Aggregate.h; the class lives in la_udfs.h
#pragma once
#include "la_udfs.h"
