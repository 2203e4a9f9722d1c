// LASillyTransposeMultiply1Join.h -- same header name as sharedLibraries/headers/LASillyTransposeMultiply1Join.h; the class lives in la_udfs.h
#pragma once
#include "la_udfs.h"
