// Query.h -- same header name as the reference's queries/headers/Query.h; the drivers of this path need nothing from it beyond the shim
#pragma once
#include "pdb_shim.h"
