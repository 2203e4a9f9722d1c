// TensorData.h -- same header name as the reference's; nothing on this path needs it beyond the shim
#pragma once
#include "pdb_shim.h"
