// PDBObject.h -- same header name as in the netsDB tree; everything lives in pdb_shim.h
#pragma once
#include "pdb_shim.h"
