/* GetVTable.h -- stand-in for the reference header of the same name: see pdb_min.h (test infrastructure) */
#pragma once
#include "pdb_min.h"
