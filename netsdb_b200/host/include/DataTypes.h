// DataTypes.h -- same header name as the reference's utilities/headers/DataTypes.h.  The reference's copy (and the
// storage headers the real PDBClient.h pulls in) carry a file-scope `using namespace std;` that the LA drivers rely on
// (unqualified cout / string / make_shared in TestLA04_Multiply.cc); reproduced here for them.
#pragma once
#include "pdb_shim.h"
using namespace std;
