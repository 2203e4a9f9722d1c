#pragma once
#include "net_udfs.h"
