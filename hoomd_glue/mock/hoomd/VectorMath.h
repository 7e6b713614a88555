#pragma once
#include "../../../oracle/shim/hoomd/VectorMath.h"
