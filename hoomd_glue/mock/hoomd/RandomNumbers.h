#pragma once
#include "../../../oracle/shim/hoomd/RandomNumbers.h"
