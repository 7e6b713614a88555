// MOCK (compile-only guard): forwards to the oracle's shim of hoomd/HOOMDMath.h and adds the HIP/CUDA types the glue names.
#pragma once
#include "../../../oracle/shim/hoomd/HOOMDMath.h"
#include <cstddef>
#include <cstdint>
#ifndef AZB_MOCK_HIP_TYPES
#define AZB_MOCK_HIP_TYPES
typedef int hipError_t;
static const hipError_t hipSuccess = 0;
static const hipError_t hipErrorInvalidValue = 1;
struct hipDeviceProp_t
    {
    int major, minor;
    };
namespace hoomd
    {
struct uchar3
    {
    unsigned char x, y, z;
    };
    }
#endif
