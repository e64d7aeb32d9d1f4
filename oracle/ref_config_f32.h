// TEST INFRASTRUCTURE ONLY. Force-included (-include) when building the NUM_OBJECTS=32 variant of
// the unmodified reference rasterizer for config 4: the reference's cuda_rasterizer/config.h:15-18
// hard-codes 16 without an #ifndef guard, so its include guard is pre-defined here and the four
// constants are supplied instead. CUB is pulled in first because its templates use NUM_CHANNELS as
// a parameter name.
#pragma once
#include <cub/cub.cuh>
#include <cub/device/device_radix_sort.cuh>
#define CUDA_RASTERIZER_CONFIG_H_INCLUDED
#define NUM_CHANNELS 3
#define NUM_OBJECTS 32
#define BLOCK_X 16
#define BLOCK_Y 16
