// launch.cuh — host-side dispatch of the lanes-per-row template parameter.
#pragma once
#include "cg_kernels.cuh"
#include "matrix.cuh"

// CGB_LAUNCH_L(lanes, kernel_template, grid, stream, args...) launches kernel_template<L><<<grid, kBlock>>>.
#define CGB_LAUNCH_L(lanes, KERNEL, grid, stream, ...)                                   \
  do {                                                                                   \
    switch (lanes) {                                                                     \
    case 1: KERNEL<1><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break;         \
    case 2: KERNEL<2><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break;         \
    case 4: KERNEL<4><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break;         \
    case 8: KERNEL<8><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break;         \
    case 16: KERNEL<16><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break;       \
    default: KERNEL<32><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break;       \
    }                                                                                    \
  } while (0)

// Same for kernels with a second bool template parameter.
#define CGB_LAUNCH_LB(lanes, flag, KERNEL, grid, stream, ...)                            \
  do {                                                                                   \
    if (flag) {                                                                          \
      switch (lanes) {                                                                   \
      case 1: KERNEL<1, true><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      case 2: KERNEL<2, true><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      case 4: KERNEL<4, true><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      case 8: KERNEL<8, true><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      case 16: KERNEL<16, true><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      default: KERNEL<32, true><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      }                                                                                  \
    } else {                                                                             \
      switch (lanes) {                                                                   \
      case 1: KERNEL<1, false><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      case 2: KERNEL<2, false><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      case 4: KERNEL<4, false><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      case 8: KERNEL<8, false><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      case 16: KERNEL<16, false><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      default: KERNEL<32, false><<<(grid), cgb::kBlock, 0, (stream)>>>(__VA_ARGS__); break; \
      }                                                                                  \
    }                                                                                    \
  } while (0)
