// Launch / shared-memory portability shim.  The product is built by nvcc for sm_100a; defining
// B200_EMU (tests/emu only) swaps in the SIMT emulator so the same kernels can be logic-tested in the
// GPU-less build container.  Nothing in the product build defines B200_EMU.
#pragma once
#ifdef B200_EMU
#include "simt_emu.h"
#define B200_LAUNCH(kernel, grid, block, smem, stream, ...) \
    emu::launch(dim3(grid), dim3(block), (smem), [=]() { kernel(__VA_ARGS__); })
#define B200_DYN_SMEM(name) unsigned char *name = emu::dyn_smem()
#else
#include <cuda_runtime.h>
#define B200_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#define B200_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#endif
#include <stdint.h>

#define B200_FULL_MASK 0xffffffffu
typedef uint64_t u64;
typedef int64_t i64;
typedef uint16_t u16;
typedef uint32_t u32;
typedef int32_t i32;
typedef uint8_t u8;
